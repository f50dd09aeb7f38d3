N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_tp_gpu.py -m gpu -x -q -s 2>&1 | tail -25
timeout 600 python tools/bench_gateway.py --gpus $N --rounds 3 --levels gateway --remote coalesce > gpurun_out/t8_gateway_coalesce_n$N.json 2> gpurun_out/t8_gateway_coalesce_n$N.err; tail -c 900 gpurun_out/t8_gateway_coalesce_n$N.json; tail -3 gpurun_out/t8_gateway_coalesce_n$N.err
timeout 600 python tools/bench_gateway.py --gpus $N --rounds 3 --levels gateway > gpurun_out/t8_gateway_plain_n$N.json 2> gpurun_out/t8_gateway_plain_n$N.err; tail -c 900 gpurun_out/t8_gateway_plain_n$N.json
