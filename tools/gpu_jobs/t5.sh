timeout 900 python -m pytest tests/test_dense_gpu.py -m gpu -x -q -s 2>&1 | tail -25
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 1500 python bench.py --preset qwen2.5-7b-bf16 --batch 32 --prompt-len 2048 --new-tokens 64 --steps 2 --warmup 3 --no-gateway > gpurun_out/t5_bench7b.json 2> gpurun_out/t5_bench7b.err
tail -c 3000 gpurun_out/t5_bench7b.json; tail -5 gpurun_out/t5_bench7b.err
