N=${N:-$(nvidia-smi -L | wc -l)}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/t7_bench_n$N.json 2> gpurun_out/t7_bench_n$N.err
python -c "
import json,sys; d=json.loads(open('gpurun_out/t7_bench_n$N.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('n_gpus','value','e2e','levels_tok_per_s','binding_level','gateway','worker_http')})"
tail -5 gpurun_out/t7_bench_n$N.err
