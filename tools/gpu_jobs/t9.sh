timeout 900 python -m pytest tests/test_dense_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 1500 python bench.py --preset qwen2.5-7b-bf16 --batch 32 --prompt-len 2048 --new-tokens 64 --steps 2 --warmup 2 --no-gateway --no-cpu-baseline --no-e2e > gpurun_out/t9_bench7b.json 2> gpurun_out/t9_bench7b.err
python -c "
import json; d=json.loads(open('gpurun_out/t9_bench7b.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('value','ms_per_step','decode','prefill','roofline')})"
tail -3 gpurun_out/t9_bench7b.err
timeout 600 python tools/bench_traffic.py --rate 60 --requests 300 --long-frac 0.15 --long-len 1536 --chunks 512,8192 > gpurun_out/t9_traffic_long.json 2> gpurun_out/t9_traffic_long.err; cat gpurun_out/t9_traffic_long.json; tail -2 gpurun_out/t9_traffic_long.err
