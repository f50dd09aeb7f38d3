timeout 300 python -m pytest tests/test_dense_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 400 python bench.py --preset qwen2.5-7b-bf16 --batch 32 --prompt-len 2048 --new-tokens 64 --steps 2 --warmup 1 --no-gateway --no-cpu-baseline --no-e2e > gpurun_out/t20_bench7b.json 2> gpurun_out/t20_bench7b.err
python -c "
import json; d=json.loads(open('gpurun_out/t20_bench7b.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('value','ms_per_step','decode','prefill','roofline')})"
tail -2 gpurun_out/t20_bench7b.err
