timeout 900 python -m pytest tests/ -m gpu -x -q 2>&1 | tail -8
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py --preset qwen2.5-7b-bf16 --batch 32 --prompt-len 2048 --new-tokens 64 --steps 2 --warmup 2 --no-gateway --no-cpu-baseline --no-e2e > gpurun_out/t16_bench7b.json 2> gpurun_out/t16_bench7b.err
python -c "
import json; d=json.loads(open('gpurun_out/t16_bench7b.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('value','ms_per_step','decode','prefill','roofline')})"
tail -3 gpurun_out/t16_bench7b.err
