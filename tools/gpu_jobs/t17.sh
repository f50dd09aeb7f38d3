timeout 900 python bench.py > gpurun_out/t17_bench_default.json 2> gpurun_out/t17_bench_default.err
python -c "
import json; d=json.loads(open('gpurun_out/t17_bench_default.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('value','ms_per_step','decode','e2e','roofline','cpu_baseline','self_check','levels_tok_per_s','binding_level','clocks','gpu_launches')})"
tail -3 gpurun_out/t17_bench_default.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 2>&1 | tail -1 | cut -c 1-700
