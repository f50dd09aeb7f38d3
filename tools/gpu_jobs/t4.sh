timeout 600 python -m pytest tests/test_umma_gpu.py tests/test_engine_gpu.py -m gpu -x -q -k "norm_tail or chunked" 2>&1 | tail -15
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/timeline.py > gpurun_out/t4_timeline.log 2>&1; head -16 gpurun_out/t4_timeline.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/t4_bench.json 2> gpurun_out/t4_bench.err
python -c "
import json; d=json.loads(open('gpurun_out/t4_bench.json').read().strip().splitlines()[-1]); print({k:d.get(k) for k in ('value','ms_per_step','decode','e2e','self_check','cpu_baseline','gpu_launches','levels_tok_per_s','binding_level','gateway')})"
tail -3 gpurun_out/t4_bench.err
timeout 600 python tools/bench_traffic.py --rate 150 --requests 500 > gpurun_out/t4_traffic.json 2> gpurun_out/t4_traffic.err; cat gpurun_out/t4_traffic.json; tail -3 gpurun_out/t4_traffic.err
