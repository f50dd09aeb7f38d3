# ncu evidence for round 2: launch list of the bench command, and --set full of every kernel of the path
python bench.py --steps 1 --warmup 1 --no-graphs --no-e2e --no-cpu-baseline --no-gateway > gpurun_out/t11_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 504 -c 700 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 1 --warmup 1 --no-graphs --no-e2e --no-cpu-baseline --no-gateway > gpurun_out/t11_ncu_launches.log 2>&1
echo "launch list rc $?"; wc -l gpurun_out/r02_launches.csv
python tools/ncu_engine_target.py > gpurun_out/t11_target_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -o gpurun_out/r02_engine_full python tools/ncu_engine_target.py > gpurun_out/t11_ncu_full.log 2>&1
echo "full rc $?"; ls -la gpurun_out/r02_engine_full.ncu-rep
for B in 8 128 256; do
  timeout 600 python bench.py --batch $B --steps 3 --warmup 3 --no-gateway --no-cpu-baseline > gpurun_out/t11_bench_b$B.json 2> gpurun_out/t11_bench_b$B.err
  python -c "
import json; d=json.loads(open('gpurun_out/t11_bench_b$B.json').read().strip().splitlines()[-1]); print($B, {k:d.get(k) for k in ('value','ms_per_step','decode')}, d['e2e']['value'])"
done
