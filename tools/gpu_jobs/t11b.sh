python bench.py --steps 1 --warmup 1 --no-graphs --no-e2e --no-cpu-baseline --no-gateway > gpurun_out/t11_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 504 -c 700 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 1 --warmup 1 --no-graphs --no-e2e --no-cpu-baseline --no-gateway > gpurun_out/t11_ncu_launches.log 2>&1
echo "launch list rc $?"; wc -l gpurun_out/r02_launches.csv
python tools/ncu_engine_target.py > gpurun_out/t11_target_plain.log 2>&1 &&
ncu --set full --clock-control none -o /tmp/r02_engine_full python tools/ncu_engine_target.py > gpurun_out/t11_ncu_full.log 2>&1
echo "full rc $?"
python tools/ncu_summarize.py /tmp/r02_engine_full.ncu-rep > gpurun_out/r02_kernels_ncu_full.txt 2> gpurun_out/t11_summarize.err; wc -l gpurun_out/r02_kernels_ncu_full.txt
du -sh gpurun_out
