export PRESET=qwen2.5-7b-bf16-l2 N_SEQS=32 PROMPT=256
python tools/ncu_engine_target.py > gpurun_out/t19_target_plain.log 2>&1 &&
timeout 500 ncu --set full --clock-control none -o /tmp/r02_dense_full python tools/ncu_engine_target.py > gpurun_out/t19_ncu_full.log 2>&1
echo "full rc $?"; tail -2 gpurun_out/t19_target_plain.log
python tools/ncu_summarize.py /tmp/r02_dense_full.ncu-rep > gpurun_out/r02_dense_kernels_ncu_full.txt 2> gpurun_out/t19_summarize.err; wc -l gpurun_out/r02_dense_kernels_ncu_full.txt
