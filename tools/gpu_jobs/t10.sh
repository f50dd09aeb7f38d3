free -g | head -2
VGATE_B200_PREFILL_CHUNK=8192 VGATE_B200_MAX_STALL_MS=1000000 timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 2 --tp 2 --preset qwen2.5-14b-bf16 --batch 64 --prompt-len 2048 --new-tokens 64 --steps 2 --warmup 1 --no-cpu-baseline --no-gateway > gpurun_out/t10_bench14b_tp2.json 2> gpurun_out/t10_bench14b_tp2.err
tail -c 2500 gpurun_out/t10_bench14b_tp2.json; tail -8 gpurun_out/t10_bench14b_tp2.err
