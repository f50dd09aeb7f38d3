timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -x -q 2>&1 | tail -8
timeout 900 python tools/tp_timeline.py 64 2048 2 4 2>&1 | tail -28
VGATE_B200_TP_ALLREDUCE=nccl timeout 900 python tools/tp_timeline.py 64 2048 2 4 2>&1 | grep -E "^tp|decode step"
