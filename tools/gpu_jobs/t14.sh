timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -x -q 2>&1 | tail -4
timeout 900 python tools/tp_timeline.py 64 2048 2 4 2>&1 | tail -36 | grep -E "^tp|phases|rmsnorm|decode step"
