timeout 900 python tools/tp_timeline.py 64 2048 2 2>&1 | tail -30
timeout 900 python tools/tp_timeline.py 64 2048 1 2>&1 | tail -30
