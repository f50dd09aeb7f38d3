timeout 900 python tools/tp_timeline.py 64 2048 2 4 2>&1 | grep -E "^tp|^warm|decode step"
timeout 900 python tools/tp_timeline.py 64 2048 2 16 2>&1 | grep -E "^tp|^warm|decode step"
