#!/bin/bash
# decode step time for several split-K settings (qkv,o,gate_up,down); run on the GPU box
for sp in "4,4,1,12" "2,4,1,12" "4,2,1,12" "4,4,1,6" "4,4,1,8" "3,3,1,12" "6,6,1,12" "2,2,1,6"; do
  VGATE_B200_SPLITS=$sp timeout 120 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$sp', round(d['decode']['tpot_ms_mean'],4), 'ms/step')"
done
