#!/bin/bash
# where the coarse pass spends its time at C2: TM_TC_DBG = 1 MMA only, 2 + TMEM reads, 3 + fast path, 0 everything
for dbg in 1 2 3 0; do
  echo -n "dbg=$dbg "
  TM_TC_DBG=$dbg timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 4 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('coarse', round(d['phases_ms_per_step']['coarse'], 2), 'TF', round(d['roofline']['achieved'], 1), d['clocks']['sm_mhz'])"
done
