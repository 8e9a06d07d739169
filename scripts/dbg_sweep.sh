#!/bin/bash
# Developer probe: time the coarse search kernel with its epilogue stages disabled (TM_TC_DBG):
# 0 = full kernel, 1 = TMA + MMA pipeline only (no TMEM reads), 2 = + TMEM reads, no compare,
# 3 = + group minima and votes, candidate slow path never taken.
# Results are wrong in modes 1/2 by design; only coarse_ms is meaningful.
for d in ${MODES:-0 1 2 3}; do
  TM_TC_DBG=$d timeout 200 python bench.py --frames ${FRAMES:-37} --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > /tmp/dbg_$d.log 2>&1
  python - "$d" <<'PY'
import sys, json
d = sys.argv[1]
for l in open(f"/tmp/dbg_{d}.log"):
    if l.startswith("{"):
        j = json.loads(l)
        print("dbg", d, j["phases_ms_per_step"], "TFLOP/s", j["roofline"]["achieved"], j["clocks"])
PY
done
