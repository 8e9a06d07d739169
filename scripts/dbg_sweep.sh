#!/bin/bash
# tuning sweep: where does the per-item overhead of the coarse pass come from (frame-sized batches)
for dbg in 0 3 2 1; do for n in 1 8; do
  echo "dbg=$dbg nsplit=$n"
  TM_TC_DBG=$dbg TM_TC_NSPLIT=$n timeout 300 python scripts/frame_call_bench.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('  ', d['case'], 'coarse', round(d['device_ms_per_call']['coarse_ms'], 4))"
done; done
