#!/bin/bash
# where the k = 8 coarse pass spends its time on the C5 shape (debug-sweep build,
#   make -C tiler_b200/csrc DEFS=-DTM_DEBUG_SWEEP B=build_dbg OUT=../libtmann_dbg.so):
# TM_TC_DBG = 1 MMA + TMA only, 2 + TMEM reads, 3 + fast path, 0 everything; k = 1 on the same shape beside it
set -u
O=gpurun_out
mkdir -p $O
export TMANN_LIB=$PWD/tiler_b200/libtmann_dbg.so
for dbg in 1 2 3 0; do
  for only in k8 k1; do
    TM_TC_DBG=$dbg timeout 300 python scripts/c5_bench.py --only $only --queries 524288 --reps 2 2> $O/sweep.err | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('dbg=$dbg', d['case'], 'coarse', round(d['phases_ms']['coarse'], 2), 'ms', round(d['tflops'], 1), 'TF (whole call)')"
  done
done
