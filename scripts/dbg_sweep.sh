#!/bin/bash
# where the coarse pass spends its time: TM_TC_DBG = 1 MMA + TMA only, 2 + TMEM reads, 3 + fast path, 0 everything.
# Needs the debug-sweep build of the library (the release build has no such switch):
#   make -C tiler_b200/csrc DEFS=-DTM_DEBUG_SWEEP B=build_dbg OUT=../libtmann_dbg.so
# usage: scripts/dbg_sweep.sh [bench.py arguments, e.g. --workload c5_4k_2000f_256k_k8 --frames 16]
export TMANN_LIB=$PWD/tiler_b200/libtmann_dbg.so
for dbg in 1 2 3 0; do
  echo -n "dbg=$dbg "
  TM_TC_DBG=$dbg timeout 300 python bench.py --no-cpu-baseline --no-e2e --no-kernels --no-extras --steps 3 --warmup 1 "$@" 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('coarse', round(d['phases_ms_per_step']['coarse'], 2), 'TF', round(d['roofline']['achieved'], 1), d['clocks']['sm_mhz'])"
done
