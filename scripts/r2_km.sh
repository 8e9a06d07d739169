#!/bin/bash
# KModes handle: tests, then one KModesIter over C4 with the polled completion (default) and the copy + synchronise one
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_kmodes_loop.py tests/test_zz_no_alloc.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_km.log 2>&1; echo "pytest rc=$?" >> $O/pytest_km.log
tail -4 $O/pytest_km.log
for p in 1 0 1 0; do TM_KM_POLL=$p timeout 300 python scripts/km_iter_bench.py 4; done
TM_KM_POLL=1 timeout 300 python scripts/km_iter_bench.py 24
TM_KM_POLL=1 timeout 300 python scripts/km_iter_bench.py 64
