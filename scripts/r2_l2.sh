#!/bin/bash
# does the k > 1 deficit come from the tileset overflowing L2?  k = 8 and k = 1 against 65 536 / 262 144 / 1 048 576 rows
O=gpurun_out; mkdir -p $O
for k in 8 1; do for tiles in 16384 65536 262144; do
python bench.py --workload c5_4k_2000f_256k_k8 --k $k --tiles $tiles --frames 8 --steps 3 --warmup 1 --no-kernels --no-e2e --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('k=$k rows', d['config']['dataset_rows'], 'coarse ms', round(d['phases_ms_per_step']['coarse'], 2), 'TF', round(d['roofline']['achieved'], 1), 'mhz', d['clocks']['sm_mhz'], 'fallback', d['fallback_queries_per_step'])"
done; done
