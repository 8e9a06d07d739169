#!/bin/bash
# chained (split-major) plan: tests, then C3 (k = 1, 262 144 rows) and C5 (k = 8, 1 048 576 rows) with and without it
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_fullsize.py -m gpu -x -q > $O/pytest_tc.log 2>&1; echo "pytest rc=$?" >> $O/pytest_tc.log
tail -4 $O/pytest_tc.log
for c in 1 0 1 0; do
  TM_TC_CHAIN_K1=$c timeout 600 python bench.py --workload c3_1080p_1000f_64k --steps 2 --warmup 1 --frames 200 --no-kernels --no-e2e --no-extras > $O/c3_chain$c.json 2> $O/c3_chain$c.err
  python - <<PY
import json
d=json.load(open("$O/c3_chain$c.json"))
print("C3 chain_k1=$c", "value %.3fM q/s" % (d["value"]/1e6), "coarse TF %.1f" % d["roofline"]["achieved"], d["phases_ms_per_step"], "fallback", d["fallback_queries_per_step"], d["clocks"]["sm_mhz"], d.get("cpu_baseline",{}).get("parity_on_sample"), d["error_model_audit"])
PY
done
