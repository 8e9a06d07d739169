#!/bin/bash
# chained (split-major) plan of the k > 1 search: tests, then C5 with TM_TC_CHAIN=0 / default, alternating on one box
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tc.py -m gpu -x -q > $O/pytest_tc.log 2>&1; echo "pytest rc=$?" >> $O/pytest_tc.log
tail -4 $O/pytest_tc.log
for c in 1 0 1 0; do
  TM_TC_CHAIN=$c timeout 300 python scripts/c5_bench.py --only k8 --queries 1048576 --reps 3 > $O/k8_chain.jsonl 2> $O/k8_chain.err
  echo "chain=$c $(cut -c1-330 $O/k8_chain.jsonl) $(tail -1 $O/k8_chain.err | cut -c1-200)"
done
for c in 1 0; do
  TM_TC_CHAIN=$c timeout 600 python bench.py --workload c5_4k_2000f_256k_k8 --steps 2 --warmup 1 --frames 32 --no-kernels --no-e2e --no-extras > $O/c5_chain$c.json 2> $O/c5_chain$c.err
  python - <<PY
import json
d=json.load(open("$O/c5_chain$c.json"))
print("chain=$c", "value %.3fM q/s" % (d["value"]/1e6), "coarse TF %.1f" % d["roofline"]["achieved"], d["phases_ms_per_step"], "fallback", d["fallback_queries_per_step"], d["clocks"]["sm_mhz"], d.get("cpu_baseline",{}).get("parity_on_sample"), d["error_model_audit"])
PY
done
