#!/bin/bash
# k > 1 kernel A/B on one B200: tensor-core tests, then the C5 bench line with the in-line and the deferred candidate tracking
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest_tc.log 2>&1; echo "pytest rc=$?" >> $O/pytest_tc.log
tail -4 $O/pytest_tc.log
for d in 1 0 1 0; do
  TM_TC_DEFER=$d timeout 600 python bench.py --workload c5_4k_2000f_256k_k8 --steps 3 --warmup 1 --frames 16 --no-kernels --no-e2e --no-cpu-baseline --no-extras > $O/c5_defer$d.json 2> $O/c5_defer$d.err
  python - <<PY
import json
d=json.load(open("$O/c5_defer$d.json"))
print("defer=$d", "value %.3fM q/s" % (d["value"]/1e6), "coarse TF %.1f" % d["roofline"]["achieved"], d["phases_ms_per_step"], "fallback", d["fallback_queries_per_step"], d["clocks"]["sm_mhz"])
PY
done
