#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
timeout 600 python bench.py --steps 3 --no-cpu-baseline --no-e2e > $O/bench_k.json 2> $O/bench_k.err; echo "rc=$?"
python - <<PY
import json
d=json.load(open("$O/bench_k.json"))
print(json.dumps(d.get("kernels"))[:2500]); print(json.dumps(d.get("call_patterns")))
PY
tail -3 $O/bench_k.err
