#!/bin/bash
# last check of the round: the whole GPU suite, smoke, C5 end to end (pipelined host batch), default bench
set -u
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -4 $O/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --workload c5_4k_2000f_256k_k8 --steps 2 --warmup 1 --no-kernels > $O/bench_c5.json 2> $O/bench_c5.err
python - <<PY
import json
d=json.load(open("$O/bench_c5.json"))
print("C5", "value %.3fM q/s" % (d["value"]/1e6), "e2e %.3fM" % (d["e2e"]["value"]/1e6), "coarse TF %.1f" % d["roofline"]["achieved"], d["phases_ms_per_step"], d["clocks"]["sm_mhz"], d.get("cpu_baseline",{}).get("parity_on_sample"))
PY
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
cut -c1-700 $O/bench.json
