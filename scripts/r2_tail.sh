#!/bin/bash
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -6 $O/pytest_gpu.log
for t in 1 0; do
TM_TC_TAIL=$t timeout 300 python bench.py --workload c1_320x240_64f_4k --steps 20 --warmup 5 --no-cpu-baseline --no-kernels --no-extras > $O/c1_tail$t.json 2>/dev/null
TM_TC_TAIL=$t timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-extras --no-e2e > $O/c2_tail$t.json 2>$O/c2_tail$t.err
python - <<PY
import json
a=json.load(open("$O/c1_tail$t.json")); b=json.load(open("$O/c2_tail$t.json"))
print("tail=$t C1 %.1fM q/s e2e %.1fM | C2 %.2fM coarse %.2f ms TF %.1f mhz %s" % (a["value"]/1e6, a["e2e"]["value"]/1e6, b["value"]/1e6, b["phases_ms_per_step"]["coarse"], b["roofline"]["achieved"], b["clocks"]["sm_mhz"]))
print(json.dumps(b["kernels"]["kmodes_iter_c4_handle"]))
PY
done
g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o /tmp/htb -ldl -lpthread || exit 1
timeout 300 /tmp/htb tiler_b200/libtmann.so 2>/dev/null | grep frame_calls
TM_TC_TAIL=0 timeout 300 /tmp/htb tiler_b200/libtmann.so 2>/dev/null | grep frame_calls
