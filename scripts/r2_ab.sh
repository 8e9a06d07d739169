#!/bin/bash
# same-box A/B of two builds of the library: tiler_b200/libtmann.so (new) vs tiler_b200/libtmann_ab.so (baseline)
set -u
O=gpurun_out
mkdir -p $O
AB=$PWD/tiler_b200/libtmann_ab.so
for rep in 1 2; do
for lib in new old; do
  if [ $lib = old ]; then export TMANN_LIB=$AB; else unset TMANN_LIB; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-kernels --no-e2e --no-extras --no-cpu-baseline > $O/c2_$lib.json 2> $O/c2_$lib.err
  python - <<PY
import json
d=json.load(open("$O/c2_$lib.json"))
print("C2 $lib", "value %.3fM q/s" % (d["value"]/1e6), "coarse TF %.1f" % d["roofline"]["achieved"], d["phases_ms_per_step"], d["clocks"]["sm_mhz"])
PY
  timeout 300 python scripts/c5_bench.py --only k8 --queries 524288 --reps 3 > $O/k8_ab.jsonl 2> $O/k8_ab.err
  echo "k8 $lib $(cut -c90-330 $O/k8_ab.jsonl)"
done
done
