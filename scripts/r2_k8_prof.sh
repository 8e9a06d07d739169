#!/bin/bash
# one ncu --set full capture with source correlation of the k = 8 search kernel (after the same command has exited 0 without ncu)
set -u
O=gpurun_out
mkdir -p $O
TAG=${1:-k8_src}
timeout 300 python bench.py --workload c5_4k_2000f_256k_k8 --frames 2 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernels > $O/k8_plain.json 2> $O/k8_plain.err; echo "rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc2_search_kernel -s 1 -c 1 \
    -o $O/tc2_search_$TAG -f python bench.py --workload c5_4k_2000f_256k_k8 --frames 2 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernels > $O/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; ls -la $O | tail -5
