#!/bin/bash
# chained plan: operand megabytes per split (TM_TC_CHAIN=<MB>), C5 shape, one box
set -u
O=gpurun_out
mkdir -p $O
for mb in 48 24 16 32 64 96 130 48 0; do
  TM_TC_CHAIN=$mb timeout 300 python scripts/c5_bench.py --only k8 --queries 1048576 --reps 3 > $O/k8_chain.jsonl 2> $O/k8_chain.err
  echo "chain_mb=$mb $(cut -c60-330 $O/k8_chain.jsonl)"
done
