#!/bin/bash
# 8-GPU evidence run: C3 strong-scaled (default for --gpus 8), C5 (k = 8) on the 64-frame subsample, one process driving 8 GPUs
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu_multi$N.txt; nproc >> $O/gpu_multi$N.txt; free -g >> $O/gpu_multi$N.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_c3_n$N.json 2> $O/bench_c3_n$N.err; echo "rc=$?" >> $O/bench_c3_n$N.err
cut -c1-3500 $O/bench_c3_n$N.json; tail -3 $O/bench_c3_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload c5_4k_2000f_256k_k8 --steps 3 --warmup 2 > $O/bench_c5_n$N.json 2> $O/bench_c5_n$N.err; echo "rc=$?" >> $O/bench_c5_n$N.err
cut -c1-3000 $O/bench_c5_n$N.json; tail -3 $O/bench_c5_n$N.err
g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o /tmp/htb -ldl -lpthread && timeout 300 /tmp/htb tiler_b200/libtmann.so 16384 14400 $N > $O/htb_n$N.json 2> $O/htb_n$N.err
cat $O/htb_n$N.json
timeout 300 python -m pytest tests/test_gpu_multidev.py -m gpu -x -q 2>&1 | tail -3
