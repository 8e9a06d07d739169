#!/bin/bash
# multi-GPU validation (run with gpurun --gpus N): one-process multi-device tests, the strong-scaled C3 bench line under torchrun
set -u
N=${1:-2}
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu_multi.txt; nproc >> $O/gpu_multi.txt; free -g >> $O/gpu_multi.txt
timeout 600 python -m pytest tests/test_gpu_multidev.py -m gpu -x -q > $O/pytest_multidev.log 2>&1; echo "pytest rc=$?" >> $O/pytest_multidev.log
tail -5 $O/pytest_multidev.log
g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o /tmp/htb -ldl -lpthread && timeout 300 /tmp/htb tiler_b200/libtmann.so 16384 14400 $N > $O/htb_n$N.json 2> $O/htb_n$N.err
cat $O/htb_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?" >> $O/bench_n$N.err
cut -c1-4000 $O/bench_n$N.json; tail -5 $O/bench_n$N.err
