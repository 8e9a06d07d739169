#!/bin/bash
# quick round-2 validation on ONE B200: GPU tests, smoke, the default bench line, the C5 line (short)
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu.txt; nproc >> $O/gpu.txt; free -g >> $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 600 python bench.py --workload c5_4k_2000f_256k_k8 --steps 2 --warmup 1 --frames 16 --no-kernels > $O/bench_c5.json 2> $O/bench_c5.err; echo "rc=$?" >> $O/bench_c5.err
timeout 300 python bench.py --workload c4_kmodes_1m_16k --steps 3 > $O/bench_c4.json 2> $O/bench_c4.err; echo "rc=$?" >> $O/bench_c4.err
tail -15 $O/pytest_gpu.log; cat $O/smoke.log; cut -c1-3000 $O/bench.json; tail -3 $O/bench.err; cut -c1-1500 $O/bench_c5.json; tail -3 $O/bench_c5.err; cut -c1-1200 $O/bench_c4.json; tail -3 $O/bench_c4.err
