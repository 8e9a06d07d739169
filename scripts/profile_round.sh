#!/bin/bash
# Round evidence run (one B200): tests, smoke, both bench arms, ncu launch list of the bench command,
# ncu --set full captures of the search kernel and of the helper / KModes / smoothing kernels.  Outputs -> gpurun_out/.
# Every ncu pass runs only after the same command has exited 0 without ncu.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu.txt; nproc >> $O/gpu.txt
timeout 600 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 400 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 400 python bench.py --kernels --no-e2e --no-cpu-baseline --steps 3 > $O/bench_kernels.json 2> $O/bench_kernels.err
timeout 400 python scripts/c5_bench.py > $O/c5_bench.json 2> $O/c5_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc2_search_kernel -s 3 -c 1 \
    -o $O/tc2_search_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $O/ncu_tc2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:"psyv_queries_kernel|rerank_kernel" -s 6 -c 2 \
    -o $O/helpers_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $O/ncu_helpers.log 2>&1
timeout 600 python scripts/kernels_bench.py --reps 1 --only kmodes > $O/kmodes_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"kmodes_assign_kernel" -c 1 \
    -o $O/kmodes_full -f python scripts/kernels_bench.py --reps 1 --only kmodes > $O/ncu_kmodes.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"temporal_smoothing_kernel" -c 1 \
    -o $O/smooth_full -f python scripts/c5_bench.py --only smooth > $O/ncu_smooth.log 2>&1
tail -2 $O/pytest_gpu.log; cat $O/smoke.log; cut -c1-600 $O/bench.json; tail -2 $O/bench.err; cat $O/c5_bench.json
