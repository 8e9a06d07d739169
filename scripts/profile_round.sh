#!/bin/bash
# Round evidence run (one B200): tests, smoke, both bench arms, the other workloads, ncu launch list of the bench command,
# ncu --set full captures of the search kernel (k = 1 and k = 8) and of the helper kernels.  Outputs -> gpurun_out/.
# Every ncu pass runs only after the same command has exited 0 without ncu.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu.txt; nproc >> $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/pytest_zz_alone.log 2>&1; echo "pytest rc=$?" >> $O/pytest_zz_alone.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 600 python bench.py --workload c3_1080p_1000f_64k --steps 2 --warmup 1 --no-kernels --no-extras > $O/bench_c3.json 2> $O/bench_c3.err
timeout 600 python bench.py --workload c5_4k_2000f_256k_k8 --steps 2 --warmup 1 --no-kernels > $O/bench_c5.json 2> $O/bench_c5.err
timeout 300 python bench.py --workload c4_kmodes_1m_16k > $O/bench_c4.json 2> $O/bench_c4.err
timeout 300 python bench.py --workload c1_320x240_64f_4k --steps 20 --warmup 5 --no-kernels --no-extras > $O/bench_c1.json 2> $O/bench_c1.err
g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o /tmp/htb -ldl -lpthread && timeout 300 /tmp/htb tiler_b200/libtmann.so > $O/htb.json 2> $O/htb.err
timeout 300 /tmp/htb tiler_b200/libtmann.so 65536 32400 > $O/htb_1080.json 2> $O/htb_1080.err
timeout 300 python scripts/km_iter_bench.py 4 > $O/km_iter.json 2> $O/km_iter.err
# launch list of the bench command, restricted to the library's kernels (+ the CUB sort it calls): the synthetic
# inputs come from the reference LCG evaluated with torch, whose thousands of small generator launches would fill -c 400
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_lib.csv \
    -k regex:"tc2_search|psyv_|rerank_|query_key|invert_order|scatter_|DeviceRadixSort|operand_kernel|gather_rows|search_simt|scan_order|kmodes_" \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernels --no-extras > $O/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc2_search_kernel -s 3 -c 1 \
    -o $O/tc2_search_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-kernels --no-extras > $O/ncu_tc2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc2_search_kernel -s 1 -c 1 \
    -o $O/tc2_search_k8_full -f python bench.py --workload c5_4k_2000f_256k_k8 --frames 8 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-kernels > $O/ncu_tc2_k8.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on \
    -k regex:"psyv_queries_kernel|rerank_kernel" -s 6 -c 2 \
    -o $O/helpers_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-kernels --no-extras > $O/ncu_helpers.log 2>&1
tail -3 $O/pytest_gpu.log; tail -3 $O/pytest_zz_alone.log; cat $O/smoke.log; cut -c1-400 $O/bench.json; tail -2 $O/bench.err; ls -la $O
