#!/bin/bash
# call-pattern experiments (native threads): combiner on / off, trace of the batches
set -u
O=gpurun_out
mkdir -p $O
g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o /tmp/htb -ldl -lpthread || exit 1
timeout 300 /tmp/htb tiler_b200/libtmann.so > $O/htb_on.json 2> $O/htb_on.err
TM_COMBINE=0 timeout 300 /tmp/htb tiler_b200/libtmann.so > $O/htb_off.json 2> $O/htb_off.err
TM_TRACE=1 timeout 300 /tmp/htb tiler_b200/libtmann.so > /dev/null 2> $O/htb_trace.err
timeout 300 /tmp/htb tiler_b200/libtmann.so 65536 32400 > $O/htb_1080.json 2> $O/htb_1080.err
cat $O/htb_on.json; echo ---; cat $O/htb_off.json; echo ---; cat $O/htb_1080.json; grep -c batch $O/htb_trace.err
