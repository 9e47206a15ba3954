#!/bin/bash
# Pass H: ncu --set full of the skinny (rank-64) GEMM groups, one launch per group.
set -u
O=gpurun_out/h
mkdir -p $O
timeout 200 python tools/bench_skinny.py once > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/status.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_streamk|gemm_grouped" -c 8 -o $O/skinny python tools/bench_skinny.py once > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/ncu.log
