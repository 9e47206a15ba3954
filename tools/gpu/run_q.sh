#!/bin/bash
set -u
O=gpurun_out/q
mkdir -p $O
timeout 200 python tools/bench_gelu_gemm.py > $O/timing.log 2>&1; echo "timing rc=$?" >> $O/status.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05_2cta -c 2 -o $O/gelu python tools/bench_gelu_gemm.py once > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; cat $O/timing.log
