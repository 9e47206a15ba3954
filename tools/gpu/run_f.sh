#!/bin/bash
set -u
O=gpurun_out/f
mkdir -p $O
B200_ATTN_DS_TMEM=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $O/pytest_ds.log 2>&1; echo "pytest ds rc=$?" >> $O/status.txt
B200_ATTN_DKV_TMEM=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $O/pytest_dkv.log 2>&1; echo "pytest dkv rc=$?" >> $O/status.txt
for sh in "" hd128; do
echo "== base $sh" >> $O/attn_timing.log; timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
echo "== DS_TMEM $sh" >> $O/attn_timing.log; B200_ATTN_DS_TMEM=1 timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
echo "== DKV_TMEM $sh" >> $O/attn_timing.log; B200_ATTN_DKV_TMEM=1 timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
echo "== both $sh" >> $O/attn_timing.log; B200_ATTN_DS_TMEM=1 B200_ATTN_DKV_TMEM=1 timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
done
cat $O/status.txt; tail -3 $O/pytest_ds.log; tail -3 $O/pytest_dkv.log; cat $O/attn_timing.log
