#!/bin/bash
# Pass J: part of the attention exponentials on the FMA pipe (degree-3 polynomial): A/B of three column patterns.
set -u
O=gpurun_out/j
mkdir -p $O
L=building_llm_from_scratch_b200/libb200llm.so
cp $L /tmp/lib_default.so
for v in default A B C; do
  if [ $v != default ]; then cp build/alt/lib_$v.so $L; fi
  for sh in "" hd128; do
    echo "== $v $sh" >> $O/attn_timing.log; timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
  done
  timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $O/pytest_$v.log 2>&1; echo "pytest $v rc=$?" >> $O/status.txt
done
cp /tmp/lib_default.so $L
cat $O/status.txt; cat $O/attn_timing.log; tail -3 $O/pytest_A.log
