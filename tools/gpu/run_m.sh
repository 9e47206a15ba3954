#!/bin/bash
# Pass M: lse2 / delta formed inside the dQ kernel (no attn_delta launch): tests both ways, timings, bench.
set -u
O=gpurun_out/m
mkdir -p $O
timeout 900 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > $O/pytest_gemm.log 2>&1; echo "pytest gemm rc=$?" >> $O/status.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -m gpu -x -q > $O/pytest_fold.log 2>&1; echo "pytest fold rc=$?" >> $O/status.txt
B200_ATTN_DELTA_FOLD=0 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k attention > $O/pytest_nofold.log 2>&1; echo "pytest nofold rc=$?" >> $O/status.txt
for sh in "" hd128; do
  echo "== nofold $sh" >> $O/attn_timing.log; B200_ATTN_DELTA_FOLD=0 timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
  echo "== fold $sh" >> $O/attn_timing.log; timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest_gemm.log; tail -3 $O/pytest_fold.log; tail -3 $O/pytest_nofold.log; cat $O/attn_timing.log
python - <<PY
import json
d=json.loads(open('$O/bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'])
PY
