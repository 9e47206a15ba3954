#!/bin/bash
# Pass L: generalised tail K split (L*s units over several rounds) + N-fastest for B up to 96 MB: tests, shapes, bench.
set -u
O=gpurun_out/l
mkdir -p $O
timeout 900 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > $O/pytest_gemm.log 2>&1; echo "pytest gemm rc=$?" >> $O/status.txt
B200_GEMM_TAIL=0 B200_GEMM_ORDER=m timeout 300 python tools/bench_gemm.py 2>&1 | grep "dX GEMM" > $O/gemm_old.log
timeout 300 python tools/bench_gemm.py 2>&1 | grep "dX GEMM" > $O/gemm_new.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 1200 python -m pytest tests -m gpu -x -q --deselect tests/test_gemm_gpu.py > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest_gemm.log; tail -3 $O/pytest.log
echo "--- tail split off, M fastest"; cat $O/gemm_old.log; echo "--- new defaults"; cat $O/gemm_new.log
python - <<PY
import json
d=json.loads(open('$O/bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'])
for o in d['roofline']['other_kernels']: print(o['kernel'][:70], o['us'], o['frac'])
PY
