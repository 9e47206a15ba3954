#!/bin/bash
# Pass O: grouped skinny GEMMs on K-splitting CTA pairs (DSMEM reduction): tests, in-step cost, bench A/B; PDL-wide tests.
set -u
O=gpurun_out/o
mkdir -p $O
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > $O/pytest_gemm.log 2>&1; echo "pytest gemm rc=$?" >> $O/status.txt
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gemm_gpu.py > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
echo "== pair off" > $O/skinny.log; B200_GEMM_SKINNY_PAIR=0 timeout 200 python tools/bench_skinny_instep.py >> $O/skinny.log 2>&1
echo "== pair on" >> $O/skinny.log; timeout 200 python tools/bench_skinny_instep.py >> $O/skinny.log 2>&1
for i in 1 2; do
  B200_GEMM_SKINNY_PAIR=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_off$i.json 2> $O/bench_off$i.err; echo "bench off$i rc=$?" >> $O/status.txt
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_on$i.json 2> $O/bench_on$i.err; echo "bench on$i rc=$?" >> $O/status.txt
done
cat $O/status.txt; tail -4 $O/pytest_gemm.log; tail -3 $O/pytest.log; cat $O/skinny.log
for f in bench_off1 bench_on1 bench_off2 bench_on2; do python - <<PY
import json
try:
    d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1])
    print('$f', {k:round(d[k],2) for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, round(d['e2e']['value']), d['clocks']['sm_mhz'], d['final_loss'])
except Exception as e: print('$f', e)
PY
done
