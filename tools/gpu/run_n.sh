#!/bin/bash
# Pass N: programmatic dependent launch for the attention and RMSNorm kernels too: tests, bench A/B on one box.
set -u
O=gpurun_out/n
mkdir -p $O
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_engine_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
for i in 1 2; do
  B200_PDL_WIDE=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_off$i.json 2> $O/bench_off$i.err; echo "bench off$i rc=$?" >> $O/status.txt
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_on$i.json 2> $O/bench_on$i.err; echo "bench on$i rc=$?" >> $O/status.txt
done
cat $O/status.txt; tail -3 $O/pytest.log
for f in bench_off1 bench_on1 bench_off2 bench_on2; do python - <<PY
import json
d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1])
print('$f', {k:round(d[k],2) for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, round(d['e2e']['value']), d['clocks']['sm_mhz'])
PY
done
