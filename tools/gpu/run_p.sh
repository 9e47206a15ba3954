#!/bin/bash
# Pass P: L2 eviction hints on the streamed epilogue traffic (MODE 1 loads, MODE 3 a|b stores), N-fastest for the vocab
# projection's input gradient: tests, bench A/B, launch list with DRAM bytes.
set -u
O=gpurun_out/p
mkdir -p $O
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_model_gpu.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
for i in 1 2; do
  B200_GEMM_L2_HINTS=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_off$i.json 2> $O/bench_off$i.err; echo "bench off$i rc=$?" >> $O/status.txt
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_on$i.json 2> $O/bench_on$i.err; echo "bench on$i rc=$?" >> $O/status.txt
done
timeout 300 python tools/bench_gemm.py 2>&1 | grep "dX GEMM" > $O/gemm_dx.log
timeout 300 python tools/profile_step.py > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/status.txt
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file $O/launches.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
  python tools/profile_step.py > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest.log; cat $O/gemm_dx.log; tail -2 $O/plain.log
for f in bench_off1 bench_on1 bench_off2 bench_on2; do python - <<PY
import json
try:
    d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1])
    print('$f', {k:round(d[k],2) for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, round(d['e2e']['value']), d['clocks']['sm_mhz'])
except Exception as e: print('$f', e)
PY
done
