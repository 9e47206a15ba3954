#!/bin/bash
# Pass G: full GPU suite + bench with the TMEM-operand attention defaults; A/B of the mbarrier suspend-time hint
# (build/alt/libb200llm_suspend.so); ncu --set full of the three attention kernels (default library).
set -u
O=gpurun_out/g
mkdir -p $O
L=building_llm_from_scratch_b200/libb200llm.so
cp $L /tmp/lib_default.so
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
for sh in "" hd128; do
  echo "== default $sh" >> $O/attn_timing.log; timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
# --- suspend-hint library
cp build/alt/libb200llm_suspend.so $L
for sh in "" hd128; do
  echo "== suspend $sh" >> $O/attn_timing.log; timeout 200 python tools/bench_attention.py $sh >> $O/attn_timing.log 2>&1
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_suspend.json 2> $O/bench_suspend.err; echo "bench suspend rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1_suspend.json 2> $O/cfg4_n1_suspend.err; echo "cfg4 suspend rc=$?" >> $O/status.txt
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_gemm_gpu.py -m gpu -x -q > $O/pytest_suspend.log 2>&1; echo "pytest suspend rc=$?" >> $O/status.txt
cp /tmp/lib_default.so $L
# --- ncu of the attention kernels (the plain command ran above and exited 0)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_ -c 6 -o $O/attn_final python tools/bench_attention.py once > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt
tail -4 $O/pytest.log; tail -3 $O/pytest_suspend.log
cat $O/attn_timing.log
for f in bench bench_suspend; do python - <<PY
import json
d=json.loads(open('$O/$f.json').read().strip().splitlines()[-1])
print('$f', {k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'])
PY
done
for f in cfg4_n1 cfg4_n1_suspend; do python -c "import json; d=json.load(open('$O/$f.json')); print('$f', d['ms_per_step'], d['tokens_per_s'])"; done
