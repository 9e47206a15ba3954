#!/bin/bash
# Pass R: final 1-GPU validation: full GPU suite, smoke(), bench (default flags, with the CPU reference leg), reference arm,
# the other configs, launch list with DRAM bytes.
set -u
O=gpurun_out/r2
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-774m-lora --steps 5 > $O/cfg2_n1.json 2> $O/cfg2_n1.err; echo "cfg2 rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
timeout 600 python tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5_n1.json 2> $O/cfg5_n1.err; echo "cfg5 rc=$?" >> $O/status.txt
timeout 200 python tools/bench_attention.py > $O/attn_timing.log 2>&1; timeout 200 python tools/bench_attention.py hd128 >> $O/attn_timing.log 2>&1
timeout 300 python tools/profile_step.py --shapes $O/shapes.json > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/status.txt
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file $O/launches.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
  python tools/profile_step.py > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest.log; tail -2 $O/smoke.log; cat $O/attn_timing.log; tail -1 $O/plain.log
python - <<PY
import json
d=json.loads(open('$O/bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s','gpu_launches')}, d['e2e']['value'], d['roofline']['achieved'], d['roofline']['frac'], d['cpu_baseline'], d['clocks'])
print(open('$O/bench_ref.json').read()[:700])
for f in ('cfg2_n1','cfg4_n1','cfg5_n1'):
    try:
        c=json.load(open('$O/%s.json'%f)); print(f, round(c['ms_per_step'],2), round(c['tokens_per_s']), round(c['frac_of_burst_bf16_peak_per_gpu'],3))
    except Exception as e: print(f, e)
PY
