#!/bin/bash
# GPU pass C (1 x B200): tests, bench, GPT-2 configs, and ncu --set full of the three attention kernels.
set -u
O=gpurun_out/c
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-774m-lora --steps 5 > $O/cfg2_n1.json 2> $O/cfg2_n1.err; echo "cfg2 rc=$?" >> $O/status.txt
timeout 300 python tools/profile_step.py --config gpt2-774m-lora --layers 2 > $O/gpt2l_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/gpt2l_launches.csv \
    python tools/profile_step.py --config gpt2-774m-lora --layers 2 > $O/gpt2l_ncu.log 2>&1; echo "ncu gpt2l rc=$?" >> $O/status.txt
timeout 300 python tools/bench_attention.py > $O/attn_timing.log 2>&1
timeout 300 python tools/bench_attention.py once > $O/attn_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_ -c 4 -o $O/attn_full \
    python tools/bench_attention.py once > $O/attn_ncu.log 2>&1; echo "ncu attn rc=$?" >> $O/status.txt
cat $O/status.txt
tail -12 $O/pytest.log
cat $O/bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}); print(d['e2e'])"
cat $O/cfg4_n1.json $O/cfg2_n1.json $O/attn_timing.log
