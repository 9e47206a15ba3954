#!/bin/bash
set -u
O=gpurun_out/d
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python tools/bench_attention.py > $O/attn_timing.log 2>&1
timeout 300 python tools/bench_attention.py hd128 >> $O/attn_timing.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
cat $O/status.txt
tail -6 $O/pytest.log
cat $O/attn_timing.log
cat $O/bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}); print(d['e2e']); print([ (o['kernel'][:40], o['us'], o['frac']) for o in d['roofline']['other_kernels']])"
cat $O/cfg4_n1.json
