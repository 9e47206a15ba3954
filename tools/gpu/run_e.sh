#!/bin/bash
set -u
O=gpurun_out/e
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
B200_ATTN_P_TMEM=1 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" > $O/pytest_ptmem.log 2>&1; echo "pytest ptmem rc=$?" >> $O/status.txt
timeout 300 python tools/bench_attention.py > $O/attn_timing.log 2>&1
B200_ATTN_P_TMEM=1 timeout 300 python tools/bench_attention.py >> $O/attn_timing.log 2>&1
timeout 300 python tools/bench_attention.py hd128 >> $O/attn_timing.log 2>&1
B200_ATTN_P_TMEM=1 timeout 300 python tools/bench_attention.py hd128 >> $O/attn_timing.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
cat $O/status.txt
tail -4 $O/pytest.log; tail -6 $O/pytest_ptmem.log
cat $O/attn_timing.log
cat $O/bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','value_full_head','eager_trainer_tokens_per_s')}); print(d['e2e']); print(d['roofline']['achieved'], d['roofline']['frac'])"
python -c "import json; d=json.load(open('$O/cfg4_n1.json')); print('cfg4', d['ms_per_step'], d['tokens_per_s'])"
