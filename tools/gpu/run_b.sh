#!/bin/bash
# GPU pass B (1 x B200): full GPU test-suite, GPT-2 configs after the norm / fusion changes, 8B config.
set -u
O=gpurun_out/b
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-774m-lora --steps 5 > $O/cfg2_n1.json 2> $O/cfg2_n1.err; echo "cfg2 rc=$?" >> $O/status.txt
timeout 600 python tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5_n1.json 2> $O/cfg5_n1.err; echo "cfg5 rc=$?" >> $O/status.txt
timeout 300 python tools/profile_step.py --config gpt2-1.5b --layers 2 --shapes $O/gpt2_shapes.json > $O/gpt2_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/gpt2_launches.csv \
    python tools/profile_step.py --config gpt2-1.5b --layers 2 > $O/gpt2_ncu.log 2>&1; echo "ncu gpt2 rc=$?" >> $O/status.txt
cat $O/status.txt
tail -15 $O/pytest.log
cat $O/cfg4_n1.json $O/cfg2_n1.json $O/cfg5_n1.json
