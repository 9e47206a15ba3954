#!/bin/bash
# Round-2 GPU pass A (1 x B200): tests, smoke, bench, reference-on-GPU, other configs, launch lists.
set -u
O=gpurun_out/a
mkdir -p $O
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
nproc > $O/nproc.txt; free -g >> $O/nproc.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/status.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/status.txt
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 300 python tools/reference_on_gpu.py --steps 5 --warmup 2 > $O/ref_gpu.json 2> $O/ref_gpu.err; echo "ref_gpu rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_n1.json 2> $O/cfg4_n1.err; echo "cfg4 rc=$?" >> $O/status.txt
timeout 300 python tools/run_config.py gpt2-774m-lora --steps 5 > $O/cfg2_n1.json 2> $O/cfg2_n1.err; echo "cfg2 rc=$?" >> $O/status.txt
timeout 600 python tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5_n1.json 2> $O/cfg5_n1.err; echo "cfg5 rc=$?" >> $O/status.txt
timeout 300 python tools/profile_step.py --config gpt2-1.5b --layers 2 --shapes $O/gpt2_shapes.json > $O/gpt2_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/gpt2_launches.csv \
    python tools/profile_step.py --config gpt2-1.5b --layers 2 > $O/gpt2_ncu.log 2>&1; echo "ncu gpt2 rc=$?" >> $O/status.txt
timeout 300 python tools/profile_step.py --layers 2 --shapes $O/llama_shapes.json > $O/llama_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file $O/llama_launches.csv \
    python tools/profile_step.py --layers 2 > $O/llama_ncu.log 2>&1; echo "ncu llama rc=$?" >> $O/status.txt
cat $O/status.txt
tail -3 $O/pytest.log
