#!/bin/bash
# Multi-GPU pass (N x B200 on one box): usage  bash tools/gpu/run_multi.sh N [tests]
set -u
N=$1
O=gpurun_out/n$N
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/gpus.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "${2:-}" = "tests" ]; then
  timeout 1200 python -m pytest tests/test_ddp_gpu.py -m gpu -q > $O/pytest_ddp.log 2>&1; echo "pytest_ddp rc=$?" >> $O/status.txt
  tail -5 $O/pytest_ddp.log
fi
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 600 $TR --master-port 29512 tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4.json 2> $O/cfg4.err; echo "cfg4 rc=$?" >> $O/status.txt
timeout 900 $TR --master-port 29513 tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5.json 2> $O/cfg5.err; echo "cfg5 rc=$?" >> $O/status.txt
cat $O/status.txt
python - <<PY
import json
for f in ("bench", "cfg4", "cfg5"):
    try:
        d = json.loads(open("$O/%s.json" % f).read().strip().splitlines()[-1])
        keep = {k: d.get(k) for k in ("value", "ms_per_step", "tokens_per_s", "frac_of_burst_bf16_peak_per_gpu", "ranks_identical", "comm_exposed_ms", "data_parallel", "losses", "max_mem_gb") if k in d}
        print(f, keep)
    except Exception as e:
        print(f, "unreadable", e)
PY
tail -3 $O/bench.err $O/cfg4.err $O/cfg5.err 2>/dev/null | tail -20
