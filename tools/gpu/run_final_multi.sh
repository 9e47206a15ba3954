#!/bin/bash
# Final multi-GPU lines: usage  bash tools/gpu/run_final_multi.sh N [cfg5]
set -u
N=$1
O=gpurun_out/final_n$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29611 bench.py --gpus $N --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/status.txt
timeout 600 $TR --master-port 29612 tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4.json 2> $O/cfg4.err; echo "cfg4 rc=$?" >> $O/status.txt
NCCL_MAX_NCHANNELS=16 timeout 600 $TR --master-port 29613 tools/run_config.py gpt2-1.5b --steps 5 --sm-limit 132 > $O/cfg4_sm132.json 2> $O/cfg4_sm132.err; echo "cfg4 sm132 rc=$?" >> $O/status.txt
if [ "${2:-}" = "cfg5" ]; then
  timeout 900 $TR --master-port 29614 tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5.json 2> $O/cfg5.err; echo "cfg5 rc=$?" >> $O/status.txt
fi
cat $O/status.txt
python - <<PY
import json
for f in ("bench", "cfg4", "cfg4_sm132", "cfg5"):
    try:
        d = json.loads(open("$O/%s.json" % f).read().strip().splitlines()[-1])
        keep = {k: d.get(k) for k in ("value", "ms_per_step", "tokens_per_s", "frac_of_burst_bf16_peak_per_gpu", "ranks_identical", "comm_exposed_ms", "data_parallel") if k in d}
        print(f, keep)
    except Exception as e:
        print(f, "unreadable", e)
PY
