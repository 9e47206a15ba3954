#!/bin/bash
# 2-GPU A/B for Llama-3-8B: default exchange vs few NCCL channels + a small SM reservation.
set -u
O=gpurun_out/x
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29811 tools/run_config.py llama3-8b --steps 3 --warmup 2 > $O/cfg5_base.json 2> $O/cfg5_base.err; echo "base rc=$?" >> $O/status.txt
NCCL_MAX_NCHANNELS=4 NCCL_MIN_NCHANNELS=4 timeout 600 $TR --master-port 29812 tools/run_config.py llama3-8b --steps 3 --warmup 2 --sm-limit 144 > $O/cfg5_ch4_sm144.json 2> $O/cfg5_ch4_sm144.err; echo "ch4 rc=$?" >> $O/status.txt
cat $O/status.txt
python - <<PY
import json
for f in ("cfg5_base", "cfg5_ch4_sm144"):
    try:
        d = json.loads(open("$O/%s.json" % f).read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"],1), round(d["tokens_per_s"]), d["ranks_identical"], round(d["comm_exposed_ms"],1))
    except Exception as e:
        print(f, "unreadable", e)
PY
