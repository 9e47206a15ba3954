#!/bin/bash
set -u
N=$1
O=gpurun_out/sml$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for cfg in "140 16" "132 16" "132 32" "124 32"; do
  set -- $cfg
  i=$((i+1))
  NCCL_MAX_NCHANNELS=$2 timeout 600 $TR --master-port $((29520+i)) tools/run_config.py gpt2-1.5b --steps 5 --sm-limit $1 > $O/cfg4_$1_$2.json 2> $O/cfg4_$1_$2.err
  python -c "
import json
d=json.loads(open('$O/cfg4_$1_$2.json').read().strip().splitlines()[-1]); print('sm$1 ch$2', d['ms_per_step'], d['comm_exposed_ms'], d['ranks_identical'])" 2>&1 | tail -1
done
timeout 300 python -m pytest tests/test_fp32_gpu.py -m gpu -x -q 2>&1 | tail -15
