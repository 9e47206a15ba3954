#!/bin/bash
set -u
N=$1
O=gpurun_out/prio$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B200_NCCL_HIGH_PRIORITY=1 timeout 600 $TR --master-port 29512 tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_hi.json 2> $O/cfg4_hi.err
B200_NCCL_HIGH_PRIORITY=0 timeout 600 $TR --master-port 29513 tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_lo.json 2> $O/cfg4_lo.err
B200_NCCL_HIGH_PRIORITY=1 NCCL_MAX_NCHANNELS=8 timeout 600 $TR --master-port 29514 tools/run_config.py gpt2-1.5b --steps 5 > $O/cfg4_hi8.json 2> $O/cfg4_hi8.err
B200_NCCL_HIGH_PRIORITY=1 timeout 600 $TR --master-port 29515 tools/run_config.py gpt2-1.5b --steps 5 --bucket-mb 256 > $O/cfg4_hi_b256.json 2> $O/cfg4_hi_b256.err
for f in hi lo hi8 hi_b256; do python -c "
import json,sys
d=json.loads(open('$O/cfg4_$f.json').read().strip().splitlines()[-1]); print('$f', d['ms_per_step'], d['comm_exposed_ms'], d['ranks_identical'])" 2>&1 | tail -1; done
tail -2 $O/*.err | tail -12
