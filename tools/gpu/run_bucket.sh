#!/bin/bash
# N-GPU A/B of the gradient-exchange bucket size on the bench workload: usage  bash tools/gpu/run_bucket.sh N
set -u
N=$1
O=gpurun_out/bucket$N
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 python -m pytest tests/test_ddp_gpu.py -m gpu -q > $O/pytest_ddp.log 2>&1; echo "pytest_ddp rc=$?" >> $O/status.txt
for mb in 64 16 8; do
  B200_DDP_BUCKET_MB=$mb timeout 600 $TR --master-port 295$mb bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > $O/bench_$mb.json 2> $O/bench_$mb.err; echo "bench $mb rc=$?" >> $O/status.txt
done
cat $O/status.txt; tail -3 $O/pytest_ddp.log
for mb in 64 16 8; do python - <<PY
import json
try:
    d=json.loads(open('$O/bench_$mb.json').read().strip().splitlines()[-1])
    print('bucket $mb MB', round(d['value']), round(d['ms_per_step'],2), d['data_parallel']['comm_exposed_ms'], d['data_parallel']['ranks_identical'])
except Exception as e: print('$mb', e)
PY
done
