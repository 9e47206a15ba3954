#!/bin/bash
# Pass K: launch list + DRAM bytes + tensor-pipe activity of one eager 2-block bench step (final build); shim GPU test.
set -u
O=gpurun_out/k
mkdir -p $O
timeout 300 python -m pytest tests/test_reference_shim_gpu.py -m gpu -x -q > $O/pytest_shim.log 2>&1; echo "pytest shim rc=$?" >> $O/status.txt
timeout 300 python tools/profile_step.py --shapes $O/shapes.json > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/status.txt
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file $O/launches.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
  python tools/profile_step.py > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; tail -3 $O/pytest_shim.log; tail -2 $O/ncu.log
