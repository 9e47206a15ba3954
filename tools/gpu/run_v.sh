#!/bin/bash
set -u
O=gpurun_out/v
mkdir -p $O
timeout 300 python tools/profile_step.py --config gpt2-774m-lora > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/status.txt
timeout 900 ncu --profile-from-start off --clock-control none --csv --log-file $O/launches.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
  python tools/profile_step.py --config gpt2-774m-lora > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/status.txt
cat $O/status.txt; tail -2 $O/plain.log
