#!/bin/bash
# Session-4 call I: launch lists (ncu gpu__time_duration) of small-batch searches on a 1.25 M-row shard.
mkdir -p gpurun_out
for B in 1 32; do
CMD="python scripts/dev_gpu_check.py time auto bf16 1250000 768 $B 100 3"
$CMD > gpurun_out/s4i_plain_b$B.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/s4i_launches_b$B.csv $CMD > gpurun_out/s4i_ncu_b$B.log 2>&1; echo "ncu B=$B rc=$?"
tail -2 gpurun_out/s4i_plain_b$B.log
done
