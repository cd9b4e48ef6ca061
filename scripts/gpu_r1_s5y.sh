#!/bin/bash
# Session-5 call Y: tensor_min_queries sweep at the reference's index shape (device planner).
mkdir -p gpurun_out
export IVF_TRAIN=100000
for t in 1 2 4; do
  IVF_TC_MIN=$t timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 32,256,1024 10 10 > gpurun_out/s5y_time_tcmin$t.log 2>&1; echo "tc_min=$t rc=$?"; tail -3 gpurun_out/s5y_time_tcmin$t.log | cut -c1-120
done
