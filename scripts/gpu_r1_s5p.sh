#!/bin/bash
# Session-5 call P: threshold (queries per list) above which a list goes to the tensor-core kernel.
mkdir -p gpurun_out
export IVF_TRAIN=100000
for T in 5 8; do
IVF_TC_MIN=$T timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 32,256,1024 10 5 > gpurun_out/s5p_time_tcmin$T.log 2>&1; echo "tc_min=$T rc=$?"; tail -3 gpurun_out/s5p_time_tcmin$T.log | cut -c1-330
done
