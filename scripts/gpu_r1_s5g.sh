#!/bin/bash
# Session-5 call G: ncu --set full of the IVF scan launches (main 16-query launch among them).
mkdir -p gpurun_out
export IVF_TRAIN=100000
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ivf_scan_kernel -s 12 -c 4 -o gpurun_out/s5g_prof_ivf $CMD > gpurun_out/s5g_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/s5g_ncu.log
