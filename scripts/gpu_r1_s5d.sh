#!/bin/bash
# Session-5 call D: IVF parity tests (fixed probe-stability check), timing with balanced lists, ncu of the 16-query scan.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5d_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -30 gpurun_out/s5d_pytest_ivf.log
export IVF_TRAIN=100000
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 3"
timeout 600 python scripts/dev_ivf_check.py 1000000 768 1024 16 1,32,256,1024 10 > gpurun_out/s5d_time_1M.log 2>&1; echo "time rc=$?"; tail -6 gpurun_out/s5d_time_1M.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ivf_scan_kernel -s 8 -c 2 -o gpurun_out/s5d_prof_ivf $CMD > gpurun_out/s5d_ncu.log 2>&1; echo "ncu rc=$?"
