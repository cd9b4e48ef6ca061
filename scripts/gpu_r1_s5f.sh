#!/bin/bash
# Session-5 call F: warp-specialised IVF scan (64-byte chunks, 5-stage ring): parity, timing, ncu full of the 16-query main launch.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5f_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -5 gpurun_out/s5f_pytest_ivf.log
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 1000000 768 1024 16 1,32,256,1024 10 > gpurun_out/s5f_time_1M.log 2>&1; echo "time rc=$?"; tail -6 gpurun_out/s5f_time_1M.log
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 2"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:ivf_scan_kernel<16>" -s 5 -c 1 -o gpurun_out/s5f_prof_ivf16 $CMD > gpurun_out/s5f_ncu.log 2>&1; echo "ncu rc=$?"
