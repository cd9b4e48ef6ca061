#!/bin/bash
# Session-5 call C: first GPU run of the IVF-Flat operator: parity tests, then timing at the reference's index shape.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -x -q > gpurun_out/s5c_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -30 gpurun_out/s5c_pytest_ivf.log
timeout 900 python scripts/dev_ivf_check.py 1000000 768 1024 16 1,32,1024 10 > gpurun_out/s5c_time_1M.log 2>&1; echo "time rc=$?"; cat gpurun_out/s5c_time_1M.log | tail -8
