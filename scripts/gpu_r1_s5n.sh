#!/bin/bash
# Session-5 call N: tensor-core path for heavily probed IVF lists: parity tests, timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -x -q > gpurun_out/s5n_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -25 gpurun_out/s5n_pytest_ivf.log
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 1000000 768 1024 16 32,256,1024 10 > gpurun_out/s5n_time_1M.log 2>&1; echo "time rc=$?"; tail -4 gpurun_out/s5n_time_1M.log
