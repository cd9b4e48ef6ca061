#!/bin/bash
# Session-5 call 6a: fp16 filter copy on both planner paths, tensor_min_queries = 1: tests + timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s6a_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -30 gpurun_out/s6a_pytest_ivf.log | cut -c1-250
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 1,8,32,256,1024 10 10 > gpurun_out/s6a_time_ref.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/s6a_time_ref.log | cut -c1-330
IVF_TRAIN=10000 timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 32,1024 10 10 > gpurun_out/s6a_time_ref10k.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/s6a_time_ref10k.log | cut -c1-330
