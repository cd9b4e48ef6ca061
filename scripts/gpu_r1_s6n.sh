#!/bin/bash
# Session-5 call 6n: coarse quantizer with register prefetch (64 x 64 x 32): IVF tests + timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s6n_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -3 gpurun_out/s6n_pytest_ivf.log | cut -c1-250
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 1,32,256,1024 10 20 > gpurun_out/s6n_time_ref.log 2>&1; echo "rc=$?"; grep -E "^train|^add" gpurun_out/s6n_time_ref.log | cut -c1-160; tail -4 gpurun_out/s6n_time_ref.log | cut -c1-150
