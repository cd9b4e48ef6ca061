#!/bin/bash
# Session-5 call J: IVF scan v4c (granule loop unrolled by 2): parity, timing, ncu full.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5j_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -5 gpurun_out/s5j_pytest_ivf.log
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 1000000 768 1024 16 1,32,256,1024 10 > gpurun_out/s5j_time_1M.log 2>&1; echo "time rc=$?"; tail -6 gpurun_out/s5j_time_1M.log
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 2"
