#!/bin/bash
# Session-5 call E: IVF with exactness window + exact rescoring, prefix/skip plan, FFMA2 scan: parity, timing, launch list.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5e_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -15 gpurun_out/s5e_pytest_ivf.log
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 1000000 768 1024 16 1,32,256,1024 10 > gpurun_out/s5e_time_1M.log 2>&1; echo "time rc=$?"; tail -6 gpurun_out/s5e_time_1M.log
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 2"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s5e_launches.csv $CMD > gpurun_out/s5e_ncu.log 2>&1; echo "ncu rc=$?"
grep -E "ivf_|select_topk|coarse" gpurun_out/s5e_launches.csv | tail -14 | awk -F'","' '{print $5, $NF}'
