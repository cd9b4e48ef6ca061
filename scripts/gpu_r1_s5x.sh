#!/bin/bash
# Session-5 call X: device planner re-check + launch list of one B=1024 search at the reference's index shape.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q -x > gpurun_out/s5x_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -3 gpurun_out/s5x_pytest_ivf.log | cut -c1-300
export IVF_TRAIN=100000
timeout 600 python scripts/dev_ivf_check.py 4500000 768 4096 64 1,32,1024 10 10 > gpurun_out/s5x_time_ref.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/s5x_time_ref.log | cut -c1-200
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s5x_launches_ivf_b1024.csv \
  -k 'regex:ivf_scan|tc_ivf|select_topk|plan_|ivf_rescore|gather_queries|ivf_eps|init_lists|ivf_finalize|row_norm2|coarse_scores' \
  python scripts/dev_ivf_check.py 4500000 768 4096 64 1024 10 1 > gpurun_out/s5x_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/s5x_ncu.log | cut -c1-200
