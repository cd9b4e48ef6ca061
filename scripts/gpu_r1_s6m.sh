#!/bin/bash
# Session-5 call 6m: launch list (per-kernel device times) of IVF searches at B = 32 and B = 1024 with the fp16 filter copy.
mkdir -p gpurun_out
export IVF_TRAIN=100000
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s6m_launches_ivf.csv \
  -k 'regex:ivf_scan|tc_ivf|select_topk|plan_|ivf_rescore|gather_queries|ivf_eps|init_lists|ivf_finalize|row_norm2|coarse_scores|rows_to_f16' \
  python scripts/dev_ivf_check.py 4500000 768 4096 64 32,1024 10 1 > gpurun_out/s6m_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/s6m_ncu.log | cut -c1-160
