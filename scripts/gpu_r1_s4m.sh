#!/bin/bash
# Session-4 call M: live thresholds A/B on cfg5 and the 12.5 M shard, ncu of the live cfg5 filter launch.
mkdir -p gpurun_out
for L in 0 1; do
FIP_LIVE=$L timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4m_time_cfg5_live$L.log 2>&1; tail -2 gpurun_out/s4m_time_cfg5_live$L.log
FIP_LIVE=$L timeout 300 python scripts/dev_gpu_check.py time auto f16 12500000 768 1024 100 10 > gpurun_out/s4m_time_shard_live$L.log 2>&1; tail -2 gpurun_out/s4m_time_shard_live$L.log
done
CMD5="python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 1"
$CMD5 > gpurun_out/s4m_plain5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 1 -c 1 -o gpurun_out/s4m_cfg5_live $CMD5 > gpurun_out/s4m_ncu5.log 2>&1; echo "ncu cfg5 rc=$?"
