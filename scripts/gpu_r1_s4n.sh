#!/bin/bash
# Session-4 call N: A/B live thresholds after the code-size fix.
mkdir -p gpurun_out
for L in 0 1; do
FIP_LIVE=$L timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4n_time_cfg5_live$L.log 2>&1; tail -2 gpurun_out/s4n_time_cfg5_live$L.log
FIP_LIVE=$L timeout 300 python scripts/dev_gpu_check.py time auto f16 12500000 768 1024 100 10 > gpurun_out/s4n_time_shard_live$L.log 2>&1; tail -2 gpurun_out/s4n_time_shard_live$L.log
FIP_LIVE=$L timeout 300 python scripts/dev_gpu_check.py time auto bf16 10000000 768 256 100 10 > gpurun_out/s4n_time_b256_live$L.log 2>&1; tail -2 gpurun_out/s4n_time_b256_live$L.log
done
timeout 900 python -m pytest tests -m gpu -x -q -k "not live_thresholds" > gpurun_out/s4n_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s4n_pytest.log
