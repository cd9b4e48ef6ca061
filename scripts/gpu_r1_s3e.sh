#!/bin/bash
# Session-3 call E: exactness certificate + single-term bf16.
mkdir -p gpurun_out
L=gpurun_out/s3e_dev.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
{
echo "=== time f16 12.5M B1024 pair"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 12500000 768 1024 100; echo "rc=$?"
echo "=== time f32 1M B32 k10 tensor"; timeout 300 python scripts/dev_gpu_check.py time tensor f32 1000000 768 32 10 20; echo "rc=$?"
echo "=== time bf16 10M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 1024 100; echo "rc=$?"
echo "=== time f16 10Mx384 B256 k1024"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 384 256 1024; echo "rc=$?"
} > $L 2>&1
grep -E "TIME|rc=[^0]|scans" $L | tail -40
