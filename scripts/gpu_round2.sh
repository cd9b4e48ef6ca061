#!/bin/bash
mkdir -p gpurun_out
{
echo "=== tensor f16 small"; timeout 180 python scripts/dev_gpu_check.py parity tensor f16 5000 128 33 100; echo "rc=$?"
echo "=== tensor bf16"; timeout 180 python scripts/dev_gpu_check.py parity tensor bf16 70001 768 200 100; echo "rc=$?"
echo "=== time tensor f16 10M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 768 1024 100; echo "rc=$?"
echo "=== time tensor f16 12.5M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 12500000 768 1024 100; echo "rc=$?"
echo "=== time tensor f16 10M B128"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 768 128 100; echo "rc=$?"
echo "=== time tensor f16 10M B32"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 768 32 100; echo "rc=$?"
echo "=== time tensor bf16 10M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 1024 100; echo "rc=$?"
echo "=== time tensor f16 10Mx384 B256 k1024"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 384 256 1024; echo "rc=$?"
} > gpurun_out/round2_dev.log 2>&1
tail -40 gpurun_out/round2_dev.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/round2_pytest.log 2>&1
tail -40 gpurun_out/round2_pytest.log
