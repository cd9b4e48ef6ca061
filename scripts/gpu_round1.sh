#!/bin/bash
# First GPU bring-up: probe, per-variant parity (each in its own process under a timeout), timing.
mkdir -p gpurun_out
{
python scripts/dev_gpu_check.py probe
echo "=== stream f32"; timeout 180 python scripts/dev_gpu_check.py parity stream f32 20000 768 32 10
echo "=== stream f16"; timeout 180 python scripts/dev_gpu_check.py parity stream f16 30011 256 3 100
echo "=== tensor f16 small"; timeout 180 python scripts/dev_gpu_check.py parity tensor f16 5000 128 33 100; echo "rc=$?"
echo "=== tensor bf16"; timeout 180 python scripts/dev_gpu_check.py parity tensor bf16 70001 768 200 100; echo "rc=$?"
echo "=== time stream f32 cfg2"; timeout 300 python scripts/dev_gpu_check.py time stream f32 1000000 768 32 10
echo "=== time stream f16 B1"; timeout 300 python scripts/dev_gpu_check.py time stream f16 4000000 768 1 100
echo "=== time tensor bf16 10M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 1024 100; echo "rc=$?"
echo "=== time tensor bf16 10M B64"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 64 100; echo "rc=$?"
} > gpurun_out/round1_dev.log 2>&1
tail -60 gpurun_out/round1_dev.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/round1_pytest.log 2>&1
tail -30 gpurun_out/round1_pytest.log
