#!/bin/bash
# Session-3 call B: new K1 (CTA pairs, tf32, branch-free epilogue): parity first, then timings.
mkdir -p gpurun_out
L=gpurun_out/s3b_dev.log
{
echo "=== parity pair f16"; timeout 120 python scripts/dev_gpu_check.py parity tensor f16 30000 768 1000 100; echo "rc=$?"
echo "=== parity single f16"; FIP_CTA_PAIR=1 timeout 120 python scripts/dev_gpu_check.py parity tensor f16 50000 768 64 100; echo "rc=$?"
echo "=== parity tf32"; timeout 120 python scripts/dev_gpu_check.py parity tensor f32 20000 768 32 10; echo "rc=$?"
} > $L 2>&1
grep -E "rc=|score_ok|===|Error|error" $L | head -30
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
{
echo "=== time f16 12.5M B1024 pair"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 12500000 768 1024 100; echo "rc=$?"
echo "=== time f16 12.5M B1024 single"; FIP_CTA_PAIR=1 timeout 300 python scripts/dev_gpu_check.py time tensor f16 12500000 768 1024 100; echo "rc=$?"
echo "=== time f32 1M B32 k10 tensor"; timeout 300 python scripts/dev_gpu_check.py time tensor f32 1000000 768 32 10 20; echo "rc=$?"
echo "=== time f32 1M B32 k10 stream"; timeout 300 python scripts/dev_gpu_check.py time stream f32 1000000 768 32 10 20; echo "rc=$?"
echo "=== time bf16 10M B1024"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 1024 100; echo "rc=$?"
echo "=== time f16 10Mx384 B256 k1024"; timeout 300 python scripts/dev_gpu_check.py time tensor f16 10000000 384 256 1024; echo "rc=$?"
echo "=== time bf16 10M B64"; timeout 300 python scripts/dev_gpu_check.py time tensor bf16 10000000 768 64 100; echo "rc=$?"
} >> $L 2>&1
grep -E "TIME|rc=|===" $L | tail -30
