#!/bin/bash
# Session-4 call O: 8 vs 16 epilogue warps (A/B builds), threshold prefetch.
mkdir -p gpurun_out
for LIB in libflatip_b200.so libflatip_b200_epi16.so; do
export FLATIP_B200_LIBRARY=$PWD/rag-inference-pipeline_b200/$LIB
echo "== $LIB"
timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4o_cfg5_$LIB.log 2>&1; tail -2 gpurun_out/s4o_cfg5_$LIB.log
timeout 300 python scripts/dev_gpu_check.py time auto f16 12500000 768 1024 100 10 > gpurun_out/s4o_shard_$LIB.log 2>&1; tail -2 gpurun_out/s4o_shard_$LIB.log
timeout 300 python scripts/dev_gpu_check.py time auto bf16 10000000 768 256 100 10 > gpurun_out/s4o_b256_$LIB.log 2>&1; tail -2 gpurun_out/s4o_b256_$LIB.log
timeout 300 python scripts/dev_gpu_check.py time auto bf16 10000000 768 32 100 10 > gpurun_out/s4o_b32_$LIB.log 2>&1; tail -2 gpurun_out/s4o_b32_$LIB.log
timeout 300 python scripts/dev_gpu_check.py time auto f32 1000000 768 32 10 20 > gpurun_out/s4o_cfg2_$LIB.log 2>&1; tail -2 gpurun_out/s4o_cfg2_$LIB.log
done
export FLATIP_B200_LIBRARY=$PWD/rag-inference-pipeline_b200/libflatip_b200_epi16.so
timeout 900 python -m pytest tests -m gpu -x -q -k "not live_thresholds" > gpurun_out/s4o_pytest16.log 2>&1; echo "pytest16 rc=$?"; tail -4 gpurun_out/s4o_pytest16.log
