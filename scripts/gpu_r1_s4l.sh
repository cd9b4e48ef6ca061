#!/bin/bash
# Session-4 call L: live thresholds -- parity + timing across configs.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4l_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/s4l_pytest.log
timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4l_time_cfg5.log 2>&1; tail -3 gpurun_out/s4l_time_cfg5.log
timeout 300 python scripts/dev_gpu_check.py time auto f32 1000000 768 32 10 20 > gpurun_out/s4l_time_cfg2.log 2>&1; tail -3 gpurun_out/s4l_time_cfg2.log
timeout 300 python scripts/dev_gpu_check.py time auto f16 12500000 768 1024 100 10 > gpurun_out/s4l_time_shard.log 2>&1; tail -3 gpurun_out/s4l_time_shard.log
timeout 300 python scripts/dev_gpu_check.py time auto bf16 1250000 768 1 100 20 > gpurun_out/s4l_time_b1.log 2>&1; tail -3 gpurun_out/s4l_time_b1.log
timeout 300 python scripts/dev_gpu_check.py time auto bf16 1250000 768 32 100 20 > gpurun_out/s4l_time_b32.log 2>&1; tail -3 gpurun_out/s4l_time_b32.log
timeout 600 python scripts/sweep.py --batches 1,2,8,64,256,1024 > gpurun_out/s4l_sweep_cfg4.jsonl 2> gpurun_out/s4l_sweep_cfg4.err; cat gpurun_out/s4l_sweep_cfg4.jsonl | cut -c1-230; tail -3 gpurun_out/s4l_sweep_cfg4.err
