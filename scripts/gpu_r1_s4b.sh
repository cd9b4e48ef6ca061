#!/bin/bash
# Session-4 call B: two-pass epilogue (no smem strip, one atomic per tile) -- parity + timing.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4b_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/s4b_pytest.log
timeout 600 python scripts/sweep.py --batches 8,128,256,512,1024 > gpurun_out/s4b_sweep_cfg4.jsonl 2> gpurun_out/s4b_sweep_cfg4.err; echo "sweep rc=$?"
cat gpurun_out/s4b_sweep_cfg4.jsonl; tail -3 gpurun_out/s4b_sweep_cfg4.err
timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4b_time_cfg5.log 2>&1; tail -3 gpurun_out/s4b_time_cfg5.log
timeout 300 python scripts/dev_gpu_check.py time auto f32 1000000 768 32 10 20 > gpurun_out/s4b_time_cfg2.log 2>&1; tail -3 gpurun_out/s4b_time_cfg2.log
timeout 300 python scripts/dev_gpu_check.py time auto f16 12500000 768 1024 100 10 > gpurun_out/s4b_time_shard.log 2>&1; tail -3 gpurun_out/s4b_time_shard.log
