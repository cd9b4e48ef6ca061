#!/bin/bash
# Session-4 call A: re-validate the restored checkpoint (GPU tests), batch sweep, cfg5/cfg2 timing.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4a_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/s4a_pytest.log
timeout 600 python scripts/sweep.py > gpurun_out/s4a_sweep_cfg4.jsonl 2> gpurun_out/s4a_sweep_cfg4.err; echo "sweep rc=$?"
cat gpurun_out/s4a_sweep_cfg4.jsonl; tail -3 gpurun_out/s4a_sweep_cfg4.err
timeout 300 python scripts/sweep.py --rows 10000000 --dim 384 --dtype f16 --k 1024 --batches 256 > gpurun_out/s4a_sweep_cfg5.jsonl 2>&1; cat gpurun_out/s4a_sweep_cfg5.jsonl
timeout 300 python scripts/sweep.py --rows 1000000 --dim 768 --dtype f32 --k 10 --batches 1,4,32 --iters 30 > gpurun_out/s4a_sweep_cfg2.jsonl 2>&1; cat gpurun_out/s4a_sweep_cfg2.jsonl
timeout 300 python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 > gpurun_out/s4a_time_cfg5.log 2>&1; tail -3 gpurun_out/s4a_time_cfg5.log
timeout 300 python scripts/dev_gpu_check.py time auto f32 1000000 768 32 10 20 > gpurun_out/s4a_time_cfg2.log 2>&1; tail -3 gpurun_out/s4a_time_cfg2.log
