#!/bin/bash
# Session-4 call D: K2 with per-warp bulk-copy rings -- parity + small-batch timing.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4d_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/s4d_pytest.log
timeout 600 python scripts/sweep.py --batches 1,2,3,4,8 > gpurun_out/s4d_sweep_bf16.jsonl 2> gpurun_out/s4d_sweep.err; cat gpurun_out/s4d_sweep_bf16.jsonl; tail -3 gpurun_out/s4d_sweep.err
timeout 600 python scripts/sweep.py --batches 4,8,16 --variant stream > gpurun_out/s4d_sweep_bf16_stream.jsonl 2> gpurun_out/s4d_sweep.err; cat gpurun_out/s4d_sweep_bf16_stream.jsonl; tail -3 gpurun_out/s4d_sweep.err
timeout 300 python scripts/sweep.py --rows 4000000 --dim 768 --dtype f32 --k 10 --batches 1,2,4 --iters 20 > gpurun_out/s4d_sweep_f32.jsonl 2>&1; cat gpurun_out/s4d_sweep_f32.jsonl
timeout 300 python scripts/sweep.py --rows 10000000 --dim 384 --dtype f16 --k 100 --batches 1,2,4 --iters 20 > gpurun_out/s4d_sweep_f16_384.jsonl 2>&1; cat gpurun_out/s4d_sweep_f16_384.jsonl
timeout 300 python scripts/sweep.py --rows 5000000 --dim 1536 --dtype f16 --k 100 --batches 1,4 --iters 20 > gpurun_out/s4d_sweep_f16_1536.jsonl 2>&1; cat gpurun_out/s4d_sweep_f16_1536.jsonl
