#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4e_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/s4e_pytest.log
timeout 300 python scripts/sweep.py --rows 10000000 --dim 384 --dtype f16 --k 100 --batches 1,2,3,4 --iters 20 > gpurun_out/s4e_sweep_f16_384.jsonl 2>&1; cat gpurun_out/s4e_sweep_f16_384.jsonl
timeout 300 python scripts/sweep.py --rows 20000000 --dim 128 --dtype f16 --k 10 --batches 1,2,3,8 --iters 20 > gpurun_out/s4e_sweep_f16_128.jsonl 2>&1; cat gpurun_out/s4e_sweep_f16_128.jsonl
timeout 600 python scripts/sweep.py --batches 1,2,3,4 > gpurun_out/s4e_sweep_bf16.jsonl 2> gpurun_out/s4e_sweep.err; cat gpurun_out/s4e_sweep_bf16.jsonl; tail -3 gpurun_out/s4e_sweep.err
