#!/bin/bash
# Session-3 call A: GPU test suite, shard-size bench, ncu source-level capture of the 4 scan phases.
mkdir -p gpurun_out
python scripts/dev_gpu_check.py probe > gpurun_out/probe.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
CMD="python bench.py --rows 12500000 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_shard.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 12 -c 4 -o gpurun_out/prof_tc_scan $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_full.log
cat gpurun_out/plain_shard.log | tail -2
ls -la gpurun_out
