#!/bin/bash
# Session-5 call B: ncu --set full of the single filter launch (live thresholds) at cfg3 N=1.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/s5b_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 1 -c 1 -o gpurun_out/s5b_prof_tc_main $CMD > gpurun_out/s5b_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/s5b_ncu_full.log
