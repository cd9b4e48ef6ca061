#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2p
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain_cfg3.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_cfg3.csv $CMD > gpurun_out/${T}_ncu_launches_cfg3.log 2>&1; echo "ncu launches cfg3 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 7 -c 1 -o gpurun_out/${T}_prof_cfg3 $CMD > gpurun_out/${T}_ncu_full_cfg3.log 2>&1; echo "ncu full cfg3 rc=$?"; tail -2 gpurun_out/${T}_ncu_full_cfg3.log
ls -la gpurun_out | head
