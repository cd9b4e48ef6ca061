#!/bin/bash
# Round-2 GPU job: ONE compute-sanitizer tool per call (B200_PROFILING.md) on the smoke-sized pass over
# every kernel family.  usage: bash scripts/gpu_job.sh <memcheck|racecheck|synccheck>
mkdir -p gpurun_out
TOOL=${1:-memcheck}
timeout 300 python scripts/sanitize_smoke.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 50 python scripts/sanitize_smoke.py > gpurun_out/sanitize_$TOOL.log 2>&1; echo "$TOOL rc=$?"
grep -c "^ok" gpurun_out/sanitize_$TOOL.log; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SANITIZE_SMOKE_OK|Error|hazard" gpurun_out/sanitize_$TOOL.log | head -20; tail -5 gpurun_out/sanitize_$TOOL.log | cut -c1-300
