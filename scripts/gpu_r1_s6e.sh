#!/bin/bash
# Session-5 call 6e: fp16 filter copy for fp32 flat storage: all flat GPU tests (parity, multi, sharded).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py tests/test_reference_integration.py -q > gpurun_out/s6e_pytest_flat.log 2>&1; echo "pytest flat rc=$?"; tail -15 gpurun_out/s6e_pytest_flat.log | cut -c1-300
