#!/bin/bash
# Session-5 call U: IVF tests (large-k case, crowded tf32 window fallback).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5u_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -25 gpurun_out/s5u_pytest_ivf.log | cut -c1-300
