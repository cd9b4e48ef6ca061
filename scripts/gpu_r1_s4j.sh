#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4j_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/s4j_pytest.log
