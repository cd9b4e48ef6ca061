#!/bin/bash
# Session-5 call 6d: fp16 filter copy for fp32 flat storage: parity tests + bench cfg2 (1M x 768 fp32, B=32, k=10).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -q -x > gpurun_out/s6d_pytest_parity.log 2>&1; echo "pytest parity rc=$?"; tail -25 gpurun_out/s6d_pytest_parity.log | cut -c1-300
timeout 900 python bench.py --workload cfg2 > gpurun_out/s6d_bench_cfg2.json 2> gpurun_out/s6d_bench_cfg2.err; echo "bench cfg2 rc=$?"; cut -c1-1500 gpurun_out/s6d_bench_cfg2.json; tail -3 gpurun_out/s6d_bench_cfg2.err
