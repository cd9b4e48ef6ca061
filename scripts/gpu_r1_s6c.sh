#!/bin/bash
# Session-5 call 6c: full verification at HEAD: every GPU test, smoke(), the default bench (both arms).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s6c_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/s6c_pytest_gpu.log | cut -c1-300
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s6c_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s6c_smoke.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/s6c_bench_default.json 2> gpurun_out/s6c_bench_default.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/s6c_bench_default.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s6c_bench_reference.json 2> gpurun_out/s6c_bench_reference.err; echo "bench ref rc=$?"; cut -c1-300 gpurun_out/s6c_bench_reference.json
