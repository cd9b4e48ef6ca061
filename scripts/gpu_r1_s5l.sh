#!/bin/bash
# Session-5 call L: full GPU suite + smoke + default bench at HEAD (flat + IVF in one library).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s5l_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s5l_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s5l_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/s5l_smoke.log
timeout 900 python bench.py > gpurun_out/s5l_bench_cfg3_n1.json 2> gpurun_out/s5l_bench_cfg3_n1.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/s5l_bench_cfg3_n1.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s5l_bench_reference.json 2>&1; echo "ref rc=$?"; cut -c1-300 gpurun_out/s5l_bench_reference.json
