#!/bin/bash
# Session-5 call 6o: smoke() at HEAD and the IVF bench line (B = 32) with the final kernels.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s6o_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s6o_smoke.log | cut -c1-200
timeout 600 python bench.py --workload ivf > gpurun_out/s6o_bench_ivf_b32.json 2> gpurun_out/s6o_bench_ivf_b32.err; echo "bench ivf rc=$?"; cut -c1-250 gpurun_out/s6o_bench_ivf_b32.json
