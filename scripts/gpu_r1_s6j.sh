#!/bin/bash
# Session-5 call 6j: final verification at HEAD: every GPU test, smoke(), default bench (both arms), IVF bench lines.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/s6j_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/s6j_pytest_gpu.log | cut -c1-300
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s6j_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s6j_smoke.log | cut -c1-300
timeout 900 python bench.py > gpurun_out/s6j_bench_default.json 2> gpurun_out/s6j_bench_default.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/s6j_bench_default.json
timeout 900 python bench.py --workload ivf > gpurun_out/s6j_bench_ivf_b32.json 2> gpurun_out/s6j_bench_ivf_b32.err; echo "bench ivf rc=$?"; cut -c1-250 gpurun_out/s6j_bench_ivf_b32.json
for B in 256 1024; do
timeout 900 python bench.py --workload ivf --batch $B --no-cpu-baseline > gpurun_out/s6j_bench_ivf_b${B}.json 2> gpurun_out/s6j_bench_ivf_b${B}.err; echo "bench ivf B=$B rc=$?"; cut -c1-250 gpurun_out/s6j_bench_ivf_b${B}.json
done
for B in 32 1024; do
timeout 900 python bench.py --workload ivf --batch $B --ivf-train 10000 --no-cpu-baseline > gpurun_out/s6j_bench_ivf_b${B}_train10k.json 2> gpurun_out/s6j_bench_ivf_b${B}_train10k.err; echo "bench ivf train10k B=$B rc=$?"; cut -c1-250 gpurun_out/s6j_bench_ivf_b${B}_train10k.json
done
timeout 900 python bench.py --workload cfg2 > gpurun_out/s6j_bench_cfg2.json 2> gpurun_out/s6j_bench_cfg2.err; echo "bench cfg2 rc=$?"; cut -c1-250 gpurun_out/s6j_bench_cfg2.json
