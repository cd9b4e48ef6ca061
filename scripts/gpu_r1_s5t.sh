#!/bin/bash
# Session-5 call T: full GPU suite, smoke, default bench, IVF bench lines at HEAD.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s5t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s5t_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s5t_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/s5t_smoke.log
timeout 900 python bench.py > gpurun_out/s5t_bench_cfg3_n1.json 2> gpurun_out/s5t_bench_cfg3_n1.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/s5t_bench_cfg3_n1.json
timeout 900 python bench.py --workload ivf > gpurun_out/s5t_bench_ivf_b32.json 2> gpurun_out/s5t_bench_ivf_b32.err; echo "bench ivf rc=$?"; cut -c1-250 gpurun_out/s5t_bench_ivf_b32.json
for B in 256 1024; do
timeout 900 python bench.py --workload ivf --batch $B --no-cpu-baseline > gpurun_out/s5t_bench_ivf_b${B}.json 2> gpurun_out/s5t_bench_ivf_b${B}.err; echo "bench ivf B=$B rc=$?"; cut -c1-250 gpurun_out/s5t_bench_ivf_b${B}.json
done
for B in 32 1024; do
timeout 900 python bench.py --workload ivf --batch $B --ivf-train 10000 --no-cpu-baseline > gpurun_out/s5t_bench_ivf_b${B}_train10k.json 2> gpurun_out/s5t_bench_ivf_b${B}_train10k.err; echo "bench ivf train10k B=$B rc=$?"; cut -c1-250 gpurun_out/s5t_bench_ivf_b${B}_train10k.json
done
