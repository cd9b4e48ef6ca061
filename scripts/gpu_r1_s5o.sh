#!/bin/bash
# Session-5 call O: bench.py --workload ivf with the tensor-core path (B = 32 / 256 / 1024, balanced and the reference's 10k-row training), ncu of tc_ivf_kernel.
mkdir -p gpurun_out
for B in 32 256 1024; do
timeout 900 python bench.py --workload ivf --batch $B --no-cpu-baseline > gpurun_out/s5o_bench_ivf_b${B}.json 2> gpurun_out/s5o_bench_ivf_b${B}.err; echo "bench ivf B=$B rc=$?"; cut -c1-250 gpurun_out/s5o_bench_ivf_b${B}.json
done
timeout 900 python bench.py --workload ivf --batch 1024 --ivf-train 10000 --no-cpu-baseline > gpurun_out/s5o_bench_ivf_b1024_train10k.json 2> gpurun_out/s5o_bench_ivf_b1024_train10k.err; echo "bench ivf train10k rc=$?"; cut -c1-250 gpurun_out/s5o_bench_ivf_b1024_train10k.json
timeout 900 python bench.py --workload ivf --batch 32 --ivf-train 10000 --no-cpu-baseline > gpurun_out/s5o_bench_ivf_b32_train10k.json 2> gpurun_out/s5o_bench_ivf_b32_train10k.err; echo "bench ivf b32 train10k rc=$?"; cut -c1-250 gpurun_out/s5o_bench_ivf_b32_train10k.json
export IVF_TRAIN=100000
CMD="python scripts/dev_ivf_check.py 1000000 768 1024 16 1024 10 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_ivf_kernel -s 3 -c 1 -o gpurun_out/s5o_prof_tc_ivf $CMD > gpurun_out/s5o_ncu.log 2>&1; echo "ncu rc=$?"
