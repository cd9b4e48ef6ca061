#!/bin/bash
# Session-5 call 6b: bench.py --workload ivf with the fp16 filter copy (B = 32 / 256 / 1024; the reference's
# 10k-row training at B = 32 / 1024), then ncu --set full of tc_ivf_kernel<0> at the reference's shape, B = 1024.
mkdir -p gpurun_out
timeout 900 python bench.py --workload ivf > gpurun_out/s6b_bench_ivf_b32.json 2> gpurun_out/s6b_bench_ivf_b32.err; echo "bench ivf rc=$?"; cut -c1-250 gpurun_out/s6b_bench_ivf_b32.json
for B in 256 1024; do
timeout 900 python bench.py --workload ivf --batch $B --no-cpu-baseline > gpurun_out/s6b_bench_ivf_b${B}.json 2> gpurun_out/s6b_bench_ivf_b${B}.err; echo "bench ivf B=$B rc=$?"; cut -c1-250 gpurun_out/s6b_bench_ivf_b${B}.json
done
for B in 32 1024; do
timeout 900 python bench.py --workload ivf --batch $B --ivf-train 10000 --no-cpu-baseline > gpurun_out/s6b_bench_ivf_b${B}_train10k.json 2> gpurun_out/s6b_bench_ivf_b${B}_train10k.err; echo "bench ivf train10k B=$B rc=$?"; cut -c1-250 gpurun_out/s6b_bench_ivf_b${B}_train10k.json
done
export IVF_TRAIN=100000
CMD="python scripts/dev_ivf_check.py 4500000 768 4096 64 1024 10 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_ivf_kernel -s 3 -c 1 -o gpurun_out/s6b_prof_tc_ivf_f16 $CMD > gpurun_out/s6b_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
