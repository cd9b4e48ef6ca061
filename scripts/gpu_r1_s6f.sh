#!/bin/bash
# Session-5 call 6f: the reference's own corpus size as a flat fp32 index (4.5M x 768, B=32, k=10; configs[0] shape)
# with the fp16 filter copy, and ncu --set full of the dominant tc_scan launch at cfg2.
mkdir -p gpurun_out
timeout 900 python bench.py --workload cfg2 --rows 4500000 > gpurun_out/s6f_bench_cfg1_4p5M.json 2> gpurun_out/s6f_bench_cfg1_4p5M.err; echo "bench 4.5M rc=$?"; cut -c1-300 gpurun_out/s6f_bench_cfg1_4p5M.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 7 -c 1 -o gpurun_out/s6f_prof_tc_scan_cfg2_f16copy python bench.py --workload cfg2 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s6f_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/s6f_prof*.ncu-rep
