#!/bin/bash
# Session-4 call C: ncu --set full of (1) cfg5 main phase, (2) cfg4 B=256 main phase, (3) K2 at B=1.
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
CMD5="python scripts/dev_gpu_check.py time auto f16 10000000 384 256 1024 1"
$CMD5 > gpurun_out/s4c_plain5.log 2>&1 && $NCU -k regex:tc_scan_kernel -s 4 -c 1 -o gpurun_out/s4c_cfg5_main $CMD5 > gpurun_out/s4c_ncu5.log 2>&1; echo "ncu cfg5 rc=$?"
CMD4="python scripts/dev_gpu_check.py time auto bf16 10000000 768 256 100 1"
$CMD4 > gpurun_out/s4c_plain4.log 2>&1 && $NCU -k regex:tc_scan_kernel -s 3 -c 1 -o gpurun_out/s4c_cfg4_b256_main $CMD4 > gpurun_out/s4c_ncu4.log 2>&1; echo "ncu cfg4 rc=$?"
CMD1="python scripts/dev_gpu_check.py time auto bf16 10000000 768 1 100 1"
$CMD1 > gpurun_out/s4c_plain1.log 2>&1 && $NCU -k regex:stream_scan_kernel -s 3 -c 1 -o gpurun_out/s4c_k2_b1_main $CMD1 > gpurun_out/s4c_ncu1.log 2>&1; echo "ncu k2 b1 rc=$?"
CMD2="python scripts/dev_gpu_check.py time auto bf16 10000000 768 2 100 1"
$CMD2 > gpurun_out/s4c_plain2.log 2>&1 && $NCU -k regex:stream_scan_kernel -s 3 -c 1 -o gpurun_out/s4c_k2_b2_main $CMD2 > gpurun_out/s4c_ncu2.log 2>&1; echo "ncu k2 b2 rc=$?"
ls -la gpurun_out/*.ncu-rep
