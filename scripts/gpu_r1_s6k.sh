#!/bin/bash
# Session-5 call 6k (2 GPUs): multi-GPU tests and the N=2 bench line at HEAD.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q > gpurun_out/s6k_pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -3 gpurun_out/s6k_pytest_multi.log | cut -c1-250
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/s6k_bench_cfg3_n2.json 2> gpurun_out/s6k_bench_cfg3_n2.err; echo "bench n2 rc=$?"; tail -1 gpurun_out/s6k_bench_cfg3_n2.json | cut -c1-400
