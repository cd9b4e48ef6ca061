#!/bin/bash
# Session-5 call K: IVF at the reference's index shape (4.5 M x 768, nlist 4096, nprobe 64; scripts/create_test_docs.py).
mkdir -p gpurun_out
free -g | head -2
IVF_TRAIN=10000 timeout 900 python scripts/dev_ivf_check.py 4500000 768 4096 64 1,32,256,1024 10 5 > gpurun_out/s5k_time_ref_train10k.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/s5k_time_ref_train10k.log
IVF_TRAIN=100000 timeout 900 python scripts/dev_ivf_check.py 4500000 768 4096 64 1,32,256,1024 10 5 > gpurun_out/s5k_time_ref_train100k.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/s5k_time_ref_train100k.log
