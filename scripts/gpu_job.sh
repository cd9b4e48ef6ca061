#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2c
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py tests/test_gpu_caller_contract.py "tests/test_gpu_parity.py::test_cfg5_full_size" -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${T}_pytest.log | cut -c1-300
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${T}_bench_cfg3_n$N.json 2> gpurun_out/${T}_bench_cfg3_n$N.err; echo "bench cfg3 n$N rc=$?"; cut -c1-3000 gpurun_out/${T}_bench_cfg3_n$N.json; tail -5 gpurun_out/${T}_bench_cfg3_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 scripts/sweep_sharded.py > gpurun_out/${T}_sweep_cfg4_n$N.jsonl 2> gpurun_out/${T}_sweep_n$N.err; echo "sweep n$N rc=$?"; cut -c1-300 gpurun_out/${T}_sweep_cfg4_n$N.jsonl; tail -3 gpurun_out/${T}_sweep_n$N.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${T}_bench_cfg3_n1.json 2> gpurun_out/${T}_bench_cfg3_n1.err; echo "bench cfg3 n1 rc=$?"; cut -c1-3200 gpurun_out/${T}_bench_cfg3_n1.json; tail -3 gpurun_out/${T}_bench_cfg3_n1.err
ls -la gpurun_out | head -30
