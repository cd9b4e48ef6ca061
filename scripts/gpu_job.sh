#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2d
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log | cut -c1-300
for g in 8 4; do
  [ $g -le $N ] || continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2954$g bench.py --gpus $g --steps 10 --warmup 3 > gpurun_out/${T}_bench_cfg3_n$g.json 2> gpurun_out/${T}_bench_cfg3_n$g.err; echo "bench cfg3 n$g rc=$?"; cut -c1-3300 gpurun_out/${T}_bench_cfg3_n$g.json; tail -3 gpurun_out/${T}_bench_cfg3_n$g.err | cut -c1-300
done
for g in 8 4; do
  [ $g -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2955$g scripts/sweep_sharded.py > gpurun_out/${T}_sweep_cfg4_n$g.jsonl 2> gpurun_out/${T}_sweep_n$g.err; echo "sweep n$g rc=$?"; cut -c1-300 gpurun_out/${T}_sweep_cfg4_n$g.jsonl; tail -2 gpurun_out/${T}_sweep_n$g.err | cut -c1-300
done
g=$N
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $g --steps 10 --warmup 3 --exchange nccl --no-cpu-baseline > gpurun_out/${T}_bench_cfg3_n${g}_nccl.json 2> gpurun_out/${T}_bench_cfg3_n${g}_nccl.err; echo "bench cfg3 n$g nccl rc=$?"; cut -c1-400 gpurun_out/${T}_bench_cfg3_n${g}_nccl.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus $g --workload cfg5 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg5_n$g.json 2> gpurun_out/${T}_bench_cfg5_n$g.err; echo "bench cfg5 n$g rc=$?"; cut -c1-1500 gpurun_out/${T}_bench_cfg5_n$g.json
ls -la gpurun_out | head -30
