#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2m
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-300
g=$N
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $g --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg3_n$g.json 2> gpurun_out/${T}_bench_cfg3_n$g.err; echo "bench cfg3 n$g rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg3_n$g.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ranks', r['per_rank_scan_ms'], 'frac', r['frac'], 'whole', r['whole_search_frac'], 'share', r['kernel_share_of_step'], d['gpu_launches'], d['verify']['ok'], d['verify']['ids_sha256'][:12], d['e2e']['result_equals_device_loop'])"; tail -2 gpurun_out/${T}_bench_cfg3_n$g.err | cut -c1-300
for g in 8 4 2; do
  [ $g -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port 2975$g scripts/sweep_sharded.py > gpurun_out/${T}_sweep_cfg4_n$g.jsonl 2> gpurun_out/${T}_sweep_n$g.err; echo "sweep n$g rc=$?"; cut -c1-260 gpurun_out/${T}_sweep_cfg4_n$g.jsonl; tail -2 gpurun_out/${T}_sweep_n$g.err | cut -c1-200
done
timeout 600 python scripts/sweep.py > gpurun_out/${T}_sweep_cfg4_n1.jsonl 2> gpurun_out/${T}_sweep_n1.err; echo "sweep n1 rc=$?"; cut -c1-260 gpurun_out/${T}_sweep_cfg4_n1.jsonl
