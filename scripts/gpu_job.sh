#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2e
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
timeout 600 python -m pytest tests/test_gpu_group.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log | cut -c1-300
for b in 1 0 1; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2957$b bench.py --gpus $N --steps 10 --warmup 3 --balance $b --no-cpu-baseline > gpurun_out/${T}_bench_cfg3_n${N}_bal$b.json 2> gpurun_out/${T}_bench_cfg3_n${N}_bal$b.err; echo "bench n$N balance=$b rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg3_n${N}_bal$b.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ranks', r['per_rank_scan_ms'], 'w', d['shard_weights'], 'whole', r['whole_search_frac'], d['verify']['ok'], d['verify']['ids_sha256'][:12], d['e2e']['result_equals_device_loop'])"; tail -2 gpurun_out/${T}_bench_cfg3_n${N}_bal$b.err | cut -c1-300
done
ls -la gpurun_out | head
