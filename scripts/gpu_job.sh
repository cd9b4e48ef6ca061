#!/bin/bash
# Round-2 GPU job: multi-GPU checks at HEAD (group / torchrun tests, default bench under torchrun).
mkdir -p gpurun_out
T=r2t
N=$(nvidia-smi -L | wc -l); echo "gpus=$N"
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py tests/test_gpu_caller_contract.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29841 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg3_n$N.json 2> gpurun_out/${T}_bench_cfg3_n$N.err; echo "bench n$N rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg3_n$N.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ranks', r['per_rank_scan_ms'], 'w', d['shard_weights'], 'whole', r['whole_search_frac'], d['verify']['ok'], d['verify']['ids_sha256'][:12], d['e2e']['result_equals_device_loop'])"; tail -2 gpurun_out/${T}_bench_cfg3_n$N.err | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29842 bench.py --impl reference --gpus $N --steps 2 --warmup 1 --rows 2000000 > gpurun_out/${T}_bench_reference_n$N.json 2> gpurun_out/${T}_bench_reference_n$N.err; echo "reference n$N rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_reference_n$N.json')); print(d['value'], d['cpu_baseline']['cores'], d['cpu_baseline']['omp_num_threads_env'])"
