#!/bin/bash
# Session-3 call H (2 GPUs): sharded path over NCCL, scaling sanity, cpu-baseline fix check.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_cfg3_n2.json 2> gpurun_out/bench_cfg3_n2.err; echo "bench n2 rc=$?"
tail -c 2200 gpurun_out/bench_cfg3_n2.json; tail -5 gpurun_out/bench_cfg3_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --rows 25000000 > gpurun_out/bench_25M_n2.json 2> gpurun_out/bench_25M_n2.err; echo "bench 25M n2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bench_25M_n2.json')); print(d['value'], d['ms_per_step'], d['roofline']['all_scan_launches_ms_per_step'], d['e2e'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 --workload cfg2 > gpurun_out/bench_cfg2_n2.json 2> gpurun_out/bench_cfg2_n2.err; echo "bench cfg2 n2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bench_cfg2_n2.json')); print(d['value'], d['ms_per_step'], d['roofline']['all_scan_launches_ms_per_step'], d['e2e'])"
python bench.py --rows 2000000 --steps 3 --warmup 3 > gpurun_out/bench_small_cpu.json 2>&1; python -c "import json; d=json.loads(open('gpurun_out/bench_small_cpu.json').read().strip().splitlines()[-1]); print(d['cpu_baseline'])"
