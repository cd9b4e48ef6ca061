#!/bin/bash
# Session-4 call G (2 GPUs): K4 fused exchange+merge -- in-process test, 2-rank parity, nccl vs p2p timing.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/s4g_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/s4g_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
P=29600
for X in nccl p2p; do for B in 1 32; do
P=$((P+1))
timeout 600 $TR --master-port $P bench.py --gpus 2 --steps 50 --warmup 5 --workload cfg4 --batch $B --exchange $X --no-cpu-baseline > gpurun_out/s4g_cfg4_b${B}_${X}.json 2> gpurun_out/s4g_cfg4_b${B}_${X}.err; echo "bench cfg4 B=$B $X rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4g_cfg4_b${B}_${X}.json')); print(d['exchange'], d['value'], d['ms_per_step'], d['roofline']['all_scan_launches_ms_per_step'], d['e2e']['value'], d['phases_per_search'])" || tail -5 gpurun_out/s4g_cfg4_b${B}_${X}.err
done; done
timeout 900 $TR --master-port 29650 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/s4g_bench_cfg3_n2.json 2> gpurun_out/s4g_bench_cfg3_n2.err; echo "bench n2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4g_bench_cfg3_n2.json')); print(d['exchange'], d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['e2e']['value'], d['verify'])"
