#!/bin/bash
# Session-4 call H (8 GPUs): 8-rank parity of both exchange paths, headline bench at N=8, small batch.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/s4h_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s4h_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29711 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/s4h_bench_cfg3_n8.json 2> gpurun_out/s4h_bench_cfg3_n8.err; echo "bench n8 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4h_bench_cfg3_n8.json')); print(d['exchange'], d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['e2e']['value'], d['clocks'], d['verify'])" || tail -20 gpurun_out/s4h_bench_cfg3_n8.err
timeout 900 $TR --master-port 29712 bench.py --gpus 8 --steps 10 --warmup 3 --exchange nccl > gpurun_out/s4h_bench_cfg3_n8_nccl.json 2> gpurun_out/s4h_bench_cfg3_n8_nccl.err; echo "bench n8 nccl rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4h_bench_cfg3_n8_nccl.json')); print(d['exchange'], d['value'], d['ms_per_step'], d['e2e']['value'])"
P=29720
for X in p2p nccl; do for B in 1 32; do
P=$((P+1))
timeout 600 $TR --master-port $P bench.py --gpus 8 --steps 50 --warmup 5 --workload cfg4 --batch $B --exchange $X --no-cpu-baseline > gpurun_out/s4h_cfg4_b${B}_${X}_n8.json 2> gpurun_out/s4h_cfg4_b${B}_${X}_n8.err; echo "bench cfg4 B=$B $X rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4h_cfg4_b${B}_${X}_n8.json')); print(d['exchange'], d['value'], d['ms_per_step'], d['roofline']['all_scan_launches_ms_per_step'], d['e2e']['value'], d['phases_per_search'])"
done; done
