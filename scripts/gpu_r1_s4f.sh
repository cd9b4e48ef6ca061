#!/bin/bash
# Session-4 call F (2 GPUs): packed single all-gather; bench cfg3 N=2 + small-batch sharded timing.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "packed or sharded or merge or auto_variant" > gpurun_out/s4f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s4f_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/s4f_bench_cfg3_n2.json 2> gpurun_out/s4f_bench_cfg3_n2.err; echo "bench n2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4f_bench_cfg3_n2.json')); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['e2e'], d['clocks'], d['verify'])"; tail -3 gpurun_out/s4f_bench_cfg3_n2.err
for B in 1 32 1024; do
timeout 600 $TR --master-port 2952$((B%10)) bench.py --gpus 2 --steps 20 --warmup 3 --workload cfg4 --batch $B --no-cpu-baseline > gpurun_out/s4f_bench_cfg4_b${B}_n2.json 2> gpurun_out/s4f_bench_cfg4_b${B}_n2.err; echo "bench cfg4 B=$B n2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/s4f_bench_cfg4_b${B}_n2.json')); print(d['value'], d['ms_per_step'], d['roofline']['all_scan_launches_ms_per_step'], d['e2e']['value'], d['verify'])"
done
