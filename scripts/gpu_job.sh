#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2n
for f in 1 0 1 0; do
timeout 300 python bench.py --rows 12500000 --steps 20 --warmup 3 --no-cpu-baseline --opt fused_rescore=$f > gpurun_out/${T}_bench_shard_f$f.json 2> gpurun_out/${T}_bench_shard_f$f.err; echo "bench shard fused=$f rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_shard_f$f.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], r['all_scan_launches_ms_per_step'], 'nonscan', d['ms_per_step']-r['all_scan_launches_ms_per_step'], d['gpu_launches'])"
done
for f in 1 0; do
CMD="python bench.py --rows 12500000 --steps 2 --warmup 3 --no-cpu-baseline --opt fused_rescore=$f"
$CMD > gpurun_out/${T}_plain_$f.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_shard_f$f.csv $CMD > gpurun_out/${T}_ncu_$f.log 2>&1; echo "ncu f=$f rc=$?"
done
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
