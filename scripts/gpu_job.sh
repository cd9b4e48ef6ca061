#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2g
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/${T}_smoke.log
for w in cfg2 cfg5; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_${w}.json 2> gpurun_out/${T}_bench_${w}.err; echo "bench $w rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_${w}.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['gpu_launches'], d['verify']['ok'])"; tail -2 gpurun_out/${T}_bench_${w}.err
done
CMD="python bench.py --workload cfg2 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain_cfg2.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_cfg2.csv $CMD > gpurun_out/${T}_ncu_launches_cfg2.log 2>&1; echo "ncu cfg2 rc=$?"
CMD="python bench.py --rows 12500000 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain_shard.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_shard.csv $CMD > gpurun_out/${T}_ncu_launches_shard.log 2>&1; echo "ncu shard rc=$?"
ls gpurun_out | head -40
