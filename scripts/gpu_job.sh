#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2q
timeout 300 python bench.py --workload cfg2 --rows 4500000 --steps 20 --warmup 3 > gpurun_out/${T}_bench_flat_4p5M.json 2> gpurun_out/${T}_bench_flat_4p5M.err; echo "flat 4.5M rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_flat_4p5M.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['cpu_baseline']['value'], d['verify']['ok'])"
timeout 600 python bench.py --workload ivf --batch 256 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_ivf_b256.json 2> gpurun_out/${T}_bench_ivf_b256.err; echo "ivf 256 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_ivf_b256.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'])"
timeout 300 python bench.py --workload cfg2 --steps 20 --warmup 3 --batch 1 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2_b1.json 2> gpurun_out/${T}_bench_cfg2_b1.err; echo "cfg2 B=1 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg2_b1.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['variant'])"
