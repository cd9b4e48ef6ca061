#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2b
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/${T}_pytest.log | cut -c1-300
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err; echo "bench cfg3 rc=$?"; cut -c1-2500 gpurun_out/${T}_bench_cfg3.json; tail -3 gpurun_out/${T}_bench_cfg3.err
for o in 0 1; do
  timeout 300 python bench.py --workload cfg5 --steps 20 --warmup 3 --no-cpu-baseline --opt defer_epilogue=$((o+1)) > gpurun_out/${T}_bench_cfg5_defer$o.json 2> gpurun_out/${T}_bench_cfg5_defer$o.err; echo "bench cfg5 defer=$o rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg5_defer$o.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['verify'])"; tail -2 gpurun_out/${T}_bench_cfg5_defer$o.err
done
for w in cfg2 cfg4; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_${w}.json 2> gpurun_out/${T}_bench_${w}.err; echo "bench $w rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_${w}.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['verify'])"; tail -2 gpurun_out/${T}_bench_${w}.err
done
timeout 300 python bench.py --workload cfg2 --steps 20 --warmup 3 --no-cpu-baseline --opt filter_copy=0 --variant stream > gpurun_out/${T}_bench_cfg2_k2.json 2> gpurun_out/${T}_bench_cfg2_k2.err; echo "bench cfg2 K2 rc=$?"; cut -c1-400 gpurun_out/${T}_bench_cfg2_k2.json
CMD="python bench.py --workload cfg5 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_cfg5.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 7 -c 1 -o gpurun_out/${T}_prof_cfg5 $CMD > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 gpurun_out/${T}_ncu_full.log
timeout 600 python scripts/sweep.py > gpurun_out/${T}_sweep_cfg4_n1.jsonl 2> gpurun_out/${T}_sweep.err; echo "sweep rc=$?"; cat gpurun_out/${T}_sweep_cfg4_n1.jsonl | cut -c1-330
ls -la gpurun_out | head -40
