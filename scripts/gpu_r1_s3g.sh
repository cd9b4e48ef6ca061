#!/bin/bash
# Session-3 call G: the contract bench (cfg3, N=1, 100 M rows), reference arm, other configs, ncu.
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_cfg3_n1.json 2> gpurun_out/bench_cfg3_n1.err; echo "bench cfg3 rc=$?"
tail -c 2500 gpurun_out/bench_cfg3_n1.json; tail -3 gpurun_out/bench_cfg3_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_n1.json | cut -c1-600
for w in cfg2 cfg4 cfg5; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "bench $w rc=$?"
  python -c "import sys,json; d=json.loads(open('gpurun_out/bench_${w}_n1.json').read()); print(d['value'], d['ms_per_step'], d['roofline'], d['e2e']['value'], d['variant'], d['phases_per_search'])"
done
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_cfg3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_cfg3.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/plain_cfg3b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 4 -c 1 -o gpurun_out/prof_tc_main $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
