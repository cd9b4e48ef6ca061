#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_cfg3_n1.json 2> gpurun_out/bench_cfg3_n1.err; echo "bench cfg3 rc=$?"; tail -c 3000 gpurun_out/bench_cfg3_n1.json; tail -5 gpurun_out/bench_cfg3_n1.err
for w in cfg2 cfg4 cfg5; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${w}_n1.json 2> gpurun_out/bench_${w}_n1.err; echo "bench $w rc=$?"; cat gpurun_out/bench_${w}_n1.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['roofline'], d['e2e']['value'], d['variant'], d['phases_per_search'])"
done
CMD="python bench.py --rows 12500000 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_shard.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_shard.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/plain_shard2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 12 -c 4 -o gpurun_out/prof_tc_scan $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
