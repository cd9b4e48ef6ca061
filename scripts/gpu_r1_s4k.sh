#!/bin/bash
# Session-4 call K: the contract bench (cfg3, N=1, 100 M rows), reference arm, other configs, ncu.
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/s4k_bench_cfg3_n1.json 2> gpurun_out/s4k_bench_cfg3_n1.err; echo "bench cfg3 rc=$?"
tail -c 1500 gpurun_out/s4k_bench_cfg3_n1.json; tail -3 gpurun_out/s4k_bench_cfg3_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s4k_bench_ref_n1.json 2> gpurun_out/s4k_bench_ref_n1.err; echo "ref rc=$?"; cut -c1-700 gpurun_out/s4k_bench_ref_n1.json
for w in cfg2 cfg4 cfg5; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/s4k_bench_${w}_n1.json 2> gpurun_out/s4k_bench_${w}_n1.err; echo "bench $w rc=$?"
  python -c "import sys,json; d=json.loads(open('gpurun_out/s4k_bench_${w}_n1.json').read()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['e2e']['value'], d['variant'], d['phases_per_search'])"
done
python bench.py --rows 12500000 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/s4k_bench_shard12p5M_n1.json 2> gpurun_out/s4k_bench_shard.err; echo "bench shard rc=$?"
python -c "import sys,json; d=json.loads(open('gpurun_out/s4k_bench_shard12p5M_n1.json').read()); print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['clocks'])"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/s4k_plain_cfg3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/s4k_launches_cfg3.csv $CMD > gpurun_out/s4k_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/s4k_plain_cfg3b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 4 -c 1 -o gpurun_out/s4k_prof_tc_main $CMD > gpurun_out/s4k_ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/s4k_ncu_full.log
CMD1="python bench.py --workload cfg4 --batch 1 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD1 > gpurun_out/s4k_plain_b1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:stream_scan_kernel -s 2 -c 1 -o gpurun_out/s4k_prof_k2_main $CMD1 > gpurun_out/s4k_ncu_full_k2.log 2>&1; echo "ncu full k2 rc=$?"
ls -la gpurun_out | tail -20
