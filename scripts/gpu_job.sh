#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2a
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc; free -g | head -2
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest.log | cut -c1-300
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${T}_bench_cfg3.json 2> gpurun_out/${T}_bench_cfg3.err; echo "bench cfg3 rc=$?"; cut -c1-1500 gpurun_out/${T}_bench_cfg3.json; tail -3 gpurun_out/${T}_bench_cfg3.err
for w in cfg5 cfg2; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_${w}.json 2> gpurun_out/${T}_bench_${w}.err; echo "bench $w rc=$?"; cut -c1-1200 gpurun_out/${T}_bench_${w}.json
done
CMD="python bench.py --workload cfg5 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_cfg5.csv $CMD > gpurun_out/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_scan_kernel -s 7 -c 2 -o gpurun_out/${T}_prof_cfg5 $CMD > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"; tail -2 gpurun_out/${T}_ncu_full.log
timeout 200 python scripts/cublas_shape_comparator.py > gpurun_out/${T}_cublas_shape.json 2> gpurun_out/${T}_cublas_shape.err; echo "cublas rc=$?"; cat gpurun_out/${T}_cublas_shape.json
timeout 200 python scripts/cublas_shape_comparator.py --rows 100000000 --seconds 3 > gpurun_out/${T}_cublas_shape_100M.json 2>> gpurun_out/${T}_cublas_shape.err; echo "cublas100M rc=$?"; cat gpurun_out/${T}_cublas_shape_100M.json
timeout 300 python scripts/survivor_stats.py > gpurun_out/${T}_survivors.json 2> gpurun_out/${T}_survivors.err; echo "survivors rc=$?"; cat gpurun_out/${T}_survivors.json
ls -la gpurun_out | head -30
