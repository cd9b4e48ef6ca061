#!/bin/bash
# Session-5 call A: HEAD verification: GPU tests, smoke, default bench (cfg3 N=1), other workloads, launch list.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/s5a_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/s5a_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s5a_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/s5a_smoke.log
timeout 900 python bench.py > gpurun_out/s5a_bench_cfg3_n1.json 2> gpurun_out/s5a_bench_cfg3_n1.err; echo "bench rc=$?"; cat gpurun_out/s5a_bench_cfg3_n1.json
for W in cfg2 cfg4 cfg5; do
timeout 600 python bench.py --workload $W --no-cpu-baseline > gpurun_out/s5a_bench_${W}_n1.json 2> gpurun_out/s5a_bench_${W}_n1.err; echo "bench $W rc=$?"; cat gpurun_out/s5a_bench_${W}_n1.json
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s5a_launches_cfg3.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/s5a_ncu.log 2>&1; echo "ncu rc=$?"
