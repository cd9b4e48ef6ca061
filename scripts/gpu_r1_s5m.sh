#!/bin/bash
# Session-5 call M: bench.py --workload ivf (both arms) at the reference's index shape; IVF tests after the profiling-stat change.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ivf.py -q > gpurun_out/s5m_pytest_ivf.log 2>&1; echo "pytest ivf rc=$?"; tail -2 gpurun_out/s5m_pytest_ivf.log
timeout 900 python bench.py --workload ivf > gpurun_out/s5m_bench_ivf_n1.json 2> gpurun_out/s5m_bench_ivf_n1.err; echo "bench ivf rc=$?"; cat gpurun_out/s5m_bench_ivf_n1.json; tail -3 gpurun_out/s5m_bench_ivf_n1.err
timeout 900 python bench.py --workload ivf --batch 1024 --no-cpu-baseline > gpurun_out/s5m_bench_ivf_b1024_n1.json 2> gpurun_out/s5m_bench_ivf_b1024_n1.err; echo "bench ivf b1024 rc=$?"; cat gpurun_out/s5m_bench_ivf_b1024_n1.json
timeout 900 python bench.py --workload ivf --impl reference --steps 3 --warmup 1 > gpurun_out/s5m_bench_ivf_reference.json 2> gpurun_out/s5m_bench_ivf_reference.err; echo "ref rc=$?"; cat gpurun_out/s5m_bench_ivf_reference.json
