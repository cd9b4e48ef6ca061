#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2o
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${T}_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/${T}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_cfg3_n1.json 2> gpurun_out/${T}_bench_cfg3_n1.err; echo "bench default rc=$?"; cut -c1-3500 gpurun_out/${T}_bench_cfg3_n1.json; tail -2 gpurun_out/${T}_bench_cfg3_n1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; echo "reference rc=$?"; cut -c1-1800 gpurun_out/${T}_bench_reference.json
timeout 300 python scripts/sweep.py --batches 1,2,3,4 > gpurun_out/${T}_sweep.jsonl 2> gpurun_out/${T}_sweep.err; cut -c1-230 gpurun_out/${T}_sweep.jsonl
