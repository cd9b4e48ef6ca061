#!/bin/bash
# Round-2 GPU job: the checks the driver runs at round end (GPU tests with -x, smoke, default bench) at HEAD.
mkdir -p gpurun_out
T=r2s
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/${T}_smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg3_n1.json 2> gpurun_out/${T}_bench_cfg3_n1.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg3_n1.json')); r=d['roofline']; print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'frac', r['frac'], 'whole', r['whole_search_frac'], d['verify']['ok'], d['verify']['ids_sha256'][:12])"
