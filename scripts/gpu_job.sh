#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2r
timeout 900 python -m pytest tests/test_gpu_ivf.py -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest ivf rc=$?"; tail -6 gpurun_out/${T}_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for B in 32 1024; do
 for c in 1 0; do
  timeout 600 python bench.py --workload ivf --batch $B --steps 20 --warmup 3 --no-cpu-baseline --ivf-coarse $c > gpurun_out/${T}_bench_ivf_b${B}_c$c.json 2> gpurun_out/${T}_bench_ivf_b${B}_c$c.err; echo "bench ivf B=$B coarse=$c rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_ivf_b${B}_c$c.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['gpu_launches'], d.get('coarse'))"; tail -2 gpurun_out/${T}_bench_ivf_b${B}_c$c.err
 done
done
