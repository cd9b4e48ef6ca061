#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2i
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest.log | cut -c1-300
timeout 300 python bench.py --workload cfg2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_cfg2.json 2> gpurun_out/${T}_bench_cfg2.err; echo "bench cfg2 rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_cfg2.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['gpu_launches'], d['verify']['ok'], d['verify']['ids_sha256'][:10])"
for B in 32 1024; do
  timeout 600 python bench.py --workload ivf --batch $B --steps 20 --warmup 3 > gpurun_out/${T}_bench_ivf_b$B.json 2> gpurun_out/${T}_bench_ivf_b$B.err; echo "bench ivf B=$B rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_ivf_b$B.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['gpu_launches'], d['verify'], d.get('coarse'))"; tail -2 gpurun_out/${T}_bench_ivf_b$B.err
done
