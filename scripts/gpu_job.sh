#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2l
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${T}_pytest.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"
for w in cfg2 cfg5 cfg4; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_${w}.json 2> gpurun_out/${T}_bench_${w}.err; echo "bench $w rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_${w}.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['gpu_launches'], d['verify']['ok'], d['verify']['ids_sha256'][:10])"; tail -2 gpurun_out/${T}_bench_${w}.err
done
timeout 300 python bench.py --rows 12500000 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_shard.json 2> gpurun_out/${T}_bench_shard.err; echo "bench shard rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_shard.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['gpu_launches'])"
timeout 600 python scripts/sweep.py --batches 1,2,4,8,32,128,256,512,1024 > gpurun_out/${T}_sweep.jsonl 2> gpurun_out/${T}_sweep.err; cut -c1-200 gpurun_out/${T}_sweep.jsonl
