#!/bin/bash
# Round-2 GPU job (overwritten per gpurun call; results land in gpurun_out/ and are summarised under profiles/).
mkdir -p gpurun_out
T=r2h
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_gpu_group.py -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${T}_pytest.log | cut -c1-300
for w in cfg2 cfg5 cfg4; do
 for f in 1 0; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --opt fused_rescore=$f > gpurun_out/${T}_bench_${w}_fused$f.json 2> gpurun_out/${T}_bench_${w}_fused$f.err; echo "bench $w fused=$f rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_${w}_fused$f.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'], d['gpu_launches'], d['verify']['ok'], d['verify']['ids_sha256'][:10])"; tail -2 gpurun_out/${T}_bench_${w}_fused$f.err
 done
done
timeout 300 python bench.py --rows 12500000 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_shard.json 2> gpurun_out/${T}_bench_shard.err; echo "bench shard rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/${T}_bench_shard.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['whole_search_frac'], d['roofline']['kernel_share_of_step'])"
