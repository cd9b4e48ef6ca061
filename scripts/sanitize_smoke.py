#!/usr/bin/env python
"""Smoke-sized pass over every kernel family, for `compute-sanitizer --tool {memcheck,racecheck,
synccheck} python scripts/sanitize_smoke.py` (one tool per gpurun call, see B200_PROFILING.md).
numpy + ctypes only (no torch import: the sanitizer instruments every CUDA call of the process).
Covers: K1 tensor-core scan single-CTA and CTA-pair, two-pass and deferred epilogue, dense and
filter phases; K2 streaming scan; K3 select (256- and 1024-thread forms) and rescore; device
dedupe; K4 exchange + merge through a 3-shard group on one device; IVF-Flat scan + tensor path.
Every result is checked against the CPU oracle, so a sanitizer-clean run is also a correct one."""
import os
import sys

os.environ["FLATIP_B200_ALLOW_SHARED_DEVICE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np  # noqa: E402

import flatip_oracle  # noqa: E402  (checker only)
import ivf_oracle  # noqa: E402
from rag_inference_pipeline_b200 import FlatIPIndex, IVFFlatL2Index, synthetic  # noqa: E402
from rag_inference_pipeline_b200.group import FlatIPGroup  # noqa: E402


def check(tag, D, I, q, corpus, k, rtol):
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    res = flatip_oracle.compare_with_oracle(D, I, Dr, Ir, rtol=rtol)
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, (tag, res)
    print("ok", tag, flush=True)


n, dim = 9000, 128
for dtype, rtol in (("f16", 1e-3), ("f32", 1e-5)):
    corpus = synthetic.synthetic_corpus(0, n, dim, dtype)
    idx = FlatIPIndex(dim, dtype, 0)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    for nq, k, variant, opts in ((3, 10, "stream", {}), (40, 10, "tensor", {}), (300, 20, "tensor", {}),
                                 (140, 300, "tensor", {"defer_epilogue": 2}), (40, 300, "tensor", {"defer_epilogue": 1})):
        q = synthetic.synthetic_queries(nq, dim)
        idx.set_variant(variant)
        for key, val in opts.items():
            idx.set_option(key, val)
        D, I = idx.search(q, k)
        check(f"{dtype} nq={nq} k={k} {variant} {opts}", D, I, q, corpus, k, rtol)
        idx.set_option("defer_epilogue", 0)
    idx.set_option("dedupe", 2)
    q = synthetic.synthetic_queries(12, dim)[[0, 1, 0, 2, 1, 11]]
    D, I = idx.search(q, 10)
    check(f"{dtype} dedupe", D, I, q, corpus, 10, rtol)
    idx.close()

corpus = synthetic.synthetic_corpus(0, n, dim, "bf16")
g = FlatIPGroup(dim, "bf16", [0, 0, 0], n)
g.add_synthetic(n, synthetic.CORPUS_SEED)
for nq, k in ((50, 10), (200, 130)):
    q = synthetic.synthetic_queries(nq, dim)
    D, I = g.search(q, k)
    check(f"group x3 nq={nq} k={k}", D, I, q, corpus, k, 1e-3)
g.close()

rng = np.random.default_rng(0)
x = rng.standard_normal((6000, 64)).astype(np.float32)
q = (x[:150] + 0.3 * rng.standard_normal((150, 64))).astype(np.float32)
ivf = IVFFlatL2Index(64, 16, device=0, nprobe=4)
ivf.train(x[:2000])
ivf.add(x)
D, I = ivf.search(q, 10)
sizes, codes, ids = ivf.export_lists()
Dr, Ir, _, D64 = ivf_oracle.ivf_search(q, ivf.centroids(), sizes, codes, ids, 11, 4)
res = ivf_oracle.compare_ivf(D, I, Dr[:, :10], Ir[:, :10], D64)
assert res["dist_bad"] == 0 and res["id_mismatch_outside_ties"] == 0, res
print("ok ivf", flush=True)
ivf.close()
print("SANITIZE_SMOKE_OK", flush=True)
