"""Developer diagnostics for a GPU box (not part of the test suite): verbose parity + timing.

usage: python scripts/dev_gpu_check.py {probe|parity|time} [args]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np


def probe():
    os.system("nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv")
    os.system("nproc; free -g | head -2")
    try:
        import faiss  # noqa
        print("faiss importable:", faiss.__version__)
    except Exception as e:
        print("faiss not importable:", e)
    print("baseline/_ref:", os.path.exists(os.path.join(ROOT, "baseline", "_ref")))


def parity(variant, dtype, n, dim, nq, k):
    import flatip_oracle, c_oracle
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.synthetic import synthetic_queries
    n, dim, nq, k = int(n), int(dim), int(nq), int(k)
    corpus = c_oracle.synthetic_corpus(0, n, dim, dtype)
    queries = synthetic_queries(nq, dim)
    idx = FlatIPIndex(dim, dtype, 0)
    idx.add_synthetic(n, 1234, 0)
    rec = idx.reconstruct_n(0, min(n, 1000))
    print("generator match:", np.array_equal(rec, corpus[:len(rec)]))
    idx.set_variant(variant)
    t = time.time()
    D, I = idx.search(queries, k)
    print(f"search {variant} {dtype} n={n} d={dim} nq={nq} k={k}: {time.time()-t:.3f}s "
          f"launches={idx.stat('launches')} phases={idx.stat('phases')} overflow={idx.stat('overflow_queries')} "
          f"cap={idx.stat('candidate_capacity')} variant={idx.last_variant()}")
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k)
    res = flatip_oracle.compare_with_oracle(D, I, Dr, Ir, rtol=flatip_oracle.SCORE_RTOL[dtype])
    print(res)
    if res["id_mismatch_total"]:
        bad = np.argwhere(I != Ir)[:5]
        for q, j in bad:
            print("  mismatch q", q, "pos", j, "gpu", I[q, j], D[q, j], "ref", Ir[q, j], Dr[q, j])
    idx.close()


def timeit(variant, dtype, n, dim, nq, k, iters=5):
    import torch
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.synthetic import synthetic_queries
    n, dim, nq, k, iters = int(n), int(dim), int(nq), int(k), int(iters)
    idx = FlatIPIndex(dim, dtype, 0)
    t = time.time()
    idx.add_synthetic(n, 1234, 0)
    print(f"generated {n}x{dim} {dtype} in {time.time()-t:.2f}s")
    idx.set_variant(variant)
    if os.environ.get("FIP_CTA_PAIR"):
        idx.set_option("cta_pair", int(os.environ["FIP_CTA_PAIR"]))
    if os.environ.get("FIP_LIVE"):
        idx.set_option("live_thresholds", int(os.environ["FIP_LIVE"]))
    if os.environ.get("FIP_SPLIT"):
        idx.set_option("query_split", int(os.environ["FIP_SPLIT"]))
    idx.set_option("profile_events", 1)
    q = torch.from_numpy(synthetic_queries(nq, dim)).cuda()
    D = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    I = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        idx.search_device(q.data_ptr(), nq, k, D.data_ptr(), I.data_ptr(), s)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        idx.search_device(q.data_ptr(), nq, k, D.data_ptr(), I.data_ptr(), s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    es = 4 if dtype == "f32" else 2
    gbs = n * dim * es / ms / 1e6
    tfs = 2.0 * nq * n * dim / ms / 1e9
    print(f"TIME {variant} {dtype} n={n} d={dim} nq={nq} k={k}: {ms:.3f} ms/search  {nq/ms*1e3:.0f} QPS  "
          f"{gbs:.0f} GB/s  {tfs:.1f} TFLOP/s  launches={idx.stat('launches')} phases={idx.stat('phases')} "
          f"overflow={idx.stat('overflow_queries')} uncert={idx.stat('uncertified_queries')} split={idx.stat('query_split_used')} cta_group={idx.stat('cta_group')} "
          f"scan_ms={idx.stat('scan_time_ns')/1e6:.3f}")
    sc = [idx.stat(f"scan_ns_{i}") / 1e6 for i in range(16) if idx.stat(f"scan_ns_{i}") >= 0]
    gp = [idx.stat(f"gap_ns_{i}") / 1e6 for i in range(16) if idx.stat(f"gap_ns_{i}") >= 0]
    print("  scans(ms):", " ".join(f"{v:.3f}" for v in sc), "| gaps(ms):", " ".join(f"{v:.3f}" for v in gp))
    idx.close()


if __name__ == "__main__":
    cmd = sys.argv[1]
    {"probe": probe, "parity": parity, "time": timeit}[cmd](*sys.argv[2:])
