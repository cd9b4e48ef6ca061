"""Developer timing of the IVF-Flat operator on a GPU box (not part of the test suite).

usage: python scripts/dev_ivf_check.py N DIM NLIST NPROBE "B1,B2,.." K [iters]
Corpus: standard-normal rows like /root/reference/scripts/create_test_docs.py:87-96, trained on the
first 10000 rows (as the reference does), queries = perturbed corpus rows.
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main():
    import torch
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    n, dim, nlist, nprobe = (int(float(a)) for a in sys.argv[1:5])
    batches = [int(b) for b in sys.argv[5].split(",")]
    k = int(sys.argv[6])
    iters = int(sys.argv[7]) if len(sys.argv) > 7 else 10
    rng = np.random.default_rng(0)
    index = IVFFlatL2Index(dim, nlist, nprobe=nprobe)
    if "IVF_TC_MIN" in os.environ:
        index.set_option("tensor_min_queries", int(os.environ["IVF_TC_MIN"]))
    t = time.time()
    first = rng.standard_normal((min(n, 100000), dim), dtype=np.float32)
    ntrain = int(os.environ.get("IVF_TRAIN", 10000))
    index.train(first[:ntrain])
    print(f"train: {time.time() - t:.2f}s", flush=True)
    t = time.time()
    index.add(first)
    done = len(first)
    keep = first[:4096].copy()
    while done < n:
        m = min(100000, n - done)
        index.add(rng.standard_normal((m, dim), dtype=np.float32))
        done += m
    sizes = index.list_sizes()
    print(f"add {n} rows: {time.time() - t:.2f}s; list sizes min/mean/max = {sizes.min()}/{sizes.mean():.0f}/{sizes.max()}", flush=True)
    for B in batches:
        q = (keep[rng.integers(0, len(keep), B)] + 0.3 * rng.standard_normal((B, dim), dtype=np.float32)).astype(np.float32)
        qd = torch.from_numpy(q).cuda()
        D = torch.empty((B, k), dtype=torch.float32, device="cuda")
        I = torch.empty((B, k), dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        for _ in range(3):
            index.search_device(qd.data_ptr(), B, k, D.data_ptr(), I.data_ptr())
        t = time.perf_counter()
        for _ in range(iters):
            index.search_device(qd.data_ptr(), B, k, D.data_ptr(), I.data_ptr())
        dt = (time.perf_counter() - t) / iters
        rows, pairs = index.stat("rows_scanned"), index.stat("pairs_scored")
        t = time.perf_counter()
        for _ in range(iters):
            index.search(q, k)
        dth = (time.perf_counter() - t) / iters
        print(f"B={B} k={k} nprobe={nprobe}: {dt * 1e3:.3f} ms/search device ({B / dt:.0f} QPS), host API {dth * 1e3:.3f} ms; "
              f"tiles={index.stat('tiles')} (16-query: {index.stat('tiles16')}) rows_scanned={rows} "
              f"({rows * dim * 4 / dt / 1e9:.0f} GB/s of list bytes) pairs={pairs} ({pairs * dim / dt / 1e12:.2f} TFMA/s) "
              f"plan={index.stat('plan_us')} us launches={index.stat('launches')} phases={index.stat('phases')} "
              f"overflow={index.stat('overflow')} tensor_items={index.stat('tensor_items')} tensor_pairs={index.stat('tensor_pairs')}", flush=True)
    index.close()


if __name__ == "__main__":
    main()
