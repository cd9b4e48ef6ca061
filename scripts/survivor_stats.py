#!/usr/bin/env python
"""Does the Irwin-Hall(4) synthetic corpus behave like the reference's Gaussian corpus
(/root/reference/scripts/create_test_docs.py:87-96, rng.standard_normal) as far as the kernels can
tell?  What the filter cost depends on is the number of candidates a filter launch appends (tail
mass above the running k-th best score) and the near-tie frequency at the top.  This runs the same
search over (a) the Philox / Irwin-Hall corpus and (b) numpy standard-normal rows of the same shape
and prints, for each, the list entries after the filter launch, the longest list, uncertified /
overflowed queries and the mean gap between the k-th and (k+1)-th score in units of the score.

usage: python scripts/survivor_stats.py [--rows 2000000] [--dim 768] [--batch 1024] [--k 100] [--dtype f16]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rag_inference_pipeline_b200 import FlatIPIndex  # noqa: E402
from rag_inference_pipeline_b200.synthetic import CORPUS_SEED, synthetic_queries  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=2_000_000)
ap.add_argument("--dim", type=int, default=768)
ap.add_argument("--batch", type=int, default=1024)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--dtype", default="f16")
args = ap.parse_args()
q = synthetic_queries(args.batch, args.dim)
out = {}
for name in ("irwin_hall", "gaussian"):
    idx = FlatIPIndex(args.dim, args.dtype, 0, capacity_rows=args.rows)
    if name == "irwin_hall":
        idx.add_synthetic(args.rows, CORPUS_SEED, 0)
    else:
        rng = np.random.default_rng(7)
        for r0 in range(0, args.rows, 100_000):
            idx.add(rng.standard_normal((min(100_000, args.rows - r0), args.dim), dtype=np.float32))
    idx.set_option("debug_counts", 1)
    D, I = idx.search(q, args.k + 1)
    D, I = idx.search(q, args.k)
    Dk1, _ = idx.search(q, args.k + 1)
    gap = (Dk1[:, args.k - 1] - Dk1[:, args.k]) / np.abs(Dk1[:, args.k - 1])
    out[name] = {"list_entries_after_filter": idx.stat("list_entries_after_filter"),
                 "per_query": idx.stat("list_entries_after_filter") / args.batch,
                 "list_max": idx.stat("list_max_after_filter"), "phases": idx.stat("phases"),
                 "uncertified": idx.stat("uncertified_queries"), "overflow": idx.stat("overflow_queries"),
                 "kth_score_mean": float(D[:, -1].mean()), "top1_score_mean": float(D[:, 0].mean()),
                 "rel_gap_k_to_k1_mean": float(gap.mean()), "rel_gap_k_to_k1_min": float(gap.min())}
    idx.close()
out["config"] = vars(args)
print(json.dumps(out))
