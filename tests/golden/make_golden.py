"""Generates tests/golden/*.npz.  Run from the repo root: python tests/golden/make_golden.py

PARITY UNPINNED: faiss-cpu (the reference's arithmetic for this path) cannot be installed in the
build image and /root/reference holds no result fixture for FAISSStore.search, so these vectors are
produced by the fp64 numpy restatement (oracle/flatip_oracle.py), NOT by the reference.  What IS
taken from the reference are the inputs of its only real-index fixture,
/root/reference/tests/test_retrieval_service.py:202-212,260-261:
    corpus  = np.random.default_rng(42).random((10, 768), dtype=float32)
    queries = np.random.default_rng(42).random((2, 768)).astype(float32)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import flatip_oracle  # noqa: E402
from rag_inference_pipeline_b200.synthetic import synthetic_corpus, synthetic_queries  # noqa: E402


def reference_fixture():
    corpus = np.random.default_rng(42).random((10, 768)).astype("float32")
    queries = np.random.default_rng(42).random((2, 768)).astype("float32")
    out = {"corpus": corpus, "queries": queries}
    for k in (1, 3, 10, 12):
        D, I = flatip_oracle.flat_ip_search(queries, corpus, k)
        out[f"D_k{k}"] = D
        out[f"I_k{k}"] = I
    np.savez_compressed(os.path.join(HERE, "reference_fixture_flatip.npz"), **out)


def synthetic_cases():
    out = {}
    cases = [("f32", 3000, 64, 5, 10), ("f16", 5000, 128, 33, 100), ("bf16", 2500, 64, 8, 7)]
    for dtype, n, dim, nq, k in cases:
        corpus = synthetic_corpus(0, n, dim, dtype)
        queries = synthetic_queries(nq, dim)
        D, I = flatip_oracle.flat_ip_search(queries, corpus, k)
        tag = f"{dtype}_{n}x{dim}_q{nq}_k{k}"
        out[f"D_{tag}"] = D
        out[f"I_{tag}"] = I
        out[f"sample_{tag}"] = corpus[:4, :8]
    np.savez_compressed(os.path.join(HERE, "synthetic_cases.npz"), **out)


if __name__ == "__main__":
    reference_fixture()
    synthetic_cases()
    print("golden fixtures written to", HERE)
