"""GPU test of ``fip_search_rows_static`` -- the search entry point for hosts that are not allowed to
allocate device memory (a retriever squeezed in beside an LLM server that already owns HBM, or a
process whose pooling allocator bans raw ``cudaMalloc``).  Its workspace is a static device array
shared by every call, so what matters is (a) that it returns what the allocating entry point
``fip_search_device`` returns and (b) that a call is not disturbed by what the PREVIOUS call left
in that array: every case below runs several calls back to back with different shapes.

Expected values come from the existing entry point through the same C ABI: bit-equal against its
streaming variant (same kernels, same arithmetic order), and inside the repo's usual score
tolerance / near-tie rule against its default (tensor-core) variant.

The file sorts last on purpose: ``pytest -x`` stops at the first failure, and a failure here must
not hide the parity tests.  It passes alone in a fresh process and after the other GPU tests.
"""
import numpy as np
import pytest

import flatip_oracle
from flatip_oracle import SCORE_RTOL
from rag_inference_pipeline_b200 import synthetic

pytestmark = pytest.mark.gpu

LOWEST = np.finfo(np.float32).min


def _torch_dtype(dtype):
    import torch
    return {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[dtype]


class _Bench:
    """A corpus held twice: inside a FlatIPIndex (the existing, allocating path) and as a plain torch
    tensor in the storage dtype (what a no-allocation caller owns)."""

    def __init__(self, dtype, n, dim, rows32=None):
        import torch
        from rag_inference_pipeline_b200 import FlatIPIndex

        self.dtype, self.n, self.dim = dtype, n, dim
        self.index = FlatIPIndex(dim, dtype, 0)
        if rows32 is None:
            self.index.add_synthetic(n, synthetic.CORPUS_SEED, 0)
            rows32 = self.index.reconstruct_n(0, n) if n else np.zeros((0, dim), np.float32)
        else:
            self.index.add(rows32)
            rows32 = self.index.reconstruct_n(0, n)
        # the stored values are exactly representable in the storage dtype: this cast is lossless
        self.rows = torch.from_numpy(rows32).to("cuda:0").to(_torch_dtype(dtype)).contiguous()
        torch.cuda.synchronize()

    def existing(self, queries, k, variant):
        import torch
        self.index.set_variant(variant)
        q = torch.from_numpy(queries).to("cuda:0")
        D = torch.empty((len(queries), k), dtype=torch.float32, device="cuda:0")
        I = torch.empty((len(queries), k), dtype=torch.int64, device="cuda:0")
        self.index.search_device(q.data_ptr(), len(queries), k, D.data_ptr(), I.data_ptr())
        torch.cuda.synchronize()
        return D.cpu().numpy(), I.cpu().numpy()

    def static(self, queries, k, id_base=0):
        import torch
        from rag_inference_pipeline_b200.index import search_rows_static

        q = torch.from_numpy(queries).to("cuda:0")
        # poisoned outputs: every slot must be written by the call
        D = torch.full((len(queries), k), float("nan"), dtype=torch.float32, device="cuda:0")
        I = torch.full((len(queries), k), -7, dtype=torch.int64, device="cuda:0")
        torch.cuda.synchronize()
        search_rows_static(0, self.rows.data_ptr(), self.n, self.dim, self.dtype, q.data_ptr(), len(queries), k,
                           D.data_ptr(), I.data_ptr(), stream=0, id_base=id_base)
        # blocking call: no synchronisation needed before the copy back
        return D.cpu().numpy(), I.cpu().numpy()

    def close(self):
        self.index.close()


def _agree(bench, queries, k):
    D, I = bench.static(queries, k)
    Ds, Is = bench.existing(queries, k, "stream")
    assert np.array_equal(I, Is)
    assert np.array_equal(D.view(np.uint32), Ds.view(np.uint32))   # same kernels: same bits
    Dt, It = bench.existing(queries, k, "auto")
    res = flatip_oracle.compare_with_oracle(D, I, Dt, It, rtol=SCORE_RTOL[bench.dtype])
    assert res["score_ok"], res
    assert res["id_mismatch_outside_window"] == 0, res
    return D, I


@pytest.mark.parametrize("dtype,n,dim", [("f16", 30000, 128), ("bf16", 20000, 64), ("f32", 25000, 96)])
def test_static_workspace_search_matches_existing_entry_point(dtype, n, dim):
    bench = _Bench(dtype, n, dim)
    try:
        q_all = synthetic.synthetic_queries(64, dim)
        first = _agree(bench, q_all, 100)            # fills the lists of all 64 queries
        _agree(bench, q_all[40:47], 10)              # fewer queries, smaller k, on top of the leftovers
        _agree(bench, q_all[:1], 1)
        again = _agree(bench, q_all, 100)            # and the first call once more
        assert np.array_equal(first[1], again[1]) and np.array_equal(first[0], again[0])
        # global ids of a shard
        D, I = bench.static(q_all[:3], 5, id_base=1_000_000)
        assert np.array_equal(I - 1_000_000, first[1][:3, :5]) and np.array_equal(D, first[0][:3, :5])
    finally:
        bench.close()


def test_static_workspace_overflow_falls_back_to_short_phases():
    """Rows sorted by ascending score: every row beats the running threshold, the candidate lists of
    the fast schedule overflow and the call must redo the search with overflow-proof phases."""
    n, dim = 300_000, 32
    rows = np.zeros((n, dim), np.float32)
    rows[:, 0] = np.arange(n, dtype=np.float32)     # score of row i for e_0 is i
    rows[:, 1] = 1.0
    q = np.zeros((3, dim), np.float32)
    q[0, 0] = 1.0
    q[1, 0] = -1.0                                  # descending for this one
    q[2, 1] = 1.0                                   # all equal: ties broken by id
    bench = _Bench("f32", n, dim, rows32=rows)
    try:
        for k in (10, 300):
            D, I = _agree(bench, q, k)
            assert np.array_equal(I[0], np.arange(n - 1, n - 1 - k, -1))
            assert np.array_equal(I[1], np.arange(k))
            assert np.array_equal(I[2], np.arange(k))
    finally:
        bench.close()


def test_static_workspace_padding_and_limits():
    from rag_inference_pipeline_b200._native import FlatIPError
    from rag_inference_pipeline_b200.index import STATIC_MAX_QUERIES

    bench = _Bench("f16", 40, 64)
    try:
        q = synthetic.synthetic_queries(STATIC_MAX_QUERIES + 1, 64)
        D, I = _agree(bench, q[:4], 50)              # k > rows: faiss padding
        assert np.all(I[:, 40:] == -1) and np.all(D[:, 40:] == LOWEST)
        with pytest.raises(FlatIPError, match="at most 64 queries"):
            bench.static(q, 10)
        with pytest.raises(FlatIPError, match=r"k must be in \[1, 2048\]"):
            bench.static(q[:2], 2049)
        D, I = _agree(bench, q[:STATIC_MAX_QUERIES], 7)   # the largest batch, right after the refusals
        bench.n = 0                                   # an empty corpus: all padding
        D, I = bench.static(q[:2], 3)
        assert np.all(I == -1) and np.all(D == LOWEST)
    finally:
        bench.close()
