"""GPU parity tests of the IVF-Flat (L2) operator (include/ivfflat_b200.h) against the
numpy restatement of faiss IndexIVFFlat search (oracle/ivf_oracle.py), through the C ABI.

The oracle searches the SAME index (centroids and inverted lists exported from the GPU index, or
lists it built itself and imported with add_preassigned), so the comparison pins the search path:
coarse probe selection, list scan, top-k, distance / id / padding contract.
Acceptance: squared L2 distances within 1e-5 relative (+1e-4 absolute: the operator evaluates
||q||^2 - (2<q,x> - ||x||^2) in fp32) of the fp64 oracle; ids bit-exact outside near-tie windows.
"""
import types

import numpy as np
import pytest

import ivf_oracle

pytestmark = pytest.mark.gpu


def _data(n, d, nq, seed=0, clustered=False):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32)
    if clustered:
        x += 4.0 * rng.standard_normal((16, d)).astype(np.float32)[rng.integers(0, 16, n)]
    q = x[rng.integers(0, n, nq)] + 0.5 * rng.standard_normal((nq, d)).astype(np.float32)
    return x, q.astype(np.float32)


def _stable_probe_rows(q, centroids, nprobe, scale=1.0):
    """Queries whose nprobe-th / (nprobe+1)-th centroid distances are not a near tie."""
    d = np.sort(ivf_oracle.sq_l2(q, centroids), axis=1)
    if nprobe >= centroids.shape[0]:
        return np.ones(len(q), dtype=bool)
    return (d[:, nprobe] - d[:, nprobe - 1]) > 1e-3 * scale * scale + 2e-5 * d[:, nprobe]


def _check(index, q, k, nprobe, scale=1.0):
    sizes, codes, ids = index.export_lists()
    cent = index.centroids()
    D, I = index.search(q, k, nprobe=nprobe)
    Dr, Ir, _, D64 = ivf_oracle.ivf_search(q, cent, sizes, codes, ids, k + 1, nprobe)
    ok = _stable_probe_rows(q, cent, nprobe, scale)
    assert ok.mean() > 0.9
    res = ivf_oracle.compare_ivf(D[ok], I[ok], Dr[ok, :k], Ir[ok, :k], D64[ok], atol=1e-4 * scale * scale)
    assert res["dist_bad"] == 0 and res["pad_bad"] == 0 and res["id_mismatch_outside_ties"] == 0, res
    assert res["id_mismatch_total"] <= max(2, D.size // 200), res
    # rows are ascending and padding sits at the end
    valid = I >= 0
    assert np.all(valid[:, :-1] >= valid[:, 1:])
    assert np.all((np.diff(D, axis=1) >= 0) | ~valid[:, 1:])
    return D, I


@pytest.mark.parametrize("n,d,nlist,nq,k,nprobe", [
    (20000, 64, 32, 37, 10, 4),
    (50000, 768, 64, 64, 100, 8),
    (30000, 128, 256, 200, 10, 64),     # many queries per list: 16-query tiles
    (6000, 96, 64, 5, 50, 1),           # k > rows of the probed list: padding
    (20000, 64, 2, 9, 10, 1),           # lists longer than the candidate capacity: split segments
    (5000, 32, 16, 1, 1, 16),           # nprobe == nlist: exhaustive
    (50000, 64, 16, 20, 1024, 8),       # large k: 16 K-entry candidate lists, dense phase of 16 K rows
])
def test_ivf_search_matches_oracle(n, d, nlist, nq, k, nprobe):
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(n, d, nq, seed=n + d)
    index = IVFFlatL2Index(d, nlist)
    index.train(x[: min(n, 40 * nlist)])
    assert index.is_trained
    index.add(x)
    assert index.ntotal == n
    D, I = _check(index, q, k, nprobe)
    if nprobe == nlist:   # exhaustive probing == exact flat L2 search
        d64 = ivf_oracle.sq_l2(q, x)
        assert np.array_equal(I[:, 0], np.argmin(d64, axis=1))
    assert index.stat("overflow") == 0
    index.close()


def test_ivf_safe_mode_and_fixed_thresholds_agree():
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(40000, 64, 50, seed=5, clustered=True)
    index = IVFFlatL2Index(64, 8)
    index.train(x[:2000])
    index.add(x)
    D0, I0 = _check(index, q, 100, 4)
    index.set_option("force_safe", 1)
    D1, I1 = index.search(q, 100, nprobe=4)
    assert index.stat("phases") >= 2 and index.stat("candidate_capacity") == 16384
    index.set_option("force_safe", 0)
    index.set_option("live_thresholds", 0)
    D2, I2 = index.search(q, 100, nprobe=4)
    index.set_option("live_thresholds", 1)
    index.set_option("max_queries_per_tile", 4)
    D3, I3 = index.search(q, 100, nprobe=4)
    for D, I in ((D1, I1), (D2, I2), (D3, I3)):
        assert np.array_equal(I, I0) and np.array_equal(D, D0)
    index.close()


def test_ivf_tensor_core_path_for_heavily_probed_lists():
    """Lists probed by >= 16 queries of the batch go through the tcgen05 tile kernel (tf32 filter +
    exactness window + exact rescoring): same results as the FFMA2-only path, bit for bit."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(60000, 256, 300, seed=21)
    index = IVFFlatL2Index(256, 64)
    index.train(x[:4000])
    index.add(x)
    D0, I0 = _check(index, q, 20, 16)
    assert index.stat("tensor_items") > 0 and index.stat("tensor_pairs") > 0
    assert index.stat("tensor_f16") == 1          # default: fp16 filter copy of the lists (kind::f16)
    index.set_option("tensor_format", 1)          # tf32 straight from the fp32 rows
    D2, I2 = index.search(q, 20, nprobe=16)
    assert index.stat("tensor_items") > 0 and index.stat("tensor_f16") == 0
    assert np.array_equal(I0, I2) and np.array_equal(D0, D2)
    index.set_option("tensor_format", 0)
    index.set_option("tensor_min_queries", 0)
    D1, I1 = index.search(q, 20, nprobe=16)
    assert index.stat("tensor_items") == 0
    assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    index.set_option("tensor_min_queries", 8)
    xc, qc = _data(30000, 64, 200, seed=22, clustered=True)   # large norms, small distances
    index2 = IVFFlatL2Index(64, 16)
    index2.train(xc[:2000])
    index2.add(xc)
    _check(index2, qc, 50, 8)
    # (cluster offsets of norm ~32 next to distances of ~16: the tf32 window may be too wide for
    # k = 50, in which case the batch was redone on the fp32 filter)
    assert index2.stat("tensor_items") > 0 or index2.stat("crowded_window") == 1
    index.close()
    index2.close()


@pytest.mark.parametrize("scale,expect_f16", [(3.0e4, 0), (1.0e-3, None), (1.0, 1)])
def test_ivf_fp16_filter_copy_respects_fp16_range(scale, expect_f16):
    """The fp16 filter copy is used only when the lists and the queries fit fp16's range (norm test);
    components in fp16's subnormal range are covered by the window's absolute term (or the crowded
    fallback).  Results match the oracle at every magnitude, and the incremental add drops and
    rebuilds the copy."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(30000, 128, 120, seed=77)
    x = (x * scale).astype(np.float32)
    q = (q * scale).astype(np.float32)
    index = IVFFlatL2Index(128, 32)
    index.train(x[:3000])
    index.add(x[:20000])
    _check(index, q, 10, 12, scale)
    assert index.stat("device_plan") == 1
    if expect_f16 is not None and not index.stat("crowded_window"):
        assert index.stat("tensor_f16") == expect_f16
    index.add(x[20000:])                       # rebuild: the copy is dropped and rebuilt on next use
    _check(index, q, 10, 12, scale)
    if expect_f16 is not None and not index.stat("crowded_window"):
        assert index.stat("tensor_f16") == expect_f16
    if scale == 1.0:                           # a query outside fp16's range switches the batch to tf32
        qbig = q.copy()
        qbig[3] *= 1.0e4
        D, I = index.search(qbig, 10, nprobe=12)
        assert index.stat("tensor_f16") == 0 or index.stat("crowded_window") == 1
        D1, I1 = index.search(q, 10, nprobe=12)
        keep = np.arange(len(q)) != 3
        assert np.array_equal(I[keep], I1[keep]) and np.array_equal(D[keep], D1[keep])
    index.close()


def test_ivf_crowded_tf32_window_falls_back_to_fp32_filter():
    """A cluster of near-duplicates tighter than tf32 resolves overflows the tensor-core filter's
    exactness window; the batch is redone with the fp32 (FFMA2) filter and still matches the oracle."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    rng = np.random.default_rng(31)
    x = (3.0 * rng.standard_normal((20000, 64))).astype(np.float32)
    centre = (3.0 * rng.standard_normal(64)).astype(np.float32)
    x[5000:5400] = centre + 0.05 * rng.standard_normal((400, 64)).astype(np.float32)
    q = (centre + 0.05 * rng.standard_normal((40, 64))).astype(np.float32)
    index = IVFFlatL2Index(64, 16)
    index.train(x[:3000])
    index.add(x)
    _check(index, q, 10, 4)
    assert index.stat("crowded_window") == 1
    index.close()


def test_ivf_duplicate_cluster_larger_than_the_window():
    """more than 2k + 64 EXACT duplicates of one row crowd even the fp32 window.  The
    batch must then be redone with the wide window (not silently truncated by approximate score):
    the k results of a query sitting on the cluster are the k lowest ids of the duplicates."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    rng = np.random.default_rng(5)
    n, d, k = 20000, 64, 10
    x = (3.0 * rng.standard_normal((n, d))).astype(np.float32)
    dup = np.sort(rng.choice(n, size=300, replace=False))      # 300 > 2k + 64 = 84
    x[dup] = x[dup[0]]
    q = np.concatenate([x[dup[:1]], x[rng.integers(0, n, 15)] + 0.5 * rng.standard_normal((15, d)).astype(np.float32)])
    index = IVFFlatL2Index(d, 16)
    index.train(x[:3000])
    index.add(x)
    D, I = _check(index, q.astype(np.float32), k, 4)
    assert index.stat("crowded_window") == 1 and index.stat("crowded_window_unresolved") == 0
    assert I[0].tolist() == dup[:k].tolist() and np.all(np.abs(D[0]) < 1e-3)
    index.close()


@pytest.mark.parametrize("n,d,nlist,nq,k,nprobe,tc_min,qb", [
    (30000, 128, 256, 200, 10, 64, 8, 16),     # tensor items + small FFMA2 tiles
    (30000, 128, 256, 200, 10, 64, 0, 16),     # FFMA2 only: 16-query tiles with remainders
    (30000, 128, 256, 333, 10, 32, 0, 4),      # 4-query tiles only
    (4000, 64, 512, 70, 20, 100, 8, 16),       # many short and empty lists
    (20000, 64, 2, 300, 10, 2, 8, 16),         # two long lists, 300 queries each: 3 groups of <= 128
    (20000, 64, 16, 1, 5, 3, 8, 16),           # single query
])
def test_ivf_device_planner_matches_host_planner(n, d, nlist, nq, k, nprobe, tc_min, qb):
    """The plan kernels (ivf_plan.cuh) and the host planner (build_plan) must describe the same
    work: equal tile / item / row / pair counts and bit-identical search results."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(n, d, nq, seed=n + nlist + nq)
    index = IVFFlatL2Index(d, nlist)
    index.train(x[: min(n, 40 * nlist)])
    index.add(x)
    index.set_option("tensor_min_queries", tc_min)
    index.set_option("max_queries_per_tile", qb)
    index.set_option("device_plan", 2)   # also for batches the default would plan on the host
    keys = ("phases", "tiles", "tiles16", "rows_scanned", "pairs_scored", "distinct_rows", "tensor_items",
            "tensor_pairs", "overflow")
    D1, I1 = _check(index, q, k, nprobe)
    crowded = index.stat("crowded_window")
    dev = {key: index.stat(key) for key in keys}
    if not crowded:   # (a crowded rerun happens with the tensor path off: both planners again below)
        assert index.stat("device_plan") == 1
    index.set_option("device_plan", 0)
    D0, I0 = index.search(q, k, nprobe=nprobe)
    assert index.stat("device_plan") == 0
    host = {key: index.stat(key) for key in keys}
    assert dev == host
    assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    if tc_min and nq >= 8:
        assert dev["tensor_items"] > 0 or crowded
    index.close()


def test_ivf_empty_index_and_zero_query_warmup():
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(3000, 64, 4, seed=2)
    index = IVFFlatL2Index(64, 8)
    index.train(x[:500])
    D, I = index.search(q, 5, nprobe=3)          # trained, nothing added yet: all padding
    assert np.all(I == -1) and np.all(D == ivf_oracle.FLT_MAX)
    index.add(x)
    D, I = index.search(np.zeros((1, 64), np.float32), 1, nprobe=8)   # FAISSStore.load's warm-up call
    assert I.shape == (1, 1) and 0 <= I[0, 0] < 3000
    assert I[0, 0] == np.argmin((x.astype(np.float64) ** 2).sum(1))
    index.close()


def test_ivf_incremental_add_ids_and_duplicates():
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(9000, 64, 20, seed=9)
    x[100] = x[7000]            # duplicate rows land in the same list: lower id first
    q[0] = x[100]
    index = IVFFlatL2Index(64, 16)
    index.train(x[:1000])
    index.add(x[:3000])
    _check(index, q, 5, 4)      # forces a rebuild; the next adds are merged into sorted storage
    index.add(x[3000:6000])
    index.add_with_ids(x[6000:], np.arange(6000, 9000, dtype=np.int64) + 1_000_000_000_000)
    assert index.ntotal == 9000
    D, I = _check(index, q, 5, 4)
    assert list(I[0, :2]) == [100, 7000 + 1_000_000_000_000] and D[0, 0] <= 1e-3 and D[0, 1] <= 1e-3
    sizes, codes, ids = index.export_lists()
    assert sizes.sum() == 9000 and len(np.unique(ids)) == 9000
    # every list keeps add order
    off = np.concatenate([[0], np.cumsum(sizes)])
    for l in range(16):
        assert np.all(np.diff(ids[off[l]:off[l + 1]]) > 0)
    index.close()


def test_ivf_assignment_is_nearest_centroid():
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, _ = _data(20000, 128, 1, seed=3)
    index = IVFFlatL2Index(128, 100)
    index.train(x[:5000], niter=5)
    cent = index.centroids()
    assert np.isfinite(cent).all()
    a = index.assign(x)
    d = ivf_oracle.sq_l2(x, cent)
    best = d.min(axis=1)
    got = d[np.arange(len(x)), a]
    assert np.all(got <= best + 1e-3 * np.maximum(1.0, best))
    assert (a == np.argmin(d, axis=1)).mean() > 0.999
    # k-means did its job: lower quantisation error than the initial (random-row) centroids
    index2 = IVFFlatL2Index(128, 100)
    index2.train(x[:5000], niter=0)
    d0 = ivf_oracle.sq_l2(x, index2.centroids()).min(axis=1)
    assert best.mean() < d0.mean()
    index.add(x)
    assert (index.list_sizes() > 0).all()
    with pytest.raises(Exception, match="at least as large as number of clusters"):
        IVFFlatL2Index(128, 100).train(x[:50])
    with pytest.raises(Exception, match="not trained"):
        IVFFlatL2Index(128, 100).add(x[:5])
    index.close()
    index2.close()


def test_store_loads_faiss_ivfflat_file(tmp_path):
    """The reference's default index type (scripts/create_test_docs.py:83-104) loads through the
    store and is searched with settings.faiss_nprobe (faiss_store.py:84-93)."""
    from rag_inference_pipeline_b200 import B200FlatIPStore, index_io

    d, nlist, n = 768, 32, 12000
    x, q = _data(n, d, 32, seed=11)
    rng = np.random.default_rng(1)
    cent = x[rng.choice(n, nlist, replace=False)].copy()
    assign = ivf_oracle.assign_nearest(x, cent)
    sizes, codes, ids = ivf_oracle.build_lists(x, np.arange(n, dtype=np.int64), assign, nlist)
    path = tmp_path / "faiss_index.bin"
    index_io.write_faiss_ivfflat(str(path), cent, sizes, codes, ids, nprobe=64)
    settings = types.SimpleNamespace(faiss_index_path=str(path), faiss_dim=d, faiss_use_mmap=False,
                                     faiss_nprobe=8)
    store = B200FlatIPStore(settings)
    store.load()
    assert store.is_loaded and store.index_size == n
    D, I = store.search(q, 10)
    assert D.dtype == np.float32 and I.dtype == np.int64 and D.shape == (32, 10)
    Dr, Ir, _, D64 = ivf_oracle.ivf_search(q, cent, sizes, codes, ids, 11, 8)
    ok = _stable_probe_rows(q, cent, 8)
    res = ivf_oracle.compare_ivf(D[ok], I[ok], Dr[ok, :10], Ir[ok, :10], D64[ok])
    assert res["dist_bad"] == 0 and res["id_mismatch_outside_ties"] == 0, res
    with pytest.raises(ValueError, match="Embedding dimension mismatch"):
        store.search(q[:, :64], 10)
    import torch
    Dt, It = store.search_device(torch.from_numpy(q).cuda(), 10)
    assert np.array_equal(It.cpu().numpy(), I) and np.array_equal(Dt.cpu().numpy(), D)
    store.unload()
    assert not store.is_loaded


def test_ivf_coarse_quantizer_on_tensor_cores_matches_ffma():
    """The coarse quantizer on the tensor cores -- fp32 values split into three bf16 terms, the six
    significant products laid side by side along K, K1's dense mode, fp32 accumulation -- ranks the
    centroids like the fp32 FFMA GEMM: identical search results wherever the nprobe-th /
    (nprobe+1)-th centroid distances are not a near tie, on normalised rows and on rows of norm ~45."""
    from rag_inference_pipeline_b200 import IVFFlatL2Index

    x, q = _data(60000, 128, 300, seed=3, clustered=True)
    xn = (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)
    qn = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    for rows, queries, scale in ((xn, qn, 0.1), (x, q, 1.0)):
        index = IVFFlatL2Index(128, 512)
        index.train(rows[:20000])
        index.add(rows)
        index.set_option("coarse_tensor", 2)                          # every batch (auto: >= 512 queries)
        D1, I1 = _check(index, queries, 10, 32, scale=scale)
        assert index.stat("coarse_tensor") == 1
        index.set_option("coarse_tensor", 0)
        D0, I0 = _check(index, queries, 10, 32, scale=scale)
        assert index.stat("coarse_tensor") == 0
        ok = _stable_probe_rows(queries, index.centroids(), 32, scale)
        assert np.array_equal(I0[ok], I1[ok]) and np.array_equal(D0[ok], D1[ok])
        index.set_option("coarse_tensor", 2)
        D2, I2 = index.search(queries[:1], 10, nprobe=32)           # single query
        assert index.stat("coarse_tensor") == 1
        index.set_option("coarse_tensor", 1)
        index.search(queries, 10, nprobe=32)
        assert index.stat("coarse_tensor") == 0                      # auto: 300 queries stay on FFMA
        if ok[0]:
            assert np.array_equal(I2[0], I1[0])
        index.close()
