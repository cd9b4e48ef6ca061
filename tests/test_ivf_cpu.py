"""CPU tests of the IVF-Flat operator's host side: the oracle against brute force, the faiss
IndexIVFFlat file layout round trip, and the C-ABI symbols declared in include/ivfflat_b200.h."""
import os
import re

import numpy as np
import pytest

import ivf_oracle
from conftest import ROOT, gpu_available


def _index(n=3000, d=32, nlist=12, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32)
    cent = x[rng.choice(n, nlist, replace=False)].copy()
    assign = ivf_oracle.assign_nearest(x, cent)
    sizes, codes, ids = ivf_oracle.build_lists(x, np.arange(n, dtype=np.int64), assign, nlist)
    q = rng.standard_normal((7, d)).astype(np.float32)
    return x, cent, sizes, codes, ids, q


def test_oracle_exhaustive_probing_is_brute_force_l2():
    x, cent, sizes, codes, ids, q = _index()
    D, I, P, D64 = ivf_oracle.ivf_search(q, cent, sizes, codes, ids, 10, nprobe=len(cent))
    d = ((q[:, None, :].astype(np.float64) - x[None].astype(np.float64)) ** 2).sum(-1)
    ref = np.argsort(d, axis=1, kind="stable")[:, :10]
    assert np.array_equal(I, ref)
    assert np.allclose(D, np.take_along_axis(d, ref, 1), rtol=1e-6)
    assert np.array_equal(ids[P], I)


def test_oracle_probes_restrict_the_scan_and_pad():
    x, cent, sizes, codes, ids, q = _index(n=200, nlist=40)
    D, I, _, _ = ivf_oracle.ivf_search(q, cent, sizes, codes, ids, 50, nprobe=2)
    probes = ivf_oracle.coarse_probes(q, cent, 2)
    off = np.concatenate([[0], np.cumsum(sizes)])
    for qi in range(len(q)):
        allowed = set()
        for l in probes[qi]:
            allowed.update(ids[off[l]:off[l + 1]].tolist())
        got = I[qi][I[qi] >= 0]
        assert set(got.tolist()) == allowed or len(got) == 50
        assert np.all(D[qi][I[qi] < 0] == ivf_oracle.FLT_MAX)
        assert np.all(np.diff(D[qi][I[qi] >= 0]) >= 0)
    # the nearest centroid of a centroid is itself
    assert np.array_equal(ivf_oracle.coarse_probes(cent, cent, 1)[:, 0], np.arange(len(cent)))


def test_faiss_ivfflat_file_round_trip(tmp_path):
    from rag_inference_pipeline_b200 import index_io

    _, cent, sizes, codes, ids, _ = _index()
    for s in (sizes, np.concatenate([[int(sizes.sum())], np.zeros(len(sizes) - 1, np.int64)])):
        path = str(tmp_path / "ivf.bin")
        index_io.write_faiss_ivfflat(path, cent, s, codes, ids, nprobe=64)
        assert index_io.is_faiss_ivfflat(path)
        f = index_io.read_faiss_ivfflat(path)
        assert (f.dim, f.ntotal, f.nlist, f.nprobe, f.metric) == (32, 3000, 12, 64, "l2")
        assert np.array_equal(f.centroids, cent) and np.array_equal(f.sizes, s)
        assert np.array_equal(f.codes, codes) and np.array_equal(f.ids, ids)
    with open(path, "r+b") as fh:   # truncated file fails loudly
        fh.truncate(os.path.getsize(path) - 100)
    with pytest.raises(ValueError, match="shorter"):
        index_io.read_faiss_ivfflat(path)
    flat = str(tmp_path / "flat.bin")
    index_io.write_faiss_flat(flat, codes)
    assert not index_io.is_faiss_ivfflat(flat)


def test_library_exports_every_declared_ivf_symbol(native_lib):
    from rag_inference_pipeline_b200 import _native

    text = open(os.path.join(ROOT, "include", "ivfflat_b200.h")).read()
    declared = set(re.findall(r"\b(fivf_[a-z0-9_]+)\s*\(", text))
    assert declared == set(_native.IVF_EXPORTED_SYMBOLS), declared ^ set(_native.IVF_EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(native_lib, name), name


@pytest.mark.skipif(gpu_available(), reason="checks the no-GPU failure mode")
def test_ivf_has_no_cpu_fallback(native_lib):
    from rag_inference_pipeline_b200 import IVFFlatL2Index
    from rag_inference_pipeline_b200._native import FIP_ERR_NODEVICE, FlatIPError

    with pytest.raises(FlatIPError) as ei:
        IVFFlatL2Index(64, 8)
    assert ei.value.code == FIP_ERR_NODEVICE


def test_c_restatement_of_faiss_ivf_matches_numpy_oracle():
    """oracle/faiss_ivf_restatement.c (the timed CPU baseline: OpenMP over queries, per-query
    coarse probe + list scan + k-max-heap) against the fp64 numpy restatement."""
    import c_oracle

    _, cent, sizes, codes, ids, q = _index(n=20000, d=64, nlist=50, seed=3)
    for k, nprobe in ((10, 5), (300, 2), (5, 50), (1, 1)):
        D, I, _, D64 = ivf_oracle.ivf_search(q, cent, sizes, codes, ids, k + 1, nprobe)
        Dc, Ic = c_oracle.ivf_search_l2(q, cent, sizes, codes, ids, k, nprobe)
        res = ivf_oracle.compare_ivf(Dc, Ic, D[:, :k], I[:, :k], D64)
        assert res["dist_bad"] == 0 and res["pad_bad"] == 0 and res["id_mismatch_outside_ties"] == 0, res
    D1, I1 = c_oracle.ivf_search_l2(q, cent, sizes, codes, ids, 10, 5, nthreads=1)
    D2, I2 = c_oracle.ivf_search_l2(q, cent, sizes, codes, ids, 10, 5, nthreads=4)
    assert np.array_equal(I1, I2) and np.array_equal(D1, D2)


@pytest.mark.parametrize("nlist,nq,nprobe,k,cap,safe,tc_min", [
    (4096, 256, 64, 84, 4096, 0, 8),     # reference shape: tensor items + FFMA2 tiles
    (4096, 32, 64, 84, 4096, 0, 8),      # few queries per list: FFMA2 tiles only
    (512, 300, 32, 84, 4096, 0, 0),      # tensor path off
    (64, 200, 16, 84, 4096, 1, 8),       # safe mode: phases of cap - k rows, partial lists
    (64, 100, 16, 300, 16384, 1, 0),
    (16, 500, 16, 84, 4096, 0, 8),       # every list probed by every query: several 128-query groups
])
def test_host_plan_covers_every_probed_list_exactly_once(native_lib, nlist, nq, nprobe, k, cap, safe, tc_min):
    """The IVF host planner (csrc/ivf_api.cuh build_plan) through its no-GPU test hook: every
    (query, probed list) pair must be scored over exactly the rows of that list, once, whatever mix
    of dense prefixes, skip-aware FFMA2 tiles, tensor-core items and safe-mode phases is planned."""
    import ctypes

    f = native_lib.fivf_debug_plan
    f.restype = ctypes.c_int
    f.argtypes = [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p] + [ctypes.c_int32] * 7 + [ctypes.c_void_p]
    rng = np.random.default_rng(nlist + nq)
    sizes = rng.poisson(900, nlist).astype(np.int64)
    sizes[0] = 0                      # an empty list
    sizes[1] = 30000                  # a list longer than the candidate capacity
    sizes[2] = 5
    probes = np.stack([rng.choice(nlist, nprobe, replace=False) for _ in range(nq)]).astype(np.int64)
    probes[0, -1] = -1                # fewer than nprobe lists for one query
    others = np.setdiff1d(np.arange(nlist), [1])
    probes[1] = np.concatenate([[1], rng.choice(others, nprobe - 1, replace=False)])   # starts in the long list
    out = np.zeros(8, np.int64)
    rc = f(nlist, sizes.ctypes.data, probes.ctypes.data, nq, nprobe, k, cap, safe, tc_min, 1, out.ctypes.data)
    assert rc == 0
    phases, tiles, items, rows, pairs, _, bad, qrows = out.tolist()
    assert bad == 0
    want_pairs = int(sum(sizes[l] for l in probes.ravel() if l >= 0))
    assert pairs >= want_pairs          # (tensor items count the skipped head rows too)
    assert phases >= 2 or rows == 0
    if tc_min == 0 or safe:
        assert items == 0 and qrows == 0
    if (nlist, tc_min, safe) == (16, 8, 0):
        assert items > 0 and qrows % 32 == 0
