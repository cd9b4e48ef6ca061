"""TEST INFRASTRUCTURE ONLY -- CPU restatement of faiss ``IndexIVFFlat`` (METRIC_L2) search.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module; the
product path (rag-inference-pipeline_b200/) never does.

PARITY UNPINNED: faiss-cpu 1.13.1 (/root/reference/uv.lock:959-960) is a third-party wheel that is
neither vendored in /root/reference nor installable here, and no reference test pins an IVF search
result.  This restates the published algorithm behind the reference's call sites:

* build: /root/reference/scripts/create_test_docs.py:83-104 -- ``IndexIVFFlat(IndexFlatL2(d), d,
  nlist, METRIC_L2)``; ``add`` appends each row to the inverted list of its NEAREST centroid
  (``quantizer.assign``, squared L2);
* search: /root/reference/src/pipeline/components/faiss_store.py:84-93,152 -- ``index.nprobe = n``;
  ``index.search(x, k)``: ``IndexIVF::search`` = ``quantizer.search(x, nprobe)`` then
  ``search_preassigned``: scan the probed lists (``IVFFlatScannerL2``: ``fvec_L2sqr(q, code)``)
  into a k-max-heap per query; results ascending by squared L2 distance, missing slots id -1 /
  distance FLT_MAX (the CMax heap's neutral element).

Arithmetic is fp64 on the fp32 inputs; ties are broken (distance ascending, list position
ascending) -- the rule the CUDA operator fixes (faiss's heap order among exact ties is unspecified).
"""

from __future__ import annotations

import numpy as np

FLT_MAX = np.float32(3.4028234663852886e38)


def sq_l2(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """[na, nb] squared L2 distances in fp64, computed as sum((a - b)^2) blockwise."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    out = np.empty((a.shape[0], b.shape[0]), dtype=np.float64)
    an = (a * a).sum(1)[:, None]
    bn = (b * b).sum(1)[None, :]
    step = max(1, (1 << 24) // max(1, b.shape[0]))
    for i in range(0, a.shape[0], step):
        out[i:i + step] = an[i:i + step] + bn - 2.0 * (a[i:i + step] @ b.T)
    return np.maximum(out, 0.0)


def assign_nearest(x: np.ndarray, centroids: np.ndarray) -> np.ndarray:
    """quantizer.assign: nearest centroid (lowest number on exact ties)."""
    return np.argmin(sq_l2(x, centroids), axis=1).astype(np.int32)


def build_lists(x: np.ndarray, ids: np.ndarray, assign: np.ndarray, nlist: int):
    """Inverted lists in add order: (sizes [nlist], codes [n, d] sorted by list, ids [n])."""
    order = np.argsort(assign, kind="stable")
    sizes = np.bincount(assign, minlength=nlist).astype(np.int64)
    return sizes, np.ascontiguousarray(x[order]), np.ascontiguousarray(ids[order])


def coarse_probes(queries: np.ndarray, centroids: np.ndarray, nprobe: int) -> np.ndarray:
    """[nq, nprobe] list numbers, nearest first (distance ascending, list number ascending)."""
    d = sq_l2(queries, centroids)
    nprobe = min(nprobe, centroids.shape[0])
    idx = np.lexsort((np.broadcast_to(np.arange(d.shape[1]), d.shape), d), axis=1)
    return idx[:, :nprobe].astype(np.int64)


def ivf_search(queries: np.ndarray, centroids: np.ndarray, sizes: np.ndarray, codes: np.ndarray,
               ids: np.ndarray, k: int, nprobe: int, probes: np.ndarray | None = None):
    """(distances fp32 [nq, k], ids int64 [nq, k], positions int64 [nq, k], fp64 distances)."""
    queries = np.asarray(queries, dtype=np.float32)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    if probes is None:
        probes = coarse_probes(queries, centroids, nprobe)
    nq = queries.shape[0]
    D = np.full((nq, k), FLT_MAX, dtype=np.float32)
    I = np.full((nq, k), -1, dtype=np.int64)
    P = np.full((nq, k), -1, dtype=np.int64)
    D64 = np.full((nq, k), np.inf, dtype=np.float64)
    for q in range(nq):
        pos = np.concatenate([np.arange(off[l], off[l + 1]) for l in probes[q] if l >= 0] or
                             [np.empty(0, dtype=np.int64)])
        if pos.size == 0:
            continue
        diff = codes[pos].astype(np.float64) - queries[q].astype(np.float64)[None, :]
        dist = (diff * diff).sum(1)
        order = np.lexsort((pos, dist))[:k]
        m = order.size
        D64[q, :m] = dist[order]
        D[q, :m] = dist[order].astype(np.float32)
        P[q, :m] = pos[order]
        I[q, :m] = ids[pos[order]]
    return D, I, P, D64


def compare_ivf(D, I, Dr, Ir, D64r, rtol=1e-5, atol=1e-4):
    """Distances within tolerance; ids bit-exact outside near-tie windows (neighbouring oracle
    distances closer than the tolerance).  Returns a dict of counts."""
    tol = atol + rtol * np.abs(Dr.astype(np.float64))
    valid = Ir >= 0
    dist_bad = int((np.abs(D.astype(np.float64) - Dr.astype(np.float64)) > tol)[valid].sum())
    pad_bad = int(((I != -1) | (D != FLT_MAX))[~valid].sum())
    mism = (I != Ir) & valid
    # a mismatch is excused when the oracle's neighbours at that rank are within tolerance
    # D64r may carry extra columns (oracle run with k + 1) so the last slot sees its successor
    k = D.shape[1]
    gap_prev = np.full(D64r.shape, np.inf)
    gap_next = np.full(D64r.shape, np.inf)
    with np.errstate(invalid="ignore"):   # padding slots hold inf: inf - inf is a NaN gap, never "near"
        diff = D64r[:, 1:] - D64r[:, :-1]
    gap_prev[:, 1:] = diff
    gap_next[:, :-1] = diff
    near = (np.minimum(gap_prev, gap_next)[:, :k] <= 2 * tol)
    excused = mism & near
    return {"dist_bad": dist_bad, "pad_bad": pad_bad, "id_mismatch_total": int(mism.sum()),
            "id_mismatch_outside_ties": int((mism & ~excused).sum())}
