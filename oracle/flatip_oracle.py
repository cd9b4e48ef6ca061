"""CPU ORACLE for the flat inner-product search path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import this module.  The product (``rag-inference-pipeline_b200``) never
does: it fails loudly when the CUDA library is missing.

PARITY UNPINNED: the arithmetic of the reference's hot path lives in the third-party wheel
``faiss-cpu==1.13.1`` (``/root/reference/uv.lock:959-960``), which is neither vendored under
``/root/reference`` nor installable in this image (no wheel, no network), and no reference test
pins a search result (``/root/reference/tests/test_components.py:136-157`` only checks that a
mocked index's return value is passed through; ``tests/test_retrieval_service.py:6-10`` says the
real search test was dropped).  This file therefore restates the *published* semantics of
``faiss.IndexFlatIP.search`` -- exact ``S[i, j] = <q_i, x_j>`` over every stored row, per-query
k largest, best first, ``-1`` / lowest-float padding when fewer than k rows exist -- anchored on
the reference's own call sites:

* ``/root/reference/src/pipeline/components/faiss_store.py:113-158`` (guards, fp32 cast,
  ``self._index.search(embeddings, k)`` at line 152, ``(distances, indices)`` return order);
* ``/root/reference/src/pipeline/services/retrieval/api.py:376-425`` (rows consumed via
  ``.tolist()``, ids used as document ids);
* ``/root/reference/tests/test_retrieval_service.py:202-212,260-261`` (the only committed
  *inputs*: ``default_rng(42).random((10, 768))`` corpus, ``default_rng(42).random((2, 768))``
  queries) -- regenerated in ``tests/golden/make_golden.py``.

Tie-break: faiss's order among equal scores depends on heap / reservoir internals; BASELINE.json's
north_star fixes it as (score descending, id ascending) and that is what is implemented here.
"""

from __future__ import annotations

import numpy as np

LOWEST_F32 = np.float32(np.finfo(np.float32).min)  # faiss fills unfilled slots with -FLT_MAX


def _topk_merge(best_s, best_i, s_blk, id0, k):
    """Merge a score block [B, nb] (ids id0..id0+nb) into the running best [B, <=k]."""
    B, nb = s_blk.shape
    ids = np.broadcast_to(np.arange(id0, id0 + nb, dtype=np.int64)[None, :], (B, nb))
    s_all = np.concatenate([best_s, s_blk], axis=1)
    i_all = np.concatenate([best_i, ids], axis=1)
    # score descending, id ascending: lexsort's last key is the primary one.
    order = np.lexsort((i_all, -s_all), axis=1)[:, :k]
    return np.take_along_axis(s_all, order, axis=1), np.take_along_axis(i_all, order, axis=1)


def flat_ip_search(queries: np.ndarray, corpus: np.ndarray, k: int, *, id_base: int = 0,
                   accum: str = "f64", block: int = 65536):
    """Exact flat inner-product top-k.

    queries [B, d] float; corpus [N, d] float (already rounded to the storage dtype and up-cast,
    see ``synthetic.round_to_storage``); returns (distances float32 [B, k], ids int64 [B, k]).

    ``accum``: "f64" (truth; products of fp32 values summed in double) or "f32" (what
    faiss-cpu/sgemm would give up to summation order).  Scores are reported as fp32 either way.
    """
    if queries.ndim != 2 or corpus.ndim != 2 or queries.shape[1] != corpus.shape[1]:
        raise ValueError("shape mismatch")
    if k <= 0:
        raise ValueError("k must be positive")
    acc = np.float64 if accum == "f64" else np.float32
    q = np.ascontiguousarray(queries, dtype=np.float32).astype(acc)
    B = q.shape[0]
    N = corpus.shape[0]
    best_s = np.empty((B, 0), dtype=acc)
    best_i = np.empty((B, 0), dtype=np.int64)
    for j0 in range(0, N, block):
        xb = np.ascontiguousarray(corpus[j0:j0 + block], dtype=np.float32).astype(acc)
        s = q @ xb.T
        best_s, best_i = _topk_merge(best_s, best_i, s, j0, k)
    have = best_s.shape[1]
    D = np.full((B, k), LOWEST_F32, dtype=np.float32)
    I = np.full((B, k), -1, dtype=np.int64)
    D[:, :have] = best_s.astype(np.float32)
    I[:, :have] = best_i + id_base
    return D, I


def exact_scores_f64(queries: np.ndarray, corpus_rows: np.ndarray) -> np.ndarray:
    """<q_i, x_i'> in double for explicit (query, row) pairs: queries [B,d], corpus_rows [B,k,d]."""
    return np.einsum("bd,bkd->bk", queries.astype(np.float64), corpus_rows.astype(np.float64))


def merge_sorted_lists(dists: np.ndarray, ids: np.ndarray, k: int):
    """Oracle for the cross-shard merge: dists/ids [g, B, k_in] (each list best-first, padded with
    id -1) -> global (dists [B,k], ids [B,k]) under (score desc, id asc); padding sorts last."""
    g, B, kin = dists.shape
    s = np.transpose(dists, (1, 0, 2)).reshape(B, g * kin).astype(np.float64)
    i = np.transpose(ids, (1, 0, 2)).reshape(B, g * kin).astype(np.int64)
    pad = i < 0
    s_key = np.where(pad, -np.inf, s)
    i_key = np.where(pad, np.iinfo(np.int64).max, i)
    order = np.lexsort((i_key, -s_key), axis=1)[:, :k]
    D = np.take_along_axis(np.transpose(dists, (1, 0, 2)).reshape(B, g * kin), order, axis=1)
    I = np.take_along_axis(i, order, axis=1)
    if D.shape[1] < k:
        padn = k - D.shape[1]
        D = np.concatenate([D, np.full((B, padn), LOWEST_F32, np.float32)], axis=1)
        I = np.concatenate([I, np.full((B, padn), -1, np.int64)], axis=1)
    return D.astype(np.float32), I


# --------------------------------------------------------------------------------------------
# Acceptance rule (BASELINE.json north_star): scores within rtol (1e-3 for fp16/bf16 storage,
# 1e-5 for fp32); ids bit-exact outside near-tie windows.
# --------------------------------------------------------------------------------------------

SCORE_RTOL = {"f32": 1e-5, "f16": 1e-3, "bf16": 1e-3}


def compare_with_oracle(D_gpu, I_gpu, D_ref, I_ref, *, rtol: float, kplus1_ref=None):
    """Compare a result with the fp64 oracle's.

    Returns a dict with ``score_ok`` (all |d_gpu - d_ref| <= rtol * max(1, |d_ref|)),
    ``id_mismatch_outside_window`` (count of positions whose ids differ although the oracle score
    at that position is separated from both neighbours -- and from the (k+1)-th score when given --
    by more than the tolerance) and ``id_mismatch_total``.
    """
    D_gpu = np.asarray(D_gpu, dtype=np.float64)
    D_ref64 = np.asarray(D_ref, dtype=np.float64)
    valid = np.asarray(I_ref) >= 0
    tol = rtol * np.maximum(1.0, np.abs(D_ref64))
    score_err = np.abs(D_gpu - D_ref64)
    score_ok = bool(np.all(score_err[valid] <= tol[valid])) and bool(
        np.array_equal(np.asarray(I_gpu) < 0, ~valid))
    B, k = D_ref64.shape
    gap_prev = np.full((B, k), np.inf)
    gap_next = np.full((B, k), np.inf)
    gap_prev[:, 1:] = np.abs(D_ref64[:, 1:] - D_ref64[:, :-1])
    gap_next[:, :-1] = gap_prev[:, 1:]
    if kplus1_ref is not None:
        gap_next[:, -1] = np.abs(D_ref64[:, -1] - np.asarray(kplus1_ref, dtype=np.float64))
    # A position is "in a near-tie window" if swapping with a neighbour is within tolerance
    # on either side (2*tol: both scores carry an error of up to tol).
    in_window = (gap_prev <= 2 * tol) | (gap_next <= 2 * tol)
    mism = (np.asarray(I_gpu) != np.asarray(I_ref)) & valid
    return {
        "score_ok": score_ok,
        "max_score_err": float(score_err[valid].max()) if valid.any() else 0.0,
        "id_mismatch_total": int(mism.sum()),
        "id_mismatch_outside_window": int((mism & ~in_window).sum()),
        "window_positions": int((in_window & valid).sum()),
    }
