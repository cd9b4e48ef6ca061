"""ctypes loader for oracle/faiss_flat_restatement.c -- TEST INFRASTRUCTURE ONLY.

Used by tests, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs.  Never imported by the product package.
"""

from __future__ import annotations

import ctypes
import glob
import os
import subprocess

import numpy as np

# Same policy the reference sets before faiss spins up its OpenMP pool
# (/root/reference/src/pipeline/runtime.py:51); without it libgomp's spinning workers starve the
# BLAS threads and the sgemm path runs ~30x slower.
os.environ.setdefault("OMP_WAIT_POLICY", "passive")

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libflatip_oracle.so")
_lib = None
_sgemm_ptr = None
_blas_handle = None


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("faiss_flat_restatement.c", "faiss_ivf_restatement.c")]
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < max(map(os.path.getmtime, srcs)):
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True)
    return _LIB_PATH


def _find_openblas_sgemm():
    """Fortran-ABI ILP64 sgemm from the OpenBLAS that numpy bundles (scipy_sgemm_64_)."""
    global _sgemm_ptr, _blas_handle
    if _sgemm_ptr is not None:
        return _sgemm_ptr
    libdir = os.path.join(os.path.dirname(np.__file__), "..", "numpy.libs")
    for path in sorted(glob.glob(os.path.join(libdir, "libscipy_openblas64_*.so"))):
        try:
            h = ctypes.CDLL(path)
            fn = getattr(h, "scipy_sgemm_64_")
        except (OSError, AttributeError):
            continue
        _blas_handle = h
        _sgemm_ptr = ctypes.cast(fn, ctypes.c_void_p)
        return _sgemm_ptr
    _sgemm_ptr = ctypes.c_void_p(None)
    return _sgemm_ptr


def load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.oracle_knn_inner_product.restype = ctypes.c_int
        _lib.oracle_knn_inner_product.argtypes = [
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
            ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_int]
        _lib.oracle_max_threads.restype = ctypes.c_int
        _lib.oracle_ivf_search_l2.restype = ctypes.c_int
        _lib.oracle_ivf_search_l2.argtypes = [
            ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        _lib.oracle_synthetic_rows.restype = ctypes.c_int
        _lib.oracle_synthetic_rows.argtypes = [
            ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_uint64,
            ctypes.c_uint32, ctypes.c_int]
    return _lib


def has_blas() -> bool:
    return bool(_find_openblas_sgemm().value)


def max_threads() -> int:
    return int(load().oracle_max_threads())


def set_threads(n: int) -> int:
    """Pin BOTH thread pools of the port -- its own OpenMP loops and the OpenBLAS behind sgemm --
    to `n` threads, whatever OMP_NUM_THREADS says (torch.distributed.run exports
    OMP_NUM_THREADS=1 to every rank, which silently single-threaded the reference arm at N >= 2).
    The reference does the same with ``faiss.omp_set_num_threads(settings.faiss_threads)``
    (/root/reference/src/pipeline/runtime.py:76).  Returns the OpenBLAS thread count in effect."""
    n = max(1, int(n))
    lib = load()
    lib.oracle_set_threads.argtypes = [ctypes.c_int]
    lib.oracle_set_threads(n)
    _find_openblas_sgemm()
    got = n
    if _blas_handle is not None:
        for name in ("scipy_openblas_set_num_threads64_", "scipy_openblas_set_num_threads_64_"):
            fn = getattr(_blas_handle, name, None)
            if fn is not None:
                fn.argtypes = [ctypes.c_int]
                fn(n)
                break
        for name in ("scipy_openblas_get_num_threads64_", "scipy_openblas_get_num_threads_64_"):
            fn = getattr(_blas_handle, name, None)
            if fn is not None:
                fn.restype = ctypes.c_int
                got = int(fn())
                break
    return got


def knn_inner_product(queries: np.ndarray, corpus: np.ndarray, k: int, *, id_base: int = 0,
                      nthreads: int = 0, use_blas: bool = True):
    """faiss-style CPU flat-IP search (fp32).  Returns (D float32 [B,k], I int64 [B,k])."""
    lib = load()
    q = np.ascontiguousarray(queries, dtype=np.float32)
    x = np.ascontiguousarray(corpus, dtype=np.float32)
    if q.ndim != 2 or x.ndim != 2 or q.shape[1] != x.shape[1]:
        raise ValueError("shape mismatch")
    B, d = q.shape
    D = np.empty((B, k), dtype=np.float32)
    I = np.empty((B, k), dtype=np.int64)
    sg = _find_openblas_sgemm() if use_blas else ctypes.c_void_p(None)
    rc = lib.oracle_knn_inner_product(q.ctypes.data, x.ctypes.data, d, B, x.shape[0], k, id_base,
                                      D.ctypes.data, I.ctypes.data, sg, int(nthreads))
    if rc != 0:
        raise RuntimeError(f"oracle_knn_inner_product failed: {rc}")
    return D, I


_STORAGE = {"f32": 0, "f16": 1, "bf16": 2}


def synthetic_corpus(row0: int, nrows: int, dim: int, dtype: str = "f32", seed: int = 1234,
                     stream: int = 0) -> np.ndarray:
    """Fast C restatement of ``synthetic.synthetic_corpus`` (same bits), for big test corpora."""
    out = np.empty((nrows, dim), dtype=np.float32)
    rc = load().oracle_synthetic_rows(out.ctypes.data, row0, nrows, dim, seed, stream,
                                      _STORAGE[dtype])
    if rc != 0:
        raise RuntimeError("oracle_synthetic_rows failed")
    return out


def ivf_search_l2(queries: np.ndarray, centroids: np.ndarray, sizes: np.ndarray, codes: np.ndarray,
                  ids: np.ndarray, k: int, nprobe: int, *, nthreads: int = 0):
    """faiss-style CPU IVF-Flat (L2) search (oracle/faiss_ivf_restatement.c): OpenMP over queries,
    per-query coarse probe + list scan + k-max-heap.  Returns (D float32 [B,k], I int64 [B,k])."""
    lib = load()
    q = np.ascontiguousarray(queries, dtype=np.float32)
    c = np.ascontiguousarray(centroids, dtype=np.float32)
    x = np.ascontiguousarray(codes, dtype=np.float32)
    i = np.ascontiguousarray(ids, dtype=np.int64)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    B, d = q.shape
    if c.shape[1] != d or x.shape != (int(off[-1]), d) or i.shape != (int(off[-1]),):
        raise ValueError("shape mismatch")
    D = np.empty((B, k), dtype=np.float32)
    I = np.empty((B, k), dtype=np.int64)
    rc = lib.oracle_ivf_search_l2(q.ctypes.data, B, d, c.ctypes.data, c.shape[0], off.ctypes.data,
                                  x.ctypes.data, i.ctypes.data, k, nprobe, D.ctypes.data,
                                  I.ctypes.data, int(nthreads))
    if rc != 0:
        raise RuntimeError(f"oracle_ivf_search_l2 failed: {rc}")
    return D, I
