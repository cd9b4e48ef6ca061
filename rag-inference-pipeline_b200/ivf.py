"""`IVFFlatL2Index`: faiss.IndexIVFFlat-like wrapper over the C ABI in include/ivfflat_b200.h.

Mirrors the slice of the faiss surface the reference touches for its default index
(``/root/reference/scripts/create_test_docs.py:83-104``: ``IndexIVFFlat(IndexFlatL2(d), d, nlist,
METRIC_L2)``, ``train``, ``add``, ``nprobe``; ``/root/reference/src/pipeline/components/
faiss_store.py:73-93,152``: ``ntotal``, ``nprobe``, ``nlist``, ``search``).  All arithmetic happens
in libflatip_b200.so (csrc/ivf.cuh, csrc/ivf_api.cuh); there is no CPU fallback.
"""

from __future__ import annotations

import ctypes

import numpy as np

from . import _native


class IVFFlatL2Index:
    metric = "l2"

    def __init__(self, dim: int, nlist: int, device: int = 0, *, nprobe: int = 1):
        self._lib = _native.load_library()
        handle = ctypes.c_void_p()
        _native.check(self._lib.fivf_create(int(dim), int(nlist), int(device), ctypes.byref(handle)))
        self._h = handle
        self.d = int(dim)
        self.nlist = int(nlist)
        self.device = int(device)
        self.nprobe = int(nprobe)   # faiss default is 1; FAISSStore.load sets settings.faiss_nprobe
        self.dtype = "f32"          # IVF-Flat codes are raw fp32 vectors

    # -- lifecycle ---------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fivf_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass

    def _handle(self):
        if not self._h:
            raise RuntimeError("index has been closed")
        return self._h

    def _rows(self, x, what="rows") -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.d:
            raise ValueError(f"expected {what} of shape (n, {self.d}), got {x.shape}")
        return x

    # -- coarse quantizer --------------------------------------------------------------------
    @property
    def is_trained(self) -> bool:
        return bool(self._lib.fivf_is_trained(self._handle()))

    def train(self, x: np.ndarray, niter: int = -1, seed: int = 1234) -> None:
        x = self._rows(x, "training rows")
        _native.check(self._lib.fivf_train(self._handle(), x.ctypes.data, x.shape[0], int(niter), int(seed)))

    def set_centroids(self, centroids: np.ndarray) -> None:
        c = np.ascontiguousarray(centroids, dtype=np.float32)
        if c.shape != (self.nlist, self.d):
            raise ValueError(f"expected centroids of shape ({self.nlist}, {self.d}), got {c.shape}")
        _native.check(self._lib.fivf_set_centroids(self._handle(), c.ctypes.data))

    def centroids(self) -> np.ndarray:
        out = np.empty((self.nlist, self.d), dtype=np.float32)
        _native.check(self._lib.fivf_get_centroids(self._handle(), out.ctypes.data))
        return out

    def assign(self, x: np.ndarray) -> np.ndarray:
        x = self._rows(x)
        out = np.empty(x.shape[0], dtype=np.int32)
        _native.check(self._lib.fivf_assign(self._handle(), x.ctypes.data, x.shape[0], out.ctypes.data))
        return out

    # -- inverted lists ----------------------------------------------------------------------
    @property
    def ntotal(self) -> int:
        return int(self._lib.fivf_ntotal(self._handle()))

    def add(self, x: np.ndarray) -> None:
        x = self._rows(x)
        _native.check(self._lib.fivf_add(self._handle(), x.ctypes.data, x.shape[0], None))

    def add_with_ids(self, x: np.ndarray, ids: np.ndarray) -> None:
        x = self._rows(x)
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        if ids.shape != (x.shape[0],):
            raise ValueError("ids must have one entry per row")
        _native.check(self._lib.fivf_add(self._handle(), x.ctypes.data, x.shape[0], ids.ctypes.data))

    def add_preassigned(self, x: np.ndarray, ids: np.ndarray, list_numbers: np.ndarray) -> None:
        """Append rows to the given lists (how an index file's lists are imported unchanged)."""
        x = self._rows(x)
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        ln = np.ascontiguousarray(list_numbers, dtype=np.int32)
        if ids.shape != (x.shape[0],) or ln.shape != (x.shape[0],):
            raise ValueError("ids and list_numbers must have one entry per row")
        _native.check(self._lib.fivf_add_preassigned(self._handle(), x.ctypes.data, x.shape[0],
                                                     ids.ctypes.data, ln.ctypes.data))

    def list_sizes(self) -> np.ndarray:
        out = np.empty(self.nlist, dtype=np.int64)
        _native.check(self._lib.fivf_list_sizes(self._handle(), out.ctypes.data))
        return out

    def export_lists(self):
        """(sizes [nlist], codes [ntotal, d] fp32, ids [ntotal]) concatenated in list order."""
        sizes = self.list_sizes()
        n = int(sizes.sum())
        codes = np.empty((n, self.d), dtype=np.float32)
        ids = np.empty(n, dtype=np.int64)
        _native.check(self._lib.fivf_export_lists(self._handle(), codes.ctypes.data, ids.ctypes.data))
        return sizes, codes, ids

    # -- search ------------------------------------------------------------------------------
    def search(self, queries: np.ndarray, k: int, nprobe: int | None = None):
        q = self._rows(queries, "queries")
        nq = q.shape[0]
        d = np.empty((nq, int(k)), dtype=np.float32)
        i = np.empty((nq, int(k)), dtype=np.int64)
        _native.check(self._lib.fivf_search(self._handle(), q.ctypes.data, nq, int(k),
                                            int(self.nprobe if nprobe is None else nprobe),
                                            d.ctypes.data, i.ctypes.data))
        return d, i

    def search_device(self, q_ptr: int, nq: int, k: int, d_ptr: int, i_ptr: int, stream: int = 0,
                      nprobe: int | None = None) -> None:
        """Device pointers on this index's GPU (fp32 queries, fp32 distances, int64 ids).  The search
        runs on the index's own stream and blocks; the caller must have finished producing the
        queries (`stream` is accepted for signature parity with FlatIPIndex and ignored)."""
        _native.check(self._lib.fivf_search_device(self._handle(), q_ptr, int(nq), int(k),
                                                   int(self.nprobe if nprobe is None else nprobe), d_ptr, i_ptr))

    def set_option(self, key: str, value: int) -> None:
        _native.check(self._lib.fivf_set_option(self._handle(), key.encode(), int(value)))

    def stat(self, key: str) -> int:
        return int(self._lib.fivf_get_stat(self._handle(), key.encode()))
