"""`FlatIPIndex`: thin object wrapper over the C ABI (one index = one GPU = one corpus shard).

Mirrors the slice of the ``faiss.Index`` surface the reference touches
(``/root/reference/src/pipeline/components/faiss_store.py:66-73,106,152,182``): ``d``, ``ntotal``,
``add``, ``search``, ``reconstruct_n``.  All arithmetic happens in libflatip_b200.so.
"""

from __future__ import annotations

import ctypes

import numpy as np

from . import _native


class FlatIPIndex:
    def __init__(self, dim: int, dtype: str = "f32", device: int = 0, *, id_base: int = 0,
                 capacity_rows: int = 0):
        if dtype not in _native.DTYPE_CODES:
            raise ValueError(f"unknown storage dtype {dtype!r}")
        self._lib = _native.load_library()
        cfg = _native.FipConfig(dim=int(dim), dtype=_native.DTYPE_CODES[dtype], device=int(device),
                                reserved=0, capacity_rows=int(capacity_rows), id_base=int(id_base))
        handle = ctypes.c_void_p()
        _native.check(self._lib.fip_create(ctypes.byref(cfg), ctypes.byref(handle)))
        self._h = handle
        self.d = int(dim)
        self.dtype = _native.DTYPE_NAMES[_native.DTYPE_CODES[dtype]]
        self.device = int(device)
        self.id_base = int(id_base)

    # -- lifecycle ---------------------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fip_destroy(self._h)
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

    # -- corpus ------------------------------------------------------------------------------
    @property
    def ntotal(self) -> int:
        return int(self._lib.fip_ntotal(self._handle()))

    def add(self, x: np.ndarray) -> None:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.d:
            raise ValueError(f"expected rows of shape (n, {self.d}), got {x.shape}")
        _native.check(self._lib.fip_add_host_f32(self._handle(), x.ctypes.data, x.shape[0]))

    def add_raw(self, x: np.ndarray, dtype: str) -> None:
        """Host rows in their OWN storage width: ``x`` is float32 / float16 [n, d], or uint16 [n, d]
        holding bf16 bit patterns (``dtype="bf16"``).  16-bit rows cross PCIe as 16 bits."""
        _native.check(self._lib.fip_add_host(self._handle(), *raw_rows(x, dtype, self.d)))

    def add_device(self, ptr: int, n: int, dtype: str) -> None:
        _native.check(self._lib.fip_add_device(self._handle(), ctypes.c_void_p(ptr), int(n),
                                               _native.DTYPE_CODES[dtype]))

    def add_synthetic(self, n: int, seed: int, global_row0: int = 0) -> None:
        _native.check(self._lib.fip_add_synthetic(self._handle(), int(n), int(seed),
                                                  int(global_row0)))

    def reset(self) -> None:
        _native.check(self._lib.fip_reset(self._handle()))

    def reconstruct_n(self, row0: int, n: int) -> np.ndarray:
        out = np.empty((n, self.d), dtype=np.float32)
        _native.check(self._lib.fip_reconstruct(self._handle(), int(row0), int(n), out.ctypes.data))
        return out

    # -- search ------------------------------------------------------------------------------
    def search(self, x: np.ndarray, k: int):
        """faiss-style: x float32 [n, d] -> (D float32 [n, k], I int64 [n, k])."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.d:
            raise ValueError(f"expected queries of shape (n, {self.d}), got {x.shape}")
        n = x.shape[0]
        D = np.empty((n, int(k)), dtype=np.float32)
        I = np.empty((n, int(k)), dtype=np.int64)
        _native.check(self._lib.fip_search(self._handle(), x.ctypes.data, n, int(k),
                                           D.ctypes.data, I.ctypes.data))
        return D, I

    def search_device(self, q_ptr: int, nq: int, k: int, d_ptr: int, i_ptr: int,
                      stream: int = 0) -> None:
        """Device-resident variant: raw device pointers (e.g. torch ``data_ptr()``)."""
        _native.check(self._lib.fip_search_device(
            self._handle(), ctypes.c_void_p(q_ptr), int(nq), int(k), ctypes.c_void_p(d_ptr),
            ctypes.c_void_p(i_ptr), ctypes.c_void_p(stream)))

    def search_device_async(self, q_ptr: int, nq: int, k: int, d_ptr: int, i_ptr: int,
                            stream: int = 0) -> None:
        """Enqueue only (no host round trip).  ``search_status()`` later tells whether any such
        search must be redone with the blocking call; see include/flatip_b200.h."""
        _native.check(self._lib.fip_search_device_async(
            self._handle(), ctypes.c_void_p(q_ptr), int(nq), int(k), ctypes.c_void_p(d_ptr),
            ctypes.c_void_p(i_ptr), ctypes.c_void_p(stream)))

    def search_status(self) -> int:
        """Synchronises; nonzero when some asynchronous search since the last call was flagged."""
        flagged = ctypes.c_int32(0)
        _native.check(self._lib.fip_search_status(self._handle(), ctypes.byref(flagged)))
        return int(flagged.value)

    def flags_device_ptr(self) -> int:
        return int(self._lib.fip_search_flags_device(self._handle()) or 0)

    # -- knobs / counters --------------------------------------------------------------------
    def set_option(self, key: str, value: int) -> None:
        _native.check(self._lib.fip_set_option(self._handle(), key.encode(), int(value)))

    def set_variant(self, name: str) -> None:
        self.set_option("variant", _native.VARIANTS[name])

    def stat(self, key: str) -> int:
        return int(self._lib.fip_get_stat(self._handle(), key.encode()))

    def last_variant(self) -> str:
        return _native.VARIANT_NAMES.get(self.stat("variant"), "none")


CUDA_STREAM_LEGACY = 1   # cudaStreamLegacy: explicit handle of the legacy default stream


def torch_stream_handle(device=None) -> int:
    """Handle of torch's current stream for the ``stream`` arguments of this package.  torch's
    default stream has handle 0, which the C ABI reads as "the index's own stream"; naming the
    legacy default stream explicitly keeps the search ordered with the caller's torch work."""
    import torch

    return torch.cuda.current_stream(device).cuda_stream or CUDA_STREAM_LEGACY


def raw_rows(x: np.ndarray, dtype: str, dim: int):
    """(pointer, n, dtype code) of host rows handed over in their own width."""
    code = _native.DTYPE_CODES[dtype]
    want = {0: np.float32, 1: np.float16, 2: np.uint16}[code]
    x = np.ascontiguousarray(x)
    if x.dtype != want:
        raise ValueError(f"{dtype} rows must be {np.dtype(want).name} (bf16 as uint16 bit patterns), got {x.dtype}")
    if x.ndim != 2 or x.shape[1] != dim:
        raise ValueError(f"expected rows of shape (n, {dim}), got {x.shape}")
    return x.ctypes.data, x.shape[0], code


STATIC_MAX_QUERIES = 64   # fip_search_rows_static: queries per call its fixed workspace holds


def search_rows_static(device: int, rows_ptr: int, n_rows: int, dim: int, dtype: str, q_ptr: int, nq: int,
                       k: int, d_ptr: int, i_ptr: int, stream: int = 0, id_base: int = 0) -> None:
    """Exact search over caller-owned device rows without a single allocator call
    (``fip_search_rows_static``): for hosts that share the GPU with a server that has claimed all
    of HBM, or whose own allocator forbids raw ``cudaMalloc``.  Raw device pointers in, results
    written to ``d_ptr`` / ``i_ptr``; at most ``STATIC_MAX_QUERIES`` queries and k <= 2048 per call
    (``FlatIPError`` otherwise).  Blocking; calls on one GPU take turns."""
    lib = _native.load_library()
    P = ctypes.c_void_p
    _native.check(lib.fip_search_rows_static(
        int(device), P(rows_ptr), int(n_rows), int(dim), _native.DTYPE_CODES[dtype], int(id_base),
        P(q_ptr), int(nq), int(k), P(d_ptr), P(i_ptr), P(stream)))


def merge_device(device: int, in_d_ptr: int, in_i_ptr: int, g: int, nq: int, k_in: int, k_out: int,
                 out_d_ptr: int, out_i_ptr: int, stream: int = 0) -> None:
    """K3 cross-shard merge on raw device pointers ([g, nq, k_in] -> [nq, k_out])."""
    lib = _native.load_library()
    P = ctypes.c_void_p
    _native.check(lib.fip_merge_device(int(device), P(in_d_ptr), P(in_i_ptr), int(g), int(nq),
                                       int(k_in), int(k_out), P(out_d_ptr), P(out_i_ptr),
                                       P(stream)))


def merge_strided_device(device: int, in_d_ptr: int, in_i_ptr: int, stride_d: int, stride_i: int,
                         g: int, nq: int, k_in: int, k_out: int, out_d_ptr: int, out_i_ptr: int,
                         stream: int = 0) -> None:
    """K3 merge over a packed exchange buffer (shard strides in elements)."""
    lib = _native.load_library()
    P = ctypes.c_void_p
    _native.check(lib.fip_merge_strided_device(
        int(device), P(in_d_ptr), P(in_i_ptr), int(stride_d), int(stride_i), int(g), int(nq),
        int(k_in), int(k_out), P(out_d_ptr), P(out_i_ptr), P(stream)))
