"""`FlatIPGroup`: ONE corpus row-sharded over several GPUs of this process (C ABI ``fip_group_*``).

Same faiss-like surface as ``FlatIPIndex`` (``d``, ``ntotal``, ``add``, ``search``), so
``B200FlatIPStore`` can put either behind the reference's single call site
``registry.get("faiss_store").search(embeddings, k)``
(``/root/reference/src/pipeline/services/retrieval/api.py:383-389``).  One host thread + stream per
device inside the native call, K4 exchange + merge over NVLink peer memory, global ids, fixed
tie-break; see ``csrc/group_api.cuh``.
"""

from __future__ import annotations

import ctypes

import numpy as np

from . import _native
from .index import raw_rows


class _ShardView:
    """Statistics / options of one shard (the handle is owned by the group)."""

    def __init__(self, lib, handle):
        self._lib, self._h = lib, handle

    def stat(self, key: str) -> int:
        return int(self._lib.fip_get_stat(self._h, key.encode()))

    def set_option(self, key: str, value: int) -> None:
        _native.check(self._lib.fip_set_option(self._h, key.encode(), int(value)))

    @property
    def ntotal(self) -> int:
        return int(self._lib.fip_ntotal(self._h))

    def reconstruct_n(self, row0: int, n: int, dim: int) -> np.ndarray:
        out = np.empty((n, dim), dtype=np.float32)
        _native.check(self._lib.fip_reconstruct(self._h, int(row0), int(n), out.ctypes.data))
        return out


def calibrate_devices(devices, dim: int, dtype: str, *, batch: int = 1024, k: int = 100,
                      rows: int = 2_000_000, seconds: float = 1.5) -> list[float]:
    """Sustained search throughput (corpus rows per second of scan-kernel time) of every device,
    measured on all of them AT THE SAME TIME: under the power cap the GPUs of one box differ by
    several per cent, and a sharded search ends with the slowest one.  Feed the result to
    ``FlatIPGroup(weights=...)`` / ``ShardedFlatIPSearcher(weights=...)``."""
    lib = _native.load_library()
    devices = [int(d) for d in devices]
    arr = (ctypes.c_int32 * len(devices))(*devices)
    out = (ctypes.c_double * len(devices))()
    _native.check(lib.fip_calibrate_devices(arr, len(devices), int(dim), _native.DTYPE_CODES[dtype], int(rows),
                                            int(batch), int(k), float(seconds), out))
    return [float(v) for v in out]


class FlatIPGroup:
    def __init__(self, dim: int, dtype: str, devices, planned_rows: int, *, id_base: int = 0, weights=None):
        if dtype not in _native.DTYPE_CODES:
            raise ValueError(f"unknown storage dtype {dtype!r}")
        devices = [int(d) for d in devices]
        if not devices:
            raise ValueError("need at least one device")
        self._lib = _native.load_library()
        cfg = _native.FipConfig(dim=int(dim), dtype=_native.DTYPE_CODES[dtype], device=devices[0],
                                reserved=0, capacity_rows=0, id_base=int(id_base))
        arr = (ctypes.c_int32 * len(devices))(*devices)
        h = ctypes.c_void_p()
        warr = None
        if weights is not None:
            if len(weights) != len(devices):
                raise ValueError("one weight per device")
            warr = (ctypes.c_double * len(devices))(*[float(w) for w in weights])
        _native.check(self._lib.fip_group_create_weighted(ctypes.byref(cfg), arr, len(devices), int(planned_rows),
                                                          warr, ctypes.byref(h)))
        self._h = h
        self.d = int(dim)
        self.dtype = _native.DTYPE_NAMES[_native.DTYPE_CODES[dtype]]
        self.devices = devices
        self.device = devices[0]
        self.id_base = int(id_base)

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fip_group_destroy(self._h)
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

    @property
    def ntotal(self) -> int:
        return int(self._lib.fip_group_ntotal(self._handle()))

    def shard_bounds(self) -> list[int]:
        return [self.stat(f"shard_begin_{r}") for r in range(len(self.devices) + 1)]

    def shard(self, r: int) -> _ShardView:
        h = self._lib.fip_group_shard(self._handle(), int(r))
        if not h:
            raise IndexError(r)
        return _ShardView(self._lib, ctypes.c_void_p(h))

    def add(self, x: np.ndarray) -> None:
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.d:
            raise ValueError(f"expected rows of shape (n, {self.d}), got {x.shape}")
        _native.check(self._lib.fip_group_add_host(self._handle(), x.ctypes.data, x.shape[0], _native.FIP_F32))

    def add_raw(self, x: np.ndarray, dtype: str) -> None:
        _native.check(self._lib.fip_group_add_host(self._handle(), *raw_rows(x, dtype, self.d)))

    def add_synthetic(self, n: int, seed: int, global_row0: int | None = None) -> None:
        """Next n global rows of the Philox corpus (global_row0, if given, must equal id_base +
        ntotal: rows are always appended in global order)."""
        if global_row0 is not None and int(global_row0) != self.id_base + self.ntotal:
            raise ValueError("a group appends rows in global order")
        _native.check(self._lib.fip_group_add_synthetic(self._handle(), int(n), int(seed)))

    def reset(self) -> None:
        _native.check(self._lib.fip_group_reset(self._handle()))

    def search(self, x: np.ndarray, k: int):
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 2 or x.shape[1] != self.d:
            raise ValueError(f"expected queries of shape (n, {self.d}), got {x.shape}")
        n = x.shape[0]
        D = np.empty((n, int(k)), dtype=np.float32)
        I = np.empty((n, int(k)), dtype=np.int64)
        _native.check(self._lib.fip_group_search(self._handle(), x.ctypes.data, n, int(k),
                                                 D.ctypes.data, I.ctypes.data))
        return D, I

    def set_option(self, key: str, value: int) -> None:
        _native.check(self._lib.fip_group_set_option(self._handle(), key.encode(), int(value)))

    def set_variant(self, name: str) -> None:
        self.set_option("variant", _native.VARIANTS[name])

    def stat(self, key: str) -> int:
        return int(self._lib.fip_group_get_stat(self._handle(), key.encode()))

    def last_variant(self) -> str:
        return _native.VARIANT_NAMES.get(self.stat("variant"), "none")
