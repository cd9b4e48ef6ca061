"""`PeerExchange`: host-side handle of the fused exchange + merge kernel (K4, csrc/exchange.cuh).

One object per rank.  The buffers are peer-mapped either through CUDA IPC handles (one process per
GPU: the handles travel over the caller's process group) or directly (several ranks inside one
process).  The reference has no counterpart (faiss-cpu in one process,
``/root/reference/src/pipeline/components/faiss_store.py:152``).
"""

from __future__ import annotations

import ctypes

from . import _native


class PeerExchange:
    def __init__(self, device: int, world_size: int, rank: int, nq_max: int, k_max: int):
        self._lib = _native.load_library()
        h = ctypes.c_void_p()
        _native.check(self._lib.fip_exchange_create(int(device), int(world_size), int(rank),
                                                    int(nq_max), int(k_max), ctypes.byref(h)))
        self._h = h
        self.device, self.world_size, self.rank = int(device), int(world_size), int(rank)
        self.nq_max, self.k_max = int(nq_max), int(k_max)

    def ipc_handle(self) -> bytes:
        buf = ctypes.create_string_buffer(_native.IPC_HANDLE_BYTES)
        _native.check(self._lib.fip_exchange_ipc_handle(self._h, buf))
        return buf.raw

    def connect_ipc(self, handles: list[bytes]) -> None:
        """handles[r] = rank r's ``ipc_handle()`` (the own entry is ignored)."""
        if len(handles) != self.world_size or any(len(h) != _native.IPC_HANDLE_BYTES for h in handles):
            raise ValueError("need one 64-byte handle per rank")
        blob = b"".join(handles)
        _native.check(self._lib.fip_exchange_connect_ipc(self._h, blob))

    @staticmethod
    def connect_local(group: list["PeerExchange"]) -> None:
        """All ranks live in this process: wire every exchange to every other one."""
        arr = (ctypes.c_void_p * len(group))(*[x._h for x in group])
        for x in group:
            _native.check(x._lib.fip_exchange_connect_local(x._h, arr))

    def fits(self, nq: int, k: int) -> bool:
        return nq <= self.nq_max and k <= self.k_max

    def merge(self, d_ptr: int, i_ptr: int, nq: int, k: int, out_d_ptr: int, out_i_ptr: int,
              stream: int = 0) -> None:
        P = ctypes.c_void_p
        _native.check(self._lib.fip_exchange_merge(self._h, P(d_ptr), P(i_ptr), int(nq), int(k),
                                                   P(out_d_ptr), P(out_i_ptr), P(stream)))

    def merge_status(self, d_ptr: int, i_ptr: int, nq: int, k: int, out_d_ptr: int, out_i_ptr: int,
                     flags_ptr: int, status_ptr: int, stream: int = 0) -> None:
        """``merge`` with the shards' search statuses travelling with the rows: ``flags_ptr`` = the
        local index's ``flags_device_ptr()``; the device word at ``status_ptr`` is OR-ed with bit 0 /
        bit 1 when ANY rank's local search overflowed / could not be certified."""
        P = ctypes.c_void_p
        _native.check(self._lib.fip_exchange_merge_status(
            self._h, P(d_ptr), P(i_ptr), int(nq), int(k), P(out_d_ptr), P(out_i_ptr), P(flags_ptr),
            P(status_ptr), P(stream)))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.fip_exchange_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover - best effort
        try:
            self.close()
        except Exception:
            pass
