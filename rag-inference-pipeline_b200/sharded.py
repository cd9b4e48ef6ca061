"""Row-sharded exact search across the GPUs of one box (one process per GPU, torch.distributed).

The corpus is split into contiguous row shards (rank r owns global rows
``[r * ceil(N/g), min(N, (r+1) * ceil(N/g)))``); every rank runs the local search on its shard
(queries replicated) straight into a packed ``[distances | ids]`` buffer, the per-shard results are
exchanged with ONE NCCL all-gather over NVLink (12 * B * k bytes per rank -- latency-bound, so one
collective instead of two), and every rank merges the g sorted lists with the K3 kernel.  Because shards are contiguous and ordered, "shard index, then position" equals
"id ascending", so the global tie-break (score desc, id asc) is preserved.

The reference has no counterpart (faiss-cpu in one process,
``/root/reference/src/pipeline/components/faiss_store.py:152``); upstream faiss calls this
IndexShards.  torch is used for the process group, the device buffers and the stream only.
"""

from __future__ import annotations

import numpy as np

from .index import FlatIPIndex, merge_strided_device, torch_stream_handle


def shard_range(total_rows: int, rank: int, world_size: int, weights=None) -> tuple[int, int]:
    """Global row range [begin, end) owned by `rank`.  ``weights`` (one positive number per rank,
    e.g. measured per-GPU throughput) makes the shards proportional -- boundaries on whole 256-row
    tiles, same rule as the C ABI's fip_group_create_weighted -- instead of equal."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    total_rows = int(total_rows)
    if weights is not None:
        if len(weights) != world_size or any(not (w > 0) for w in weights):
            raise ValueError("need one positive weight per rank")
        wsum = float(sum(weights))
        bounds = [0]
        acc = 0.0
        for r in range(world_size):
            acc += float(weights[r])
            b = int(total_rows * (acc / wsum) / 256.0 + 0.5) * 256
            bounds.append(max(bounds[-1], min(total_rows, b)))
        bounds[-1] = total_rows
        return bounds[rank], bounds[rank + 1]
    per = -(-total_rows // world_size)
    begin = min(total_rows, rank * per)
    end = min(total_rows, begin + per)
    return begin, end


def pack_layout(nq: int, k: int) -> tuple[int, int]:
    """Byte layout of one rank's exchange buffer: ``[nq*k fp32 | pad to 16 B | nq*k int64]``.
    Returns (offset of the ids, total bytes); total is a multiple of 16 so that the ids of every
    rank stay 8-byte aligned inside the gathered ``[g, total]`` buffer."""
    off_i = (int(nq) * int(k) * 4 + 15) // 16 * 16
    total = (off_i + int(nq) * int(k) * 8 + 15) // 16 * 16
    return off_i, total


def gather_and_merge(pack, nq: int, k: int, *, world_size: int, all_gather, merge):
    """The exchange step, backend-agnostic.  ``pack`` is this rank's packed result (pack_layout);
    ``all_gather(pack) -> [g, total]`` stacks the ranks' buffers in rank order (ONE collective for
    distances and ids); ``merge(all_pack, off_i, total) -> (D, I)``.  Split out so the
    world_size > 1 plumbing can be exercised on CPU with gloo (tests) while production passes the
    NCCL / K3 callables."""
    off_i, total = pack_layout(nq, k)
    if world_size == 1:
        return merge(pack.reshape(1, total), off_i, total)
    return merge(all_gather(pack), off_i, total)


class ShardedFlatIPSearcher:
    """One instance per rank.  ``search_tensor`` is the device-resident hot path."""

    def __init__(self, dim: int, dtype: str, total_rows: int, *, rank: int = 0,
                 world_size: int = 1, device: int = 0, group=None, exchange: str = "auto", index=None,
                 weights=None):
        """exchange: how the per-shard results meet.  "p2p" = the fused exchange + merge kernel
        over NVLink peer memory (K4), "nccl" = one packed all-gather + the K3 merge kernel,
        "auto" = p2p when the peers' buffers can be mapped (CUDA IPC), else nccl."""
        if exchange not in ("auto", "p2p", "nccl"):
            raise ValueError("exchange must be 'auto', 'p2p' or 'nccl'")
        self.exchange_mode = exchange
        self._xchg = None
        self._xchg_failed = False   # an attempt to map the peers' buffers failed: stay on NCCL
        self._status = None         # device word: OR of every shard's search status (p2p path)
        self.exchange_used = "none"
        self.dim = int(dim)
        self.dtype = dtype
        self.total_rows = int(total_rows)
        self.rank = int(rank)
        self.world_size = int(world_size)
        self.device = int(device)
        self.group = group
        self.row_begin, self.row_end = shard_range(total_rows, rank, world_size, weights)
        # index: an existing shard index to drive (e.g. the one a B200FlatIPStore holds) instead of
        # a new, empty one; it stays owned by whoever made it.
        self._owns_index = index is None
        self.index = index if index is not None else FlatIPIndex(
            dim, dtype, device, id_base=self.row_begin, capacity_rows=self.row_end - self.row_begin)

    @property
    def local_rows(self) -> int:
        return self.row_end - self.row_begin

    def load_synthetic(self, seed: int) -> None:
        """Generate this rank's rows of the Philox corpus on its own GPU."""
        self.index.reset()
        self.index.add_synthetic(self.local_rows, seed, self.row_begin)

    def load_rows(self, rows_global_slice: np.ndarray) -> None:
        """rows_global_slice: fp32 [local_rows, dim] = this rank's slice of the corpus."""
        if rows_global_slice.shape[0] != self.local_rows:
            raise ValueError("slice does not match this rank's shard")
        self.index.reset()
        self.index.add(rows_global_slice)

    # ------------------------------------------------------------------ device-resident path
    def search_tensor(self, queries, k: int):
        """queries: torch.float32 CUDA tensor [B, dim] (same on every rank).
        Returns (D [B, k] float32, I [B, k] int64) CUDA tensors holding the GLOBAL result.

        DEFERRED VERDICT: on a single rank and on the p2p exchange this only ENQUEUES the search on
        the caller's stream -- the tensors are filled once that stream reaches them, and they are
        final only if ``check()`` (one synchronisation, any time later) returns 0.  A nonzero
        status means some search since the previous ``check()`` overflowed a candidate list or
        could not be certified exact and has to be redone; ``search_tensor_checked`` does
        search + check + redo in one call and is the form to use when the result is consumed
        right away.  (``exchange="nccl"`` runs the blocking, self-checking local search: its
        results need no ``check()``, which then returns 0.)"""
        import torch

        if queries.dtype != torch.float32 or not queries.is_cuda or queries.dim() != 2:
            raise ValueError("queries must be a 2-D float32 CUDA tensor")
        queries = queries.contiguous()
        B = queries.shape[0]
        stream = torch_stream_handle(queries.device)
        off_i, total = pack_layout(B, k)
        pack = torch.empty((total,), dtype=torch.uint8, device=queries.device)
        d_loc = pack[:B * k * 4].view(torch.float32).view(B, k)
        i_loc = pack[off_i:off_i + B * k * 8].view(torch.int64).view(B, k)
        xchg = self._ensure_exchange(B, k) if self.world_size > 1 else None
        if self.world_size == 1 or xchg is not None:
            # Nothing between the local search and its consumer needs the host: the search is only
            # ENQUEUED, the K4 kernel ships every shard's status words with its rows, and check()
            # (one synchronisation, whenever the caller wants the verdict) says whether a search
            # since the last check has to be redone.  The GPU queue never drains between searches.
            self.index.search_device_async(queries.data_ptr(), B, k, d_loc.data_ptr(),
                                           i_loc.data_ptr(), stream)
            if self.world_size == 1:
                self._last = (d_loc, i_loc)
                return d_loc, i_loc
            if self._status is None:
                self._status = torch.zeros((1,), dtype=torch.int32, device=queries.device)
            d_out = torch.empty((B, k), dtype=torch.float32, device=queries.device)
            i_out = torch.empty((B, k), dtype=torch.int64, device=queries.device)
            xchg.merge_status(d_loc.data_ptr(), i_loc.data_ptr(), B, k, d_out.data_ptr(),
                              i_out.data_ptr(), self.index.flags_device_ptr(),
                              self._status.data_ptr(), stream)
            pack.record_stream(torch.cuda.current_stream(queries.device))
            self.exchange_used = "p2p"
            self._last = (d_out, i_out)
            return d_out, i_out
        # NCCL exchange: the local search is the blocking, self-checking form
        self.index.search_device(queries.data_ptr(), B, k, d_loc.data_ptr(), i_loc.data_ptr(),
                                 stream)
        self.exchange_used = "nccl"

        def all_gather(t):
            import torch.distributed as dist

            out = torch.empty((self.world_size, t.numel()), dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(out, t, group=self.group)
            return out

        def merge(all_pack, off, tot):
            d_out = torch.empty((B, k), dtype=torch.float32, device=queries.device)
            i_out = torch.empty((B, k), dtype=torch.int64, device=queries.device)
            merge_strided_device(self.device, all_pack.data_ptr(), all_pack.data_ptr() + off,
                                 tot // 4, tot // 8, self.world_size, B, k, k,
                                 d_out.data_ptr(), i_out.data_ptr(), stream)
            return d_out, i_out

        return gather_and_merge(pack, B, k, world_size=self.world_size, all_gather=all_gather,
                                merge=merge)

    # ------------------------------------------------------------------ deferred verdict
    _last = None

    def check(self) -> int:
        """Synchronises this rank's device and returns the OR of the search statuses since the last
        check (0 = every result stands; bit 0: some shard's candidate list overflowed, bit 1: some
        query could not be certified exact).  On the p2p path the status is GLOBAL (every rank sees
        the same value, it travelled with the rows), so every rank takes the same branch.  A
        nonzero status means the searches since the last check must be redone with
        ``search_tensor_checked`` -- ``search`` (the host-buffer path) does that itself."""
        import torch

        local = self.index.search_status()
        status = local
        if self._status is not None:
            status |= int(self._status.item())
            if status:
                self._status.zero_()
        elif self.world_size > 1:
            torch.cuda.synchronize(self.device)
        return status

    def search_tensor_checked(self, queries, k: int):
        """``search_tensor`` + ``check``; a flagged search is redone with the blocking local search
        (which redoes the flagged queries on its own) and exchanged again.  Collective on p2p."""
        d, i = self.search_tensor(queries, k)
        if self.exchange_used == "nccl" or not self.check():
            return d, i
        import torch

        B = queries.shape[0]
        stream = torch_stream_handle(queries.device)
        d_loc = torch.empty((B, k), dtype=torch.float32, device=queries.device)
        i_loc = torch.empty((B, k), dtype=torch.int64, device=queries.device)
        self.index.search_device(queries.data_ptr(), B, k, d_loc.data_ptr(), i_loc.data_ptr(), stream)
        if self.world_size == 1:
            return d_loc, i_loc
        d_out = torch.empty_like(d_loc)
        i_out = torch.empty_like(i_loc)
        self._xchg.merge(d_loc.data_ptr(), i_loc.data_ptr(), B, k, d_out.data_ptr(), i_out.data_ptr(), stream)
        torch.cuda.synchronize(self.device)
        return d_out, i_out

    # ------------------------------------------------------------------ peer-memory exchange
    def _ensure_exchange(self, nq: int, k: int):
        """Collective: (re)creates the K4 exchange when this search does not fit the current one.
        Every rank sees the same (nq, k), so every rank takes the same branch."""
        if self.exchange_mode == "nccl" or self._xchg_failed:
            return None                  # (an earlier attempt failed on some rank: stay on NCCL)
        if self._xchg is not None and self._xchg.fits(nq, k):
            return self._xchg
        import torch
        import torch.distributed as dist

        from .exchange import PeerExchange

        if self._xchg is not None:
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)   # nobody still writes into the old buffers
            self._xchg.close()
            self._xchg = None
        nq_max = max(1024, 1 << (int(nq) - 1).bit_length())
        k_max = max(128, 1 << (int(k) - 1).bit_length())
        k_max = min(k_max, 16384 // self.world_size)
        # Every rank walks the same collective sequence (all_gather_object, all_reduce) whether or
        # not its own steps succeed, so a failure on one rank cannot strand the others.
        x, err, handle = None, "", b""
        try:
            if k > k_max:
                raise ValueError("world_size * k exceeds the merge capacity")
            x = PeerExchange(self.device, self.world_size, self.rank, nq_max, k_max)
            handle = x.ipc_handle()
        except Exception as e:  # noqa: BLE001 - any failure means "no peer mapping on this box"
            err = f"{type(e).__name__}: {e}"
        handles = [None] * self.world_size
        dist.all_gather_object(handles, handle, group=self.group)
        if not err:
            try:
                if any(not h for h in handles):
                    raise RuntimeError("a peer could not allocate its exchange buffer")
                x.connect_ipc(handles)
            except Exception as e:  # noqa: BLE001
                err = f"{type(e).__name__}: {e}"
        # agreement on the host (one-time set-up, no device work): everybody learns whether every
        # rank could map every peer
        oks = [None] * self.world_size
        dist.all_gather_object(oks, not err, group=self.group)
        if not all(oks):
            if x is not None:
                x.close()
            self._xchg = None
            self._xchg_failed = True
            if self.exchange_mode == "p2p":
                raise RuntimeError(f"peer-memory exchange unavailable on rank {self.rank}: {err or 'a peer failed'}")
            return None
        self._xchg = x
        return x

    # ------------------------------------------------------------------ host-buffer path
    def search(self, queries: np.ndarray, k: int):
        """numpy in, numpy out (H2D / D2H inside), same contract as FAISSStore.search."""
        import torch

        dev = torch.device("cuda", self.device)
        q = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32)).to(dev)
        d, i = self.search_tensor_checked(q, k)
        return d.cpu().numpy(), i.cpu().numpy()

    def close(self) -> None:
        if self._xchg is not None:
            import torch
            import torch.distributed as dist

            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)   # peers may still be reading / writing our buffer
            self._xchg.close()
        self._xchg = None
        if self._owns_index:
            self.index.close()
