"""Row-sharded exact search across the GPUs of one box (one process per GPU, torch.distributed).

The corpus is split into contiguous row shards (rank r owns global rows
``[r * ceil(N/g), min(N, (r+1) * ceil(N/g)))``); every rank runs the local search on its shard
(queries replicated), the per-shard ``[B, k]`` results are exchanged with ONE NCCL all-gather over
NVLink (12 * B * k bytes per rank -- latency-bound), and every rank merges the g sorted lists with
the K3 kernel.  Because shards are contiguous and ordered, "shard index, then position" equals
"id ascending", so the global tie-break (score desc, id asc) is preserved.

The reference has no counterpart (faiss-cpu in one process,
``/root/reference/src/pipeline/components/faiss_store.py:152``); upstream faiss calls this
IndexShards.  torch is used for the process group, the device buffers and the stream only.
"""

from __future__ import annotations

import numpy as np

from .index import FlatIPIndex, merge_device


def shard_range(total_rows: int, rank: int, world_size: int) -> tuple[int, int]:
    """Global row range [begin, end) owned by `rank`."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world_size")
    per = -(-int(total_rows) // world_size)
    begin = min(int(total_rows), rank * per)
    end = min(int(total_rows), begin + per)
    return begin, end


def gather_and_merge(local_d, local_i, k: int, *, world_size: int, all_gather, merge):
    """The exchange step, backend-agnostic: ``all_gather(t) -> [g, ...]`` stacked in rank order,
    ``merge(all_d, all_i, k) -> (D, I)``.  Split out so the world_size > 1 plumbing can be
    exercised on CPU with gloo (tests) while production passes the NCCL / K3 callables."""
    if world_size == 1:
        return local_d, local_i
    all_d = all_gather(local_d)
    all_i = all_gather(local_i)
    return merge(all_d, all_i, k)


class ShardedFlatIPSearcher:
    """One instance per rank.  ``search_tensor`` is the device-resident hot path."""

    def __init__(self, dim: int, dtype: str, total_rows: int, *, rank: int = 0,
                 world_size: int = 1, device: int = 0, group=None):
        self.dim = int(dim)
        self.dtype = dtype
        self.total_rows = int(total_rows)
        self.rank = int(rank)
        self.world_size = int(world_size)
        self.device = int(device)
        self.group = group
        self.row_begin, self.row_end = shard_range(total_rows, rank, world_size)
        self.index = FlatIPIndex(dim, dtype, device, id_base=self.row_begin,
                                 capacity_rows=self.row_end - self.row_begin)

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
        Returns (D [B, k] float32, I [B, k] int64) CUDA tensors holding the GLOBAL result."""
        import torch

        if queries.dtype != torch.float32 or not queries.is_cuda or queries.dim() != 2:
            raise ValueError("queries must be a 2-D float32 CUDA tensor")
        queries = queries.contiguous()
        B = queries.shape[0]
        stream = torch.cuda.current_stream(queries.device).cuda_stream
        d_loc = torch.empty((B, k), dtype=torch.float32, device=queries.device)
        i_loc = torch.empty((B, k), dtype=torch.int64, device=queries.device)
        self.index.search_device(queries.data_ptr(), B, k, d_loc.data_ptr(), i_loc.data_ptr(),
                                 stream)

        def all_gather(t):
            import torch.distributed as dist

            out = torch.empty((self.world_size,) + tuple(t.shape), dtype=t.dtype, device=t.device)
            dist.all_gather_into_tensor(out, t, group=self.group)
            return out

        def merge(all_d, all_i, kk):
            d_out = torch.empty((B, kk), dtype=torch.float32, device=queries.device)
            i_out = torch.empty((B, kk), dtype=torch.int64, device=queries.device)
            merge_device(self.device, all_d.data_ptr(), all_i.data_ptr(), self.world_size, B, kk,
                         kk, d_out.data_ptr(), i_out.data_ptr(),
                         torch.cuda.current_stream(queries.device).cuda_stream)
            return d_out, i_out

        return gather_and_merge(d_loc, i_loc, k, world_size=self.world_size,
                                all_gather=all_gather, merge=merge)

    # ------------------------------------------------------------------ host-buffer path
    def search(self, queries: np.ndarray, k: int):
        """numpy in, numpy out (H2D / D2H inside), same contract as FAISSStore.search."""
        import torch

        dev = torch.device("cuda", self.device)
        q = torch.from_numpy(np.ascontiguousarray(queries, dtype=np.float32)).to(dev)
        d, i = self.search_tensor(q, k)
        return d.cpu().numpy(), i.cpu().numpy()

    def close(self) -> None:
        self.index.close()
