"""world_size-2 CPU test (gloo) of the row-sharded plumbing: shard ranges, rank-ordered
all-gather, merge contract.  The local search and the merge are the ORACLE here (test
infrastructure) -- on a GPU box the same `gather_and_merge` is fed the CUDA search and K3."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rag_inference_pipeline_b200.sharded import gather_and_merge, pack_layout, shard_range


def test_shard_ranges_cover_exactly():
    for total, g in [(100, 8), (7, 8), (0, 2), (100_000_000, 8), (12_500_001, 3)]:
        covered = []
        for r in range(g):
            b, e = shard_range(total, r, g)
            assert 0 <= b <= e <= total
            covered.append((b, e))
        assert covered[0][0] == 0 and covered[-1][1] == total
        for (b0, e0), (b1, e1) in zip(covered, covered[1:]):
            assert e0 == b1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_pack_layout_alignment():
    for nq, k in [(1, 1), (3, 5), (7, 30), (1024, 100), (256, 1024)]:
        off_i, total = pack_layout(nq, k)
        assert off_i >= nq * k * 4 and off_i % 16 == 0
        assert total >= off_i + nq * k * 8 and total % 16 == 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, total, dim, nq, k, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "oracle"))
    import flatip_oracle
    from rag_inference_pipeline_b200 import synthetic

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = shard_range(total, rank, world)
    corpus = synthetic.synthetic_corpus(b, e - b, dim, "f16")
    q = synthetic.synthetic_queries(nq, dim)
    d_loc, i_loc = flatip_oracle.flat_ip_search(q, corpus, k, id_base=b)

    # this rank's packed exchange buffer, laid out exactly like the CUDA path's
    off_i, total = pack_layout(nq, k)
    pack = np.zeros(total, dtype=np.uint8)
    pack[:nq * k * 4] = np.ascontiguousarray(d_loc).view(np.uint8).ravel()
    pack[off_i:off_i + nq * k * 8] = np.ascontiguousarray(i_loc).view(np.uint8).ravel()
    calls = []

    def all_gather(t):
        calls.append(1)
        t = torch.from_numpy(t)
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return torch.stack(out).numpy()

    def merge(all_pack, off, tot):
        assert all_pack.shape == (world, tot) and off == off_i
        all_d = np.stack([all_pack[r, :nq * k * 4].view(np.float32).reshape(nq, k) for r in range(world)])
        all_i = np.stack([all_pack[r, off:off + nq * k * 8].view(np.int64).reshape(nq, k) for r in range(world)])
        return flatip_oracle.merge_sorted_lists(all_d, all_i, k)

    D, I = gather_and_merge(pack, nq, k, world_size=world, all_gather=all_gather, merge=merge)
    assert len(calls) == 1          # distances and ids travel in ONE collective
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), D=D, I=I)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_and_merge(tmp_path):
    import flatip_oracle
    from rag_inference_pipeline_b200 import synthetic

    total, dim, nq, k, world = 5001, 32, 7, 30, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, total, dim, nq, k, str(tmp_path)), nprocs=world, join=True)
    corpus = synthetic.synthetic_corpus(0, total, dim, "f16")
    q = synthetic.synthetic_queries(nq, dim)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    for rank in range(world):
        got = np.load(tmp_path / f"rank{rank}.npz")
        assert np.array_equal(got["I"], Ir)      # every rank holds the global answer
        assert np.array_equal(got["D"], Dr)


def _xchg_fail_worker(rank, world, port, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    from rag_inference_pipeline_b200 import exchange as exchange_mod
    from rag_inference_pipeline_b200.sharded import ShardedFlatIPSearcher

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    attempts = []

    class Boom:
        def __init__(self, *a, **kw):
            attempts.append(1)
            if rank == 1:   # only ONE rank fails: the others must follow it onto NCCL
                raise RuntimeError("no peer mapping on this box")

        def ipc_handle(self):
            return b"\x01" * 64

        def connect_ipc(self, handles):
            pass

        def close(self):
            pass

    exchange_mod.PeerExchange = Boom
    s = object.__new__(ShardedFlatIPSearcher)   # no GPU here: only the exchange set-up is under test
    s.exchange_mode, s._xchg, s._xchg_failed, s._status = "auto", None, False, None
    s.device, s.world_size, s.rank, s.group = 0, world, rank, None
    first = s._ensure_exchange(8, 10)
    second = s._ensure_exchange(8, 10)     # used to raise AttributeError ('bool' has no 'fits')
    third = s._ensure_exchange(4096, 100)
    ok = first is None and second is None and third is None and len(attempts) == 1 and s._xchg_failed
    s2 = object.__new__(ShardedFlatIPSearcher)
    s2.exchange_mode, s2._xchg, s2._xchg_failed, s2._status = "p2p", None, False, None
    s2.device, s2.world_size, s2.rank, s2.group = 0, world, rank, None
    try:
        s2._ensure_exchange(8, 10)
        ok = False                         # an explicit "p2p" request must fail loudly
    except RuntimeError as e:
        ok = ok and "peer-memory exchange unavailable" in str(e)
    with open(os.path.join(out_dir, f"ok{rank}"), "w") as f:
        f.write("1" if ok else "0")
    dist.barrier()
    dist.destroy_process_group()


def test_failed_peer_mapping_falls_back_to_nccl_for_good(tmp_path):
    """after a failed peer-memory attempt in exchange="auto" every later search must stay
    on NCCL (no second attempt, no crash), on EVERY rank, also when only one rank failed."""
    world = 2
    mp.spawn(_xchg_fail_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        assert (tmp_path / f"ok{rank}").read_text() == "1"


def test_weighted_shard_ranges():
    """Shards proportional to per-GPU throughput: contiguous, ordered, on 256-row tiles, covering
    the corpus exactly; equal weights reproduce (to a tile) the equal split."""
    total, w = 100_000_000, [1.00, 1.12, 1.01, 1.06, 1.05, 1.08, 1.09, 1.05]
    rng = [shard_range(total, r, 8, w) for r in range(8)]
    assert rng[0][0] == 0 and rng[-1][1] == total
    for (b0, e0), (b1, e1) in zip(rng, rng[1:]):
        assert e0 == b1 and b1 % 256 == 0
    sizes = [e - b for b, e in rng]
    for r in range(8):
        assert abs(sizes[r] / total - w[r] / sum(w)) < 1e-5
    eq = [shard_range(10_000, r, 4, [1, 1, 1, 1]) for r in range(4)]
    assert [b for b, _ in eq] == [0, 2560, 5120, 7424] and eq[-1][1] == 10_000
    with pytest.raises(ValueError):
        shard_range(100, 0, 2, [1.0, 0.0])


def test_shard_ranges_property():
    """Any positive weights, any size: contiguous, ordered, exact cover; interior boundaries on tiles."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=200, deadline=None)
    @given(st.integers(0, 3_000_000_000), st.lists(st.floats(0.05, 20.0), min_size=1, max_size=16))
    def check(total, weights):
        g = len(weights)
        prev = 0
        for r in range(g):
            b, e = shard_range(total, r, g, weights)
            assert b == prev and b <= e <= total
            if 0 < r and b < total:
                assert b % 256 == 0
            prev = e
        assert prev == total

    check()
