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

from rag_inference_pipeline_b200.sharded import gather_and_merge, shard_range


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

    def all_gather(t):
        t = torch.from_numpy(np.ascontiguousarray(t))
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return torch.stack(out).numpy()

    D, I = gather_and_merge(d_loc, i_loc, k, world_size=world, all_gather=all_gather,
                            merge=flatip_oracle.merge_sorted_lists)
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
