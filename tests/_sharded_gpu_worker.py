"""torchrun worker for the multi-GPU parity test (tests/test_gpu_multi.py): every rank searches its
row shard on its own GPU, the results meet through the requested exchange, rank 0 checks the
global answer against the oracle and against the other exchange mode."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import numpy as np
import torch
import torch.distributed as dist


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    import flatip_oracle
    from rag_inference_pipeline_b200 import synthetic
    from rag_inference_pipeline_b200.sharded import ShardedFlatIPSearcher

    dim, total, dtype = 128, 200_003, "f16"
    results = {}
    for mode in ("nccl", "p2p"):
        s = ShardedFlatIPSearcher(dim, dtype, total, rank=rank, world_size=world, device=local,
                                  exchange=mode)
        s.load_synthetic(synthetic.CORPUS_SEED)
        out = []
        for nq, k in [(48, 100), (1, 10), (300, 7), (48, 100), (1100, 130)]:
            q = torch.from_numpy(synthetic.synthetic_queries(nq, dim)).to(dev)
            D, I = s.search_tensor(q, k)
            status = s.check()     # deferred verdict (p2p: global, the same on every rank)
            assert status == 0, (mode, nq, k, status)
            assert s.exchange_used == mode, (s.exchange_used, mode)
            out.append((nq, k, D.cpu().numpy(), I.cpu().numpy()))
        results[mode] = out
        s.close()
    if rank == 0:
        corpus = synthetic.synthetic_corpus(0, total, dim, dtype)
        for (nq, k, Dn, In), (_, _, Dp, Ip) in zip(results["nccl"], results["p2p"]):
            assert np.array_equal(In, Ip) and np.array_equal(Dn, Dp), ("nccl vs p2p", nq, k)
            Dr, Ir = flatip_oracle.flat_ip_search(synthetic.synthetic_queries(nq, dim), corpus, k)
            res = flatip_oracle.compare_with_oracle(Dp, Ip, Dr, Ir, rtol=1e-3)
            assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, (nq, k, res)
        print("MULTI_GPU_PARITY_OK world=%d" % world, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
