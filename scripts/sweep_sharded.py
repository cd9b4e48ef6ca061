#!/usr/bin/env python
"""Batch-size sweep of BASELINE configs[3] on g GPUs: 10 M x 768 bf16, k = 100, B = 1 .. 4096, the
corpus row-sharded over the ranks (one process per GPU, K4 exchange + merge over NVLink).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node g --master-addr 127.0.0.1 \
        --master-port P scripts/sweep_sharded.py > profiles/r02_sweep_cfg4_..._n<g>.jsonl

Device-resident queries; every batch size: 3 warm-up searches, then `--iters` searches enqueued back
to back (asynchronous local search + exchange kernel, verdict read once at the end), CUDA events,
barrier on both sides, MAX over ranks.  One JSON line per batch size on rank 0: QPS, ms, the binding
roof (per-GPU algorithmic bytes or flops over the whole search) and its fraction of the measured peak.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=768)
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--batches", default="1,2,4,8,16,32,64,128,256,512,1024,2048,4096")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--exchange", default="auto")
    ap.add_argument("--balance", type=int, default=0, help="1: shards proportional to measured per-GPU throughput")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from bench import load_peaks
    from rag_inference_pipeline_b200.sharded import ShardedFlatIPSearcher
    from rag_inference_pipeline_b200.synthetic import CORPUS_SEED, synthetic_queries

    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    peaks = load_peaks()
    weights = None
    if world > 1 and args.balance:
        from rag_inference_pipeline_b200.group import calibrate_devices
        dist.barrier()
        rate = calibrate_devices([local], args.dim, args.dtype, batch=1024, k=args.k, rows=2_000_000, seconds=1.5)[0]
        every = [None] * world
        dist.all_gather_object(every, rate)
        weights = [float(r) for r in every]
    s = ShardedFlatIPSearcher(args.dim, args.dtype, args.rows, rank=rank, world_size=world, device=local,
                              exchange=args.exchange, weights=weights)
    s.load_synthetic(CORPUS_SEED)
    s.index.set_option("profile_events", 1)
    es = 4 if args.dtype == "f32" else 2
    n_local = -(-args.rows // world)
    for B in (int(b) for b in args.batches.split(",")):
        q = torch.from_numpy(synthetic_queries(B, args.dim)).to(dev)
        for _ in range(3):
            s.search_tensor(q, args.k)
        warm = s.check()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(args.iters):
            s.search_tensor(q, args.k)
        e1.record()
        barrier()
        status = s.check() | warm
        t = torch.tensor([e0.elapsed_time(e1) / args.iters, s.index.stat("scan_time_ns") / 1e6 / max(1, s.index.stat("scan_searches"))],
                         dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, scan_ms = (float(v) for v in t.tolist())
        if rank == 0:
            gbs = n_local * args.dim * es / ms / 1e6
            tfs = 2.0 * B * n_local * args.dim / ms / 1e9
            f_hbm, f_tc = gbs / peaks["hbm_gbs"], tfs / peaks["tflops_sustained"]
            line = {"gpus": world, "batch": B, "ms": round(ms, 4), "qps": round(B / ms * 1e3, 1),
                    "per_gpu_GBps": round(gbs, 1), "per_gpu_TFLOPs": round(tfs, 2),
                    "bound": "hbm" if f_hbm >= f_tc else "tensor", "frac": round(max(f_hbm, f_tc), 4),
                    "scan_ms": round(scan_ms, 4), "fixed_ms": round(ms - scan_ms, 4),
                    "variant": s.index.last_variant(), "exchange": s.exchange_used, "phases": s.index.stat("phases"),
                    "launches": s.index.stat("launches") + (1 if world > 1 else 0), "flagged": status}
            os.dup2(saved, 1)
            print(json.dumps(line), flush=True)
            os.dup2(2, 1)
    s.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
