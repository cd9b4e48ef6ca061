"""Batch-size sweep (BASELINE configs[3]): 10 M x 768 bf16, k = 100, B = 1 .. 4096 on one GPU.

One index, many batches; device-resident queries, CUDA events on the launching stream.  Prints one
JSON object per batch (QPS, ms, binding roofline and fraction of the measured peak).

usage: python scripts/sweep.py [--rows N] [--dim D] [--dtype bf16] [--k 100] [--batches 1,2,4,...]
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
    ap.add_argument("--variant", default="auto")
    ap.add_argument("--opt", action="append", default=[], help="key=value index option")
    args = ap.parse_args()

    import torch

    from bench import load_peaks
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.synthetic import synthetic_queries

    peaks = load_peaks()
    idx = FlatIPIndex(args.dim, args.dtype, 0, capacity_rows=args.rows)
    idx.add_synthetic(args.rows, 1234, 0)
    idx.set_variant(args.variant)
    for kv in args.opt:
        key, val = kv.split("=")
        idx.set_option(key, int(val))
    idx.set_option("profile_events", 1)
    es = 4 if args.dtype == "f32" else 2
    from rag_inference_pipeline_b200.index import torch_stream_handle
    s = torch_stream_handle()
    for B in (int(b) for b in args.batches.split(",")):
        q = torch.from_numpy(synthetic_queries(B, args.dim)).cuda()
        D = torch.empty((B, args.k), dtype=torch.float32, device="cuda")
        I = torch.empty((B, args.k), dtype=torch.int64, device="cuda")
        for _ in range(3):
            idx.search_device(q.data_ptr(), B, args.k, D.data_ptr(), I.data_ptr(), s)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            idx.search_device(q.data_ptr(), B, args.k, D.data_ptr(), I.data_ptr(), s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        gbs = args.rows * args.dim * es / ms / 1e6
        tfs = 2.0 * B * args.rows * args.dim / ms / 1e9
        f_hbm, f_tc = gbs / peaks["hbm_gbs"], tfs / peaks["tflops_sustained"]
        print(json.dumps({
            "batch": B, "ms": round(ms, 4), "qps": round(B / ms * 1e3, 1), "GBps": round(gbs, 1),
            "TFLOPs": round(tfs, 2), "bound": "hbm" if f_hbm >= f_tc else "tensor",
            "frac": round(max(f_hbm, f_tc), 4), "variant": idx.last_variant(),
            "phases": idx.stat("phases"), "launches": idx.stat("launches"),
            "scan_ms": round(idx.stat("scan_time_ns") / 1e6, 4),
            "uncert": idx.stat("uncertified_queries"), "overflow": idx.stat("overflow_queries"),
            "cta_group": idx.stat("cta_group"),
        }), flush=True)
    idx.close()


if __name__ == "__main__":
    main()
