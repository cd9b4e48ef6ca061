#!/usr/bin/env python
"""COMPARATOR ONLY (never on the product path): how fast is cuBLAS on the headline scan's GEMM
shape?  VERDICT r1 item 7: K1 runs at 0.87 of the sustained cuBLAS bf16 peak measured on 8192^3;
this times the library on the shape K1 actually has -- [B x dim] x [dim x rows] in fp16 with fp32
accumulation, B = 1024, dim = 768, rows swept in chunks -- under the same power cap, with the SM
clock sampled during the run, so the two can be compared like for like.

The library must WRITE its [B x rows] fp16 result (K1 never does); rows per call are chosen so
that this write stays in the noise of a tensor-bound call (2 bytes per 2*768 flops).

usage: python scripts/cublas_shape_comparator.py [--rows 12500000] [--batch 1024] [--dim 768]
Prints one JSON line.
"""
import argparse
import json
import os
import subprocess
import threading
import time

import torch

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=12_500_000)
ap.add_argument("--batch", type=int, default=1024)
ap.add_argument("--dim", type=int, default=768)
ap.add_argument("--chunk", type=int, default=2_500_000)
ap.add_argument("--seconds", type=float, default=4.0)
ap.add_argument("--dtype", default="f16")
args = ap.parse_args()

dt = {"f16": torch.float16, "bf16": torch.bfloat16}[args.dtype]
dev = torch.device("cuda", 0)
torch.manual_seed(0)
x = torch.randn((args.rows, args.dim), device=dev, dtype=dt)
q = torch.randn((args.batch, args.dim), device=dev, dtype=dt)
out = torch.empty((args.batch, args.chunk), device=dev, dtype=dt)

lines = []
proc = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits",
                         "-lms", "100", "-i", "0"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
threading.Thread(target=lambda: [lines.append((time.time(), l.strip())) for l in proc.stdout], daemon=True).start()


def one_pass():
    for r0 in range(0, args.rows, args.chunk):
        r1 = min(args.rows, r0 + args.chunk)
        torch.matmul(q, x[r0:r1].t(), out=out[:, : r1 - r0])


for _ in range(3):
    one_pass()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
passes = 0
t_begin = time.time()
e0.record()
while time.time() - t_begin < args.seconds:
    one_pass()
    passes += 1
    torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
t_end = time.time()
ms = e0.elapsed_time(e1) / passes
proc.terminate()
clk = sorted(float(l.split(",")[0]) for t, l in lines if t_begin + 0.5 <= t <= t_end and "," in l)
pw = sorted(float(l.split(",")[1]) for t, l in lines if t_begin + 0.5 <= t <= t_end and "," in l)
flops = 2.0 * args.batch * args.rows * args.dim
peaks = {}
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    peaks = json.load(open(p))
res = {"what": "cuBLAS (torch.matmul) on the scan's GEMM shape, comparator only",
       "shape": f"[{args.batch} x {args.dim}] x [{args.dim} x {args.rows}] {args.dtype}, fp32 accumulate, "
                f"{args.chunk}-row calls, result written as {args.dtype}",
       "ms_per_pass": ms, "tflops": flops / (ms * 1e-3) / 1e12, "passes": passes,
       "sm_mhz_median": clk[len(clk) // 2] if clk else None, "power_w_median": pw[len(pw) // 2] if pw else None,
       "of_sustained_8192_cube_peak": (flops / (ms * 1e-3) / 1e12) / peaks["bf16_tflops_sustained"] if peaks else None,
       "output_write_gb_per_pass": args.batch * args.rows * 2 / 1e9}
print(json.dumps(res))
