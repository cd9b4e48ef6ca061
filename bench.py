#!/usr/bin/env python
"""Benchmark of the flat inner-product retrieval hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU algorithm

Workload (default `cfg3`, the configuration the metric is quoted on): exact flat-IP search over a
synthetic 100 M x 768 fp16 corpus, batch 1024, k = 100, the corpus row-sharded over the N GPUs of
one box (strong scaling: total rows fixed).  A "step" is one search of one batch.  Prints ONE JSON
line on rank 0.  Other workloads (`--workload cfg2|cfg4|cfg5`, `--rows`, `--batch`) are for the
parity / sweep numbers quoted in DESIGN.md, not bench lines.

For N > 1 launch with torchrun (one rank per GPU); RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* are
read from the environment.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

# Must be set before ANY OpenMP runtime starts (torch loads one): spinning libgomp workers starve
# the BLAS threads of the CPU baseline (same policy as the reference, src/pipeline/runtime.py:51).
os.environ.setdefault("OMP_WAIT_POLICY", "passive")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (total_rows, dim, dtype, batch, k)
    "cfg2": (1_000_000, 768, "f32", 32, 10),
    "cfg3": (100_000_000, 768, "f16", 1024, 100),
    "cfg4": (10_000_000, 768, "bf16", 1024, 100),
    "cfg5": (10_000_000, 384, "f16", 256, 1024),
}
METRIC = "QPS (100M x 768 fp16 flat-IP, k=100, batch 1024)"


def metric_name(args, total_rows, dim, dtype, batch, k):
    """BASELINE.json's metric for the default workload; the other workloads name their own shape."""
    if args.workload == "cfg3" and not (args.rows or args.batch or args.k):
        return METRIC
    return f"QPS ({total_rows} x {dim} {dtype} flat-IP, k={k}, batch {batch})"
UNIT = "queries/s"


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "tflops_burst": float(p["bf16_tflops"]),
                "tflops_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0,
            "source": "fallback"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.gpu_index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[4:8]):
                if val.lower() == "active":
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# CPU legs (oracle = test infrastructure; used ONLY as the reported baseline / reference arm)
# ------------------------------------------------------------------------------------------
def host_mem_available_bytes():
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable:"):
                    return int(line.split()[1]) * 1024
    except OSError:
        pass
    return 16 << 30


def cpu_oracle_qps(total_rows, dim, dtype, batch, k, *, step_budget_s=12.0, steps=1, warmup=0,
                   max_rows=10_000_000):
    """Times the C/OpenMP/BLAS restatement of faiss-cpu's IndexFlatIP.search (oracle/) on a bounded
    sample of the workload and scales linearly in corpus rows (a flat scan is linear in N).

    Threads are set EXPLICITLY to every host core (torch.distributed.run exports OMP_NUM_THREADS=1,
    which must not leak into the baseline); the reference's own default is min(16, cpu_count)
    threads per node (/root/reference/src/pipeline/config/__init__.py:32-46, applied with
    faiss.omp_set_num_threads, runtime.py:76) -- when that differs from all cores it is timed too
    and reported as `reference_default_threads`.  The sample is the largest row count <= max_rows
    (10 M: SURVEY 8(d)) that keeps one step within step_budget_s and the fp32 copy within 40 % of
    the free host memory."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import c_oracle
    from rag_inference_pipeline_b200.synthetic import synthetic_queries

    ncores = os.cpu_count() or 1
    threads = c_oracle.set_threads(ncores)
    q = synthetic_queries(batch, dim)
    pilot_rows = min(total_rows, 100_000)
    x = c_oracle.synthetic_corpus(0, pilot_rows, dim, dtype)
    c_oracle.knn_inner_product(q, x, k)  # warm (thread pool, BLAS init)
    t0 = time.perf_counter()
    c_oracle.knn_inner_product(q, x, k)
    pilot = time.perf_counter() - t0
    mem_rows = int(0.4 * host_mem_available_bytes() / (dim * 4))
    rows = int(min(total_rows, max_rows, mem_rows,
                   max(pilot_rows, pilot_rows * step_budget_s / max(pilot, 1e-6))))
    if rows != pilot_rows:
        del x
        x = c_oracle.synthetic_corpus(0, rows, dim, dtype)
    for _ in range(warmup):
        c_oracle.knn_inner_product(q, x, k)
    times = []
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        c_oracle.knn_inner_product(q, x, k)
        times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    factor = total_rows / rows
    scaled = t * factor
    ref_default = None
    ref_threads = min(16, ncores)
    if ref_threads != ncores:
        c_oracle.set_threads(ref_threads)
        c_oracle.knn_inner_product(q, x, k)
        t0 = time.perf_counter()
        c_oracle.knn_inner_product(q, x, k)
        t16 = time.perf_counter() - t0
        ref_default = {"threads": ref_threads, "value": batch / (t16 * factor), "unit": UNIT}
        c_oracle.set_threads(ncores)
    else:
        ref_default = {"threads": ref_threads, "value": batch / scaled, "unit": UNIT,
                       "note": "equals all cores on this host"}
    return {
        "value": batch / scaled, "unit": UNIT, "cores": threads, "kind": "port",
        "sample": (f"oracle/faiss_flat_restatement.c (the builder's C port of faiss-cpu's algorithm: sgemm "
                   f"blocks + {'reservoir' if k >= 100 else 'heap'}, OpenBLAS; NOT faiss itself) on {rows} of "
                   f"{total_rows} rows (fp32 copies of the {dtype} values), batch {batch}, k {k}, "
                   f"{threads} threads set explicitly; {t:.2f} s per search of the sample, scaled "
                   f"linearly x{factor:.2f} to {total_rows} rows; host cpu_count={ncores}"),
        "rows_timed": rows, "extrapolation_factor": factor, "ms_per_step_sample": t * 1e3,
        "ms_per_step_scaled": scaled * 1e3, "reference_default_threads": ref_default,
        "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
    }


CPU_KEYS = ("value", "unit", "cores", "kind", "sample", "rows_timed", "extrapolation_factor",
            "reference_default_threads", "omp_num_threads_env")


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.
    faiss-cpu cannot be installed here (no wheel, no network), so this is the oracle port of its
    algorithm -- the one other place bench.py may execute oracle/.  Each step is one search of a
    bounded row sample (sized so steps + warm-up fit ~200 s); `value` is scaled linearly to the full
    corpus, `ms_per_step` is the time of one SAMPLE step as measured."""
    if rank != 0:
        return
    total_rows, dim, dtype, batch, k = workload(args)
    warm = min(args.warmup, 1)
    r = cpu_oracle_qps(total_rows, dim, dtype, batch, k,
                       step_budget_s=max(2.0, 200.0 / (args.steps + warm)), steps=args.steps, warmup=warm)
    line = {
        "impl": "reference", "metric": metric_name(args, total_rows, dim, dtype, batch, k), "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step_sample"], "ms_per_step_scaled_to_full_corpus": r["ms_per_step_scaled"],
        "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {k_: r[k_] for k_ in CPU_KEYS},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def workload(args):
    total_rows, dim, dtype, batch, k = WORKLOADS[args.workload]
    if args.rows:
        total_rows = args.rows
    if args.batch:
        batch = args.batch
    if args.k:
        k = args.k
    return total_rows, dim, dtype, batch, k


def workload_config(args, world):
    total_rows, dim, dtype, batch, k = workload(args)
    return {
        "workload": (f"exact flat-IP search, synthetic {total_rows} x {dim} {dtype} corpus "
                     f"row-sharded over {world} GPU(s), batch {batch}, k {k}"
                     + (" (BASELINE configs[2])" if args.workload == "cfg3" and not args.rows else "")),
        "total_rows": total_rows, "dim": dim, "storage": dtype, "batch": batch, "k": k,
        "sharding": f"rows/{world}", "l2": "inputs larger than L2 (corpus shard >> 126 MB), no flush",
    }


def run_b200(args, rank, local_rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist

    from rag_inference_pipeline_b200.sharded import ShardedFlatIPSearcher
    from rag_inference_pipeline_b200.synthetic import CORPUS_SEED, synthetic_queries

    total_rows, dim, dtype, batch, k = workload(args)
    # stdout carries exactly ONE JSON line: anything libraries print while we set up and run (NCCL
    # prints its version on stdout) is diverted to stderr until the line is ready.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # The component a user of the reference talks to: registry.get("faiss_store").search(...)
    # (api.py:383-389).  One GPU: the store is built first and the device-resident timing loop
    # drives the very index it holds.  N GPUs (torchrun): the timing loop runs one rank per GPU;
    # the end-to-end leg afterwards runs the store in ONE process over all N devices.
    import types as _types

    from rag_inference_pipeline_b200 import B200FlatIPStore
    store_settings = _types.SimpleNamespace(
        faiss_index_path=f"synthetic:{total_rows}?dim={dim}&dtype={dtype}&seed={CORPUS_SEED}", faiss_dim=dim)
    gloo = dist.new_group(backend="gloo") if world > 1 else None   # host-only barriers for the e2e leg
    # Shard sizes follow what every GPU sustains (N > 1): a search ends with the slowest GPU, and
    # under the power cap the B200s of one box differ by up to 10 %.  Every rank measures its own
    # device while all the others are under the same load (`--balance 0` = equal shards).
    weights = None
    if world > 1 and args.balance:
        from rag_inference_pipeline_b200.group import calibrate_devices
        dist.barrier()
        rate = calibrate_devices([local_rank], dim, dtype, batch=batch, k=k, rows=2_000_000, seconds=1.5)[0]
        every = [None] * world
        dist.all_gather_object(every, rate)
        if all(r and r > 0 for r in every):
            weights = [float(r) for r in every]
    t0 = time.time()
    store = None
    if world == 1:
        store = B200FlatIPStore(store_settings, {"device": local_rank})
        store.load()
        searcher = ShardedFlatIPSearcher(dim, dtype, total_rows, rank=0, world_size=1, device=local_rank,
                                         index=store.native_index)
    else:
        searcher = ShardedFlatIPSearcher(dim, dtype, total_rows, rank=rank, world_size=world,
                                         device=local_rank, exchange=args.exchange, weights=weights)
        searcher.load_synthetic(CORPUS_SEED)
    load_s = time.time() - t0
    if args.variant != "auto":
        searcher.index.set_variant(args.variant)
    searcher.index.set_option("profile_events", 1)
    for kv in args.opt:
        key, val = kv.split("=")
        searcher.index.set_option(key, int(val))

    q_np = synthetic_queries(batch, dim)
    q_dev = torch.from_numpy(q_np).to(dev)
    q_pinned = torch.from_numpy(q_np).pin_memory()

    # ---- device-resident timing ("value") ----
    # search_tensor only ENQUEUES (local search + K4 exchange/merge): the verdict on the searches
    # (did any shard have to redo a query?) is read once, after the timed region, by check().
    for _ in range(max(args.warmup, 3)):
        D, I = searcher.search_tensor(q_dev, k)
    warm_status = searcher.check()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        D, I = searcher.search_tensor(q_dev, k)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    timed_status = searcher.check()   # synchronises; collects the per-launch CUDA-event times
    if warm_status or timed_status:
        # some search flagged a query: its result is only valid after a redo, which the timed loop
        # did not pay for -- say so instead of reporting a number for skipped work
        sys.exit(f"bench: a search was flagged (status warm={warm_status} timed={timed_status}); "
                 "the asynchronous timing loop is not valid for this input")
    launches = args.steps * (searcher.index.stat("launches") + (1 if world > 1 else 0))  # + exchange/merge kernel
    scan_ns = searcher.index.stat("scan_time_ns")
    scan_launches = searcher.index.stat("scan_launches")
    searches_timed = max(1, searcher.index.stat("scan_searches"))
    per_scan = {}   # scan launch index -> [rows, total ns, count]
    for i in range(64):
        ns = searcher.index.stat(f"scan_ns_{i}")
        if ns < 0:
            break
        per_scan[i] = [searcher.index.stat(f"scan_rows_{i}"), ns, searches_timed]

    tms = torch.tensor([ms, float(scan_ns)], dtype=torch.float64, device=dev)
    per_rank_scan_ms = [scan_ns / 1e6 / searches_timed]
    if world > 1:
        mine = torch.tensor([scan_ns / 1e6 / searches_timed], dtype=torch.float64, device=dev)
        every = torch.empty((world,), dtype=torch.float64, device=dev)
        dist.all_gather_into_tensor(every, mine)
        per_rank_scan_ms = [round(float(v), 3) for v in every.tolist()]
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max, scan_ns_max = (float(v) for v in tms.tolist())
    variant = searcher.index.last_variant()

    # ---- verification of the result the timed loop produced (rank 0 holds the GLOBAL result):
    #   * rows best-first, ids in range and distinct per row;
    #   * 9 returned scores recomputed in fp64 from regenerated rows;
    #   * corpus blocks of 4096 rows regenerated on the host (first, middle, last and
    #     every shard boundary region) and scored against ALL queries: no row outside the result
    #     may beat the k-th returned score (beyond the dtype's tolerance), and every block row
    #     that beats it by more than the tolerance must be IN the result;
    #   * sha256 of the id matrix and of the distance matrix, so the N = 1/2/4/8 lines of a scaling
    #     run can be compared with each other (rescoring is exact fp32 in a fixed order and the
    #     tie-break is fixed, so sharding must not change a single bit).
    verify = None
    if rank == 0:
        import hashlib

        from rag_inference_pipeline_b200.synthetic import synthetic_corpus
        Dh, Ih = D.cpu().numpy(), I.cpu().numpy()
        ok_sorted = bool(np.all(np.diff(Dh, axis=1) <= 0))
        ok_range = bool(Ih.min() >= 0 and Ih.max() < total_rows) if total_rows >= k else True
        ok_distinct = all(len(set(r.tolist())) == len(r) for r in Ih[:: max(1, batch // 64)]) if total_rows >= k else True
        errs = []
        for qi in (0, batch // 2, batch - 1):
            for j in (0, k // 2, k - 1):
                row = synthetic_corpus(int(Ih[qi, j]), 1, dim, dtype)[0].astype(np.float64)
                errs.append(abs(float(row @ q_np[qi].astype(np.float64)) - float(Dh[qi, j])))
        tol = 1e-5 if dtype == "f32" else 1e-3
        blk = min(4096, total_rows)
        starts = {0, max(0, total_rows // 2 - blk // 2), max(0, total_rows - blk)}
        from rag_inference_pipeline_b200.sharded import shard_range
        for r_ in range(1, world):      # rows on both sides of every shard boundary
            b_, _ = shard_range(total_rows, r_, world, weights)
            starts.add(max(0, min(total_rows - blk, b_ - blk // 2)))
        # ... plus 1024-row blocks at seeded random places, 64 blocks in all:
        # the structural blocks alone covered 0.01-0.04 % of the corpus and only its edges
        blocks = [(b0, blk) for b0 in sorted(starts)]
        small = min(1024, total_rows)
        pick = np.random.default_rng(20240607)
        while len(blocks) < 64 and total_rows > small:
            blocks.append((int(pick.integers(0, total_rows - small + 1)), small))
        missed = 0
        rows_checked = 0
        kth = Dh[:, -1].astype(np.float64)
        margin = tol * np.maximum(1.0, np.abs(kth))
        for b0, nb in blocks:
            xb = synthetic_corpus(b0, nb, dim, dtype).astype(np.float32)
            sc = (q_np @ xb.T).astype(np.float64)                 # [batch, nb]
            rows_checked += nb
            better = sc > (kth + margin)[:, None]
            for qi, ci in zip(*np.nonzero(better)):
                if (b0 + int(ci)) not in set(Ih[qi].tolist()):
                    missed += 1
        verify = {"sorted": ok_sorted, "ids_in_range": ok_range, "ids_distinct_per_row": ok_distinct,
                  "max_abs_err_of_9_recomputed_scores": max(errs),
                  "h5_blocks": len(blocks), "h5_rows_regenerated": rows_checked,
                  "h5_rows_missed": missed,
                  "ids_sha256": hashlib.sha256(np.ascontiguousarray(Ih).tobytes()).hexdigest(),
                  "distances_sha256": hashlib.sha256(np.ascontiguousarray(Dh).tobytes()).hexdigest(),
                  "ok": bool(ok_sorted and ok_range and ok_distinct and missed == 0
                             and max(errs) <= 10 * tol)}

    # ---- end-to-end ("e2e"): the same metric through the component's public call,
    #      B200FlatIPStore.search(numpy queries, k) -> numpy (distances, ids): host buffers, H2D and
    #      D2H inside the timed region, every step.  N > 1: the torchrun ranks release their shards,
    #      rank 0 loads the corpus into ONE store over all N devices (fip_group: a host thread +
    #      stream per device, K4 exchange + merge) and times it alone; the others wait on a
    #      host-side (gloo) barrier so that no NCCL kernel occupies an SM meanwhile.
    st = {"variant": searcher.index.last_variant(), "phases": searcher.index.stat("phases"),
          "overflow": searcher.index.stat("overflow_queries"),
          "uncertified": searcher.index.stat("uncertified_queries"),
          "cta_group": searcher.index.stat("cta_group"),
          "deferred": searcher.index.stat("deferred_epilogue"),
          "filter_f16": searcher.index.stat("filter_f16"), "local_rows": searcher.local_rows,
          "exchange": searcher.exchange_used}
    e2e = None
    if world > 1:
        searcher.close()
        searcher = None
        torch.cuda.empty_cache()
        dist.barrier(group=gloo)
        if rank == 0:
            store = B200FlatIPStore(store_settings, {"devices": list(range(world)),
                                                     "shard_weights": ",".join(repr(w) for w in weights) if weights else "equal"})
            store.load()
    if rank == 0:
        for kv in args.opt:
            key, val = kv.split("=")
            store.native_index.set_option(key, int(val))
        if args.variant != "auto":
            store.native_index.set_variant(args.variant)
        for _ in range(2):
            De, Ie = store.search(q_np, k)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            De, Ie = store.search(q_np, k)
        e2e_s = time.perf_counter() - t0
        e2e = {"value": batch / (e2e_s / args.steps), "unit": UNIT,
               "h2d_bytes_per_step": batch * dim * 4 * world, "d2h_bytes_per_step": batch * k * 12,
               "api": ("B200FlatIPStore.search (fip_search, host buffers)" if world == 1 else
                       f"B200FlatIPStore.search, devices 0..{world - 1} in ONE process (fip_group_search: "
                       "per-device H2D of the queries, local search, K4 exchange + merge, D2H from device 0)"),
               "ms_per_step": e2e_s / args.steps * 1e3,
               "result_equals_device_loop": bool(np.array_equal(Ie, I.cpu().numpy())
                                                 and np.array_equal(De, D.cpu().numpy()))}
        if world > 1:
            e2e["flagged_searches"] = store.native_index.stat("flagged_searches")
            store.unload()
            store = None
    if world > 1:
        dist.barrier(group=gloo)

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        r = cpu_oracle_qps(total_rows, dim, dtype, batch, k, step_budget_s=10.0)
        cpu_baseline = {k_: r[k_] for k_ in CPU_KEYS}

    if rank == 0:
        peaks = load_peaks()
        ms_per_step = ms_max / args.steps
        n_local = st["local_rows"]  # rank 0's shard (the largest)
        esize = 4 if dtype == "f32" else 2
        if dtype == "f32" and st["filter_f16"] == 1:
            esize = 2   # the tensor-core scan reads the fp16 FILTER COPY of the fp32 corpus
        scan_s = scan_ns_max * 1e-9
        scans = max(1, scan_launches)
        # dominant kernel = the scan launch that takes the longest (the last, geometrically
        # largest phase); its algorithmic work is that phase's rows only.
        dom = max(per_scan, key=lambda i: per_scan[i][1]) if per_scan else None
        dom_rows = per_scan[dom][0] if dom is not None else n_local
        dom_s = per_scan[dom][1] * 1e-9 / per_scan[dom][2] if dom is not None else scan_s / scans
        tensor_bound = variant == "tensor" and batch * (4 / esize) > 420   # 2*B/esize > P/BW ~ 210
        # Which measured tensor peak applies: a timed region of a second or more runs under the
        # board's power cap (~1.05-1.15 GHz) like the 4 s back-to-back cuBLAS loop the SUSTAINED
        # figure comes from; a short one (small corpus / small batch: tens of ms in all) runs at
        # the boost clock, and only the BURST figure is a ceiling for it (against the sustained
        # one such a run reads 1.0-1.1, which says nothing).
        long_region = ms_max >= 1000.0
        tensor_peak = peaks["tflops_sustained"] if long_region else peaks["tflops_burst"]
        if tensor_bound:
            flops = 2.0 * batch * dom_rows * dim            # algorithmic: no padding / split credit
            achieved = flops / dom_s / 1e12 if dom_s > 0 else None
            roof = {"bound": "tensor", "achieved": achieved, "peak": tensor_peak,
                    "unit": "TFLOP/s",
                    "peak_source": (f"{peaks['source']} (sustained cuBLAS bf16, kernel timed inside a long step)"
                                    if long_region else
                                    f"{peaks['source']} (burst cuBLAS bf16: the timed region lasted {ms_max:.0f} ms, "
                                    "too short for the power cap to bite)")}
        else:
            nbytes = float(dom_rows) * dim * esize + batch * dim * 4
            achieved = nbytes / dom_s / 1e9 if dom_s > 0 else None
            roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "peak_source": f"{peaks['source']} (copy bandwidth)"}
        roof["frac"] = (roof["achieved"] / roof["peak"]) if roof["achieved"] else None
        roof["kernel"] = "tc_scan_kernel" if variant == "tensor" else "stream_scan_kernel"
        roof["launch"] = {"phase": dom, "rows": dom_rows, "avg_ms": dom_s * 1e3,
                          "phases_per_search": len(per_scan)}
        roof["scan_bytes_per_element"] = esize
        roof["launches_timed"] = scan_launches
        roof["all_scan_launches_ms_per_step"] = scan_s * 1e3 / searches_timed
        roof["searches_timed"] = searches_timed
        # every rank's own scan time per search: the step ends with the SLOWEST GPU (boxes differ by
        # several per cent in the clock their power cap allows)
        roof["per_rank_scan_ms"] = per_rank_scan_ms
        roof["whole_search_frac"] = None
        if tensor_bound and ms_per_step > 0:
            roof["whole_search_frac"] = (2.0 * batch * n_local * dim / (ms_per_step * 1e-3) / 1e12
                                         / tensor_peak)
        elif ms_per_step > 0:
            roof["whole_search_frac"] = ((float(n_local) * dim * esize) / (ms_per_step * 1e-3) / 1e9
                                         / peaks["hbm_gbs"])
        roof["kernel_share_of_step"] = (scan_s * 1e3 / searches_timed) / ms_per_step if ms_per_step > 0 else None
        traffic_path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        roof["traffic"] = None
        if os.path.exists(traffic_path):
            with open(traffic_path) as f:
                roof["traffic"] = json.load(f).get(f"{args.workload}_gpus{world}")
        line = {
            "metric": metric_name(args, total_rows, dim, dtype, batch, k), "value": batch / (ms_per_step * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": {"f16": "f16", "bf16": "bf16", "f32": "f32"}[dtype],
            "data": "synthetic", "config": workload_config(args, world),
            "clocks": clocks,
            "e2e": e2e,
            "gpu_launches": launches,
            "roofline": roof,
            "cpu_baseline": cpu_baseline,
            "variant": variant, "phases_per_search": st["phases"], "overflow_queries": st["overflow"],
            "uncertified_queries": st["uncertified"], "cta_group": st["cta_group"],
            "deferred_epilogue": st["deferred"], "options": args.opt, "exchange": st["exchange"],
            "shard_weights": ([round(w / (sum(weights) / world), 4) for w in weights] if weights else "equal"),
            "corpus_load_s": load_s, "verify": verify,
        }
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if store is not None:
        store.unload()
    elif searcher is not None:
        searcher.close()
    if world > 1:
        dist.barrier(group=gloo)
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------
# Second operator: IVF-Flat (L2) at the reference's own index shape.  Not the
# driver's default line; same JSON contract, run with `--workload ivf`.
# ------------------------------------------------------------------------------------------
IVF_SHAPE = dict(rows=4_500_000, dim=768, nlist=4096, nprobe=64, batch=32, k=10, train=100_000)


def run_ivf(args):
    """`--workload ivf`: IndexIVFFlat(IndexFlatL2(768), 768, 4096, METRIC_L2), nprobe 64, over 4.5 M
    standard-normal rows (/root/reference/scripts/create_test_docs.py:12,83-104), batch 32, k 10
    (configs/retrieval_faiss_only.yaml: retrieval_batch_size / retrieval_k).  The quantizer is trained
    on 100 000 rows (the reference trains on 10 000, which on structure-free Gaussian data yields one
    545 k-row list that every query probes; `--ivf-train 10000` reproduces that).  value = QPS with
    device-resident queries (wall clock: the call blocks; it reads the plan's sizes back), e2e = the host API.
    roofline: HBM -- algorithmic bytes = the rows of every list probed by at least one query, read
    once (2 bytes per element when the fp16 filter copy is in use, else 4), over the device time of
    the scan launches.  cpu_baseline / --impl reference: the C
    restatement of faiss's IVF search (oracle/faiss_ivf_restatement.c) on the same index and queries."""
    import numpy as np
    import torch

    from rag_inference_pipeline_b200 import IVFFlatL2Index

    sh = dict(IVF_SHAPE)
    if args.rows:
        sh["rows"] = args.rows
    if args.batch:
        sh["batch"] = args.batch
    if args.k:
        sh["k"] = args.k
    if args.ivf_train:
        sh["train"] = args.ivf_train
    n, dim, nlist, nprobe, B, k = sh["rows"], sh["dim"], sh["nlist"], sh["nprobe"], sh["batch"], sh["k"]
    metric = f"QPS ({n} x {dim} fp32 IVF-Flat L2, nlist {nlist}, nprobe {nprobe}, k={k}, batch {B})"
    config = {"workload": f"IVF-Flat (L2) search, standard-normal {n} x {dim} fp32 corpus, nlist {nlist} "
                          f"(trained on {sh['train']} rows), nprobe {nprobe}, batch {B}, k {k}, 1 GPU",
              "l2": "probed lists >> L2 (126 MB), no flush"}
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    rng = np.random.default_rng(0)
    reference = args.impl == "reference"
    keep_host = reference or not args.no_cpu_baseline
    t0 = time.time()
    first = rng.standard_normal((min(n, 100_000), dim), dtype=np.float32)
    # (the reference arm too builds the index with the GPU library, then times the CPU port on it)
    index = IVFFlatL2Index(dim, nlist, nprobe=nprobe)
    index.train(first[:sh["train"]] if sh["train"] <= len(first) else
                np.concatenate([first, rng.standard_normal((sh["train"] - len(first), dim), dtype=np.float32)]))
    index.add(first)
    done = len(first)
    while done < n:
        m = min(100_000, n - done)
        index.add(rng.standard_normal((m, dim), dtype=np.float32))
        done += m
    sizes = index.list_sizes()
    build_s = time.time() - t0
    q = (first[rng.integers(0, min(len(first), 4096), B)] + 0.3 * rng.standard_normal((B, dim), dtype=np.float32)).astype(np.float32)

    def cpu_leg(steps, warmup):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import c_oracle
        szs, codes, ids = index.export_lists()
        cent = index.centroids()
        for _ in range(max(1, warmup)):
            c_oracle.ivf_search_l2(q, cent, szs, codes, ids, k, nprobe)
        ts = []
        for _ in range(max(1, steps)):
            t = time.perf_counter()
            Dc, Ic = c_oracle.ivf_search_l2(q, cent, szs, codes, ids, k, nprobe)
            ts.append(time.perf_counter() - t)
        t = sum(ts) / len(ts)
        return {"value": B / t, "unit": UNIT, "cores": c_oracle.max_threads(), "kind": "port",
                "sample": (f"oracle/faiss_ivf_restatement.c (faiss-cpu IVF algorithm: OpenMP over queries, "
                           f"coarse probe + list scan + k-heap) on the full index and the same {B} queries; "
                           f"{t * 1e3:.1f} ms per search; host cpu_count={os.cpu_count()}"),
                "ms": t * 1e3}, (Dc, Ic)

    if reference:
        cb, _ = cpu_leg(args.steps, min(args.warmup, 1))
        line = {"impl": "reference", "metric": metric, "value": cb["value"], "unit": UNIT, "n_gpus": 1,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {k_: cb[k_] for k_ in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        return

    index.set_option("profile_events", 1)
    index.set_option("coarse_tensor", args.ivf_coarse)
    qd = torch.from_numpy(q).cuda()
    D = torch.empty((B, k), dtype=torch.float32, device="cuda")
    I = torch.empty((B, k), dtype=torch.int64, device="cuda")
    warm = max(args.warmup, 3)
    for _ in range(warm):
        index.search_device(qd.data_ptr(), B, k, D.data_ptr(), I.data_ptr())
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    launches = scan_ns = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        index.search_device(qd.data_ptr(), B, k, D.data_ptr(), I.data_ptr())
        launches += index.stat("launches")
        scan_ns += index.stat("scan_ns")
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    t0 = time.perf_counter()
    for _ in range(args.steps):
        Dh, Ih = index.search(q, k)
    dt_e2e = (time.perf_counter() - t0) / args.steps
    clocks = sampler.stop()
    peaks = load_peaks()
    distinct = index.stat("distinct_rows")
    # the tensor-core filter reads the fp16 copy of the lists when it is in use (stat "tensor_f16")
    ebytes = 2 if index.stat("tensor_f16") else 4
    alg_bytes = distinct * dim * ebytes + nlist * dim * 4 + B * dim * 4 + B * k * 12
    scan_s = scan_ns / args.steps * 1e-9
    roof = {"bound": "hbm", "achieved": alg_bytes / scan_s / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "peak_source": f"{peaks['source']} (copy bandwidth)", "kernel": "ivf_scan_kernel",
            "launch": {"scan_ms_per_search": scan_s * 1e3, "distinct_probed_rows": distinct,
                       "list_bytes_per_element": ebytes,
                       "rows_scanned": index.stat("rows_scanned"), "pairs_scored": index.stat("pairs_scored"),
                       "tiles": index.stat("tiles"), "tiles16": index.stat("tiles16")},
            "kernel_share_of_step": scan_s / dt, "whole_search_frac": alg_bytes / dt / 1e9 / peaks["hbm_gbs"],
            "traffic": None}
    roof["frac"] = roof["achieved"] / roof["peak"]
    cb = None
    verify = None
    if keep_host:
        cb, (Dc, Ic) = cpu_leg(1, 1)
        verify = {"ids_equal_cpu_port": float((Ih == Ic).mean()),
                  "max_rel_dist_err_vs_cpu_port": float(np.max(np.abs(Dh - Dc) / np.maximum(Dc, 1e-6)))}
        cb = {k_: cb[k_] for k_ in ("value", "unit", "cores", "kind", "sample")}
    line = {"metric": metric, "value": B / dt, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": warm,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config, "clocks": clocks,
            "e2e": {"value": B / dt_e2e, "unit": UNIT, "h2d_bytes_per_step": B * dim * 4,
                    "d2h_bytes_per_step": B * k * 12, "api": "IVFFlatL2Index.search (fivf_search, host buffers)"},
            "gpu_launches": launches, "roofline": roof, "cpu_baseline": cb,
            "timing": "wall clock around the blocking search call (planning and both synchronisations included)",
            "list_sizes": {"min": int(sizes.min()), "mean": float(sizes.mean()), "max": int(sizes.max())},
            "index_build_s": build_s, "verify": verify,
            "coarse": {"tensor": index.stat("coarse_tensor")}}
    index.close()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=sorted(WORKLOADS) + ["ivf"], default="cfg3")
    ap.add_argument("--ivf-train", type=int, default=0, help="--workload ivf: rows the quantizer is trained on")
    ap.add_argument("--ivf-coarse", type=int, default=1, help="--workload ivf: 1 = coarse quantizer on the tensor cores, 0 = FFMA")
    ap.add_argument("--rows", type=int, default=0, help="override total corpus rows")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--variant", choices=["auto", "stream", "tensor"], default="auto")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--balance", type=int, default=1,
                    help="N > 1: 1 = shard sizes proportional to measured per-GPU throughput, 0 = equal shards")
    ap.add_argument("--opt", action="append", default=[], help="index option key=value (A/B runs)")
    ap.add_argument("--exchange", choices=["auto", "p2p", "nccl"], default="auto",
                    help="N > 1: fused peer-memory exchange + merge kernel, or NCCL all-gather + merge")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world != args.gpus and world > 1:
        args.gpus = world
    if args.workload == "ivf":
        if rank == 0:
            run_ivf(args)
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.gpus > 1 and world == 1:
        sys.exit("for --gpus N > 1 launch with: python -m torch.distributed.run --nnodes=1 "
                 "--nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...")
    run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
