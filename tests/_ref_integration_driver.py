"""Drives the REFERENCE's own RetrievalExecutor + BatchScheduler + ComponentRegistry + component
factory (imported from /root/reference/src, read-only) against B200FlatIPStore, to prove the
drop-in boundary: the store is created by the reference's `create_component`, registered under
"faiss_store", load()ed by the registry's load hook and searched from the executor's worker thread.

Runs in its own process because the reference captures `settings` at import time and several of
its third-party imports (faiss, lz4, orjson, zstandard, sentence_transformers, opentelemetry
instrumentation) are absent from this image and have to be stubbed in sys.modules first
(SURVEY.md Appendix C.3).  No GPU in the build container: with --backend oracle the store's native
index is replaced by an oracle-backed stand-in (TEST ONLY); with --backend native the real CUDA
library is used (on the B200 box the reference comes from baseline/_ref/, see REF_SRC).
--devices 0,0 puts the corpus behind a two-shard group; --cache on runs the reference's LRU path
(api.py:392-425) and sends every request twice, so the second round is answered from its cache.
"""
import argparse
import asyncio
import json
import os
import sys
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
# The reference's sources: where they are mounted in the build container, else the untracked,
# read-only copy __graft_entry__.build() stages under baseline/_ref/ (git-ignored; it travels to the
# GPU box with the snapshot like the built .so does).
REF_SRC = next((p for p in ("/root/reference/src", os.path.join(ROOT, "baseline", "_ref", "src"))
                if os.path.isdir(os.path.join(p, "pipeline"))), "/root/reference/src")


def install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    mod("faiss", IO_FLAG_MMAP=8, read_index=lambda *a, **k: None, Index=object,
        omp_set_num_threads=lambda n: None)
    lz4 = mod("lz4")
    lz4.frame = mod("lz4.frame", compress=lambda b, **k: b, decompress=lambda b, **k: b)
    mod("orjson", dumps=lambda o, **k: json.dumps(o).encode(), loads=json.loads,
        OPT_SERIALIZE_NUMPY=0, OPT_NON_STR_KEYS=0)

    class _Z:
        def __init__(self, *a, **k):
            pass

        def compress(self, b):
            return b

        def decompress(self, b, *a, **k):
            return b

    mod("zstandard", ZstdCompressor=_Z, ZstdDecompressor=_Z)
    mod("sentence_transformers", SentenceTransformer=object)
    try:
        import opentelemetry.instrumentation  # noqa: F401
    except Exception:
        mod("opentelemetry.instrumentation")

    class _Inst:
        @staticmethod
        def instrument_app(*a, **k):
            pass

        def instrument(self, *a, **k):
            pass

    for name, cls in (("fastapi", "FastAPIInstrumentor"), ("httpx", "HTTPXClientInstrumentor")):
        try:
            __import__(f"opentelemetry.instrumentation.{name}")
        except Exception:
            mod(f"opentelemetry.instrumentation.{name}", **{cls: _Inst})


class OracleBackedIndex:
    """TEST-ONLY stand-in for FlatIPIndex (no GPU in the build container)."""

    def __init__(self, rows, dtype="f32"):
        self.rows, self.d, self.dtype = rows, rows.shape[1], dtype

    @property
    def ntotal(self):
        return self.rows.shape[0]

    def search(self, x, k):
        import flatip_oracle
        return flatip_oracle.flat_ip_search(x, self.rows, k)

    def set_variant(self, name):
        pass

    def close(self):
        pass


async def run(args):
    import numpy as np

    from rag_inference_pipeline_b200 import B200FlatIPStore, index_io, install, synthetic
    import flatip_oracle

    dim, n, k, nreq = args.dim, args.rows, args.k, args.requests
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32")
    tmp = tempfile.mkdtemp()
    path = os.path.join(tmp, "faiss_index.bin")
    index_io.write_faiss_flat(path, corpus)

    os.environ.update({"NODE_NUMBER": "1", "FAISS_INDEX_PATH": path, "faiss_dim": str(dim),
                       "retrieval_k": str(k), "retrieval_batch_size": str(args.batch),
                       "retrieval_max_batch_delay_ms": "20",
                       "DISABLE_CACHE_FOR_PROFILING": "false" if args.cache == "on" else "true"})
    devices = [int(v) for v in args.devices.split(",")] if args.devices not in ("", "none") else None
    if devices and len(set(devices)) < len(devices):
        os.environ["FLATIP_B200_ALLOW_SHARED_DEVICE"] = "1"   # one-GPU box: both shards on cuda:0
    install_stubs()
    sys.path.insert(0, REF_SRC)
    from pipeline import component_factory
    from pipeline.component_registry import ComponentRegistry
    from pipeline.config import get_settings
    from pipeline.enums import ComponentType
    from pipeline.services.retrieval.api import RetrievalExecutor
    from pipeline.services.retrieval.schemas import RetrievalRequestItem

    settings = get_settings()
    assert settings.faiss_dim == dim and settings.retrieval_k == k
    if args.backend == "oracle":
        B200FlatIPStore._load_file = lambda self: OracleBackedIndex(
            np.asarray(index_io.open_rows(str(self.index_path),
                                          index_io.read_header(str(self.index_path)))))
    # 1. the reference's factory map now builds the B200 store for type "faiss"
    install(component_factory.COMPONENT_FACTORIES)
    assert bool(getattr(settings, "disable_cache_for_profiling", True)) == (args.cache != "on")
    store = component_factory.create_component(ComponentType.FAISS, settings,
                                               {"devices": devices} if devices else {})
    assert isinstance(store, B200FlatIPStore)
    # 2. the reference's registry runs load() through its load hook
    registry = ComponentRegistry()
    registry.register("faiss_store", store, load_hook=store.load, unload_hook=store.unload)
    assert store.is_loaded and store.index_size == n
    # 3. the reference's executor + batch scheduler call search() from a worker thread
    executor = RetrievalExecutor(registry)
    await executor.start()
    queries = synthetic.synthetic_queries(nreq, dim)
    items = [RetrievalRequestItem(request_id=f"r{i}", query="q", embedding=queries[i].tolist())
             for i in range(nreq)]
    calls = []
    real_search = store.search

    def counting_search(x, kk):
        calls.append(int(len(x)))
        return real_search(x, kk)

    store.search = counting_search
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k)
    ids_ok, order_ok, max_err, total = True, True, 0.0, 0
    rounds = 2 if args.cache == "on" else 1
    searched = []
    for _ in range(rounds):
        before = sum(calls)
        responses = await asyncio.gather(*[executor.process_request(it) for it in items])
        searched.append(sum(calls) - before)
        total += len(responses)
        ids_ok &= all([d.doc_id for d in r.docs] == Ir[i].tolist() for i, r in enumerate(responses))
        # north_star's fp32 rule: 1e-5 relative (to max(1, |score|)); reported in those units x 1e-5 .. 1e-4
        max_err = max(max_err, max(abs(d.score - float(Dr[i, j])) / max(1.0, abs(float(Dr[i, j]))) * 10.0
                                   for i, r in enumerate(responses) for j, d in enumerate(r.docs)))
        order_ok &= all(r.request_id == f"r{i}" for i, r in enumerate(responses))
    await executor.stop()
    # cache on: the second round must have been answered by the reference's LRU, not by a search
    cache_ok = args.cache != "on" or (searched[0] == nreq and searched[1] == 0)
    registry.unload_all()
    print(json.dumps({"ok": bool(ids_ok and order_ok and cache_ok and max_err < 1e-4), "ids_ok": bool(ids_ok),
                      "max_score_err": max_err, "responses": total, "search_calls": len(calls),
                      "largest_batch": max(calls), "queries_searched_per_round": searched,
                      "docs_per_response": len(responses[0].docs), "unloaded": not store.is_loaded,
                      "backend": args.backend, "devices": devices, "cache": args.cache,
                      "store": type(store).__name__, "group": bool(devices and len(devices) > 1),
                      "reference_src": REF_SRC}))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", choices=["oracle", "native"], default="oracle")
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--rows", type=int, default=2000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--requests", type=int, default=70)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--devices", default="none", help="'none' or a device list such as 0,0 / 0,1")
    ap.add_argument("--cache", choices=["off", "on"], default="off")
    asyncio.run(run(ap.parse_args()))
