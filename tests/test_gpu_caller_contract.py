"""The reference's ONLY caller of the path, driven against the CUDA store on the GPU box.

FALLBACK for a box without the staged reference: tests/test_gpu_reference_caller.py runs the
reference's real RetrievalExecutor / BatchScheduler / ComponentRegistry against the CUDA store
(from the untracked copy under baseline/_ref/).  This file keeps the caller's contract restated by
hand from ``/root/reference/src/pipeline/services/retrieval/api.py``, so the same properties are
still checked where that copy is absent:

  * ``:383``      the store is looked up as ``registry.get("faiss_store")`` (factory.install);
  * ``:348-349``  ``_process_batch_sync`` runs on ``loop.run_in_executor(None, ...)`` threads, several
                  batches at once;
  * ``:387-390``  cache off: ``distances, indices = store.search(embeddings, retrieval_k)`` and every
                  row is consumed through ``.tolist()`` (Python ints / floats);
  * ``:392-425``  cache on: rows are keyed by ``sha256(embedding.tobytes())``, only the misses are
                  searched (as ONE ``np.array(missing)`` batch) and written back.

The store under test is the single-device one and the sharded one (``devices``) -- the caller
cannot tell them apart."""
import hashlib
import threading
import types
from collections import OrderedDict
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

import flatip_oracle
from rag_inference_pipeline_b200 import synthetic

pytestmark = pytest.mark.gpu


class _Registry:
    """The two lines of ComponentRegistry the caller relies on (component_registry.py:80-83)."""

    def __init__(self):
        self._c = {}

    def register(self, name, component):
        component.load()            # runtime_factory.py:215-232: load() inside register
        self._c[name] = component

    def get(self, name):
        return self._c[name]


class _Caller:
    """Restatement of RetrievalExecutor._search_faiss_with_cache (api.py:376-425)."""

    def __init__(self, registry, k, use_cache, capacity=4096):
        self.registry, self.k, self.use_cache = registry, k, use_cache
        self.cache = OrderedDict()
        self.capacity = capacity
        self._lock = threading.Lock()
        self.searched_rows = 0

    def search(self, embeddings):
        store = self.registry.get("faiss_store")
        n = len(embeddings)
        ids = [[] for _ in range(n)]
        dists = [[] for _ in range(n)]
        if not self.use_cache:
            D, I = store.search(embeddings, self.k)
            self.searched_rows += n
            return [r.tolist() for r in I], [r.tolist() for r in D]
        miss_i, miss_e = [], []
        for i in range(n):
            key = hashlib.sha256(embeddings[i].tobytes()).hexdigest()
            with self._lock:
                hit = self.cache.get(key)
            if hit:
                ids[i], dists[i] = hit
            else:
                miss_i.append(i)
                miss_e.append(embeddings[i])
        if miss_e:
            D, I = store.search(np.array(miss_e), self.k)
            with self._lock:
                self.searched_rows += len(miss_e)
            for j, i in enumerate(miss_i):
                ids[i], dists[i] = I[j].tolist(), D[j].tolist()
                key = hashlib.sha256(miss_e[j].tobytes()).hexdigest()
                with self._lock:
                    self.cache[key] = (ids[i], dists[i])
                    while len(self.cache) > self.capacity:
                        self.cache.popitem(last=False)
        return ids, dists


@pytest.mark.parametrize("devices", [None, [0, 0, 0]])
@pytest.mark.parametrize("use_cache", [False, True])
def test_caller_contract_concurrent_batches(monkeypatch, devices, use_cache):
    from rag_inference_pipeline_b200 import B200FlatIPStore, factory

    monkeypatch.setenv("FLATIP_B200_ALLOW_SHARED_DEVICE", "1")
    dim, n, k, batch = 768, 50_000, 10, 32      # retrieval_faiss_only.yaml: batch 32, k 10
    settings = types.SimpleNamespace(faiss_index_path=f"synthetic:{n}?dim={dim}&dtype=f32", faiss_dim=dim)
    factories = {}
    factory.install(factories)                   # component_factory.py:26-27 -> COMPONENT_FACTORIES
    store = factories["faiss_store"](settings, {"devices": devices} if devices else {})
    assert isinstance(store, B200FlatIPStore)
    reg = _Registry()
    reg.register("faiss_store", store)
    caller = _Caller(reg, k, use_cache)
    q_all = synthetic.synthetic_queries(8 * batch, dim)
    # batches overlap (requests repeat across users): batch b = queries [16 b, 16 b + 32)
    batches = [q_all[16 * b:16 * b + batch] for b in range(14)]
    with ThreadPoolExecutor(max_workers=8) as pool:    # default loop executor threads
        results = list(pool.map(caller.search, batches))
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32")
    Dr, Ir = flatip_oracle.flat_ip_search(q_all, corpus, k + 1)
    for b, (ids, dists) in enumerate(results):
        assert len(ids) == batch and all(type(v) is int for v in ids[0]) and all(type(v) is float for v in dists[0])
        rows = slice(16 * b, 16 * b + batch)
        res = flatip_oracle.compare_with_oracle(np.array(dists, np.float32), np.array(ids, np.int64),
                                                Dr[rows, :k], Ir[rows, :k], rtol=1e-5, kplus1_ref=Dr[rows, k])
        assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, (b, res)
    if use_cache:
        again = caller.search(batches[3])             # everything cached by now
        assert again == results[3] and caller.searched_rows <= 14 * batch
    store.unload()
    assert not store.is_loaded
