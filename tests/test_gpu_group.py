"""GPU tests (-m gpu) of the round-2 additions behind the C ABI:
  * fip_group_*  -- one corpus row-sharded over several devices INSIDE one process, behind one search
    call and behind B200FlatIPStore (the reference's registry.get("faiss_store").search call site,
    /root/reference/src/pipeline/services/retrieval/api.py:383-389).  On a single-GPU box the
    shards share device 0 (FLATIP_B200_ALLOW_SHARED_DEVICE): routing, id bases, K4, status words and
    the second exchange after a redo are all exercised; with >= 2 GPUs the real layout runs too.
  * fip_search_device_async / fip_search_status -- searches without a host round trip.
  * fip_add_host -- 16-bit corpus rows uploaded as 16 bits.
  * K4 (exchange_merge_kernel) beyond what is resident at once: nq = 4096, k = 1024, 4 ranks.
Results are compared with the single-index path bit for bit and with the CPU oracle."""
import os
import threading
import types

import numpy as np
import pytest

import c_oracle
import flatip_oracle
from flatip_oracle import SCORE_RTOL
from rag_inference_pipeline_b200 import synthetic

pytestmark = pytest.mark.gpu


@pytest.fixture()
def shared_device(monkeypatch):
    monkeypatch.setenv("FLATIP_B200_ALLOW_SHARED_DEVICE", "1")


def _single(dim, dtype, n, q, k, id_base=0):
    from rag_inference_pipeline_b200 import FlatIPIndex
    idx = FlatIPIndex(dim, dtype, 0, id_base=id_base, capacity_rows=n)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, id_base)
    out = idx.search(q, k)
    idx.close()
    return out


@pytest.mark.parametrize("dtype,n,dim,nq,k,ndev", [
    ("f16", 60_001, 256, 130, 100, 3),     # ragged last shard, two query tiles
    ("bf16", 40_000, 128, 1, 10, 2),       # single query (streaming kernel on every shard)
    ("f32", 30_000, 96, 33, 7, 4),         # fp32 storage (fp16 filter copy per shard)
    ("f16", 900, 64, 5, 400, 4),           # k larger than a shard: -1 padding inside shard results
])
def test_group_on_shared_device_equals_single_index(shared_device, dtype, n, dim, nq, k, ndev):
    from rag_inference_pipeline_b200.group import FlatIPGroup
    q = synthetic.synthetic_queries(nq, dim)
    base = 1_000_000
    Ds, Is = _single(dim, dtype, n, q, k, id_base=base)
    g = FlatIPGroup(dim, dtype, [0] * ndev, n, id_base=base)
    g.add_synthetic(n, synthetic.CORPUS_SEED)
    assert g.ntotal == n and g.stat("devices") == ndev
    per = -(-n // ndev)
    assert [g.shard(r).ntotal for r in range(ndev)] == [max(0, min(per, n - r * per)) for r in range(ndev)]
    D, I = g.search(q, k)
    assert np.array_equal(I, Is) and np.array_equal(D, Ds)        # bit for bit
    D2, I2 = g.search(q, k)                                        # second epoch (other buffer parity)
    assert np.array_equal(I2, Is) and np.array_equal(D2, Ds)
    assert g.stat("flagged_searches") == 0 and g.stat("exchange_launches") == 2 * ndev
    corpus = c_oracle.synthetic_corpus(base, n, dim, dtype)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k, id_base=base)
    res = flatip_oracle.compare_with_oracle(D, I, Dr, Ir, rtol=SCORE_RTOL[dtype])
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    g.close()


def test_group_host_rows_in_16_bits_and_growth_beyond_plan(shared_device):
    """Host rows reach the shards in their own width (fp16 as fp16, bf16 bit patterns as uint16),
    added in several calls that straddle shard boundaries; rows beyond the plan go to the last
    shard.  Same result as ONE index fed the same rows as fp32."""
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.group import FlatIPGroup
    dim, n, nq, k = 128, 25_000, 40, 20
    q = synthetic.synthetic_queries(nq, dim)
    for dtype in ("f16", "bf16"):
        rows32 = synthetic.synthetic_corpus(0, n, dim, dtype)           # values representable in dtype
        raw = rows32.astype(np.float16) if dtype == "f16" else synthetic.round_to_bf16_bits(rows32)
        ref = FlatIPIndex(dim, dtype, 0)
        ref.add(rows32)
        Ds, Is = ref.search(q, k)
        one = FlatIPIndex(dim, dtype, 0)
        one.add_raw(raw[:7000], dtype)
        one.add_raw(raw[7000:], dtype)
        assert np.array_equal(one.reconstruct_n(0, n), rows32)
        D1, I1 = one.search(q, k)
        assert np.array_equal(I1, Is) and np.array_equal(D1, Ds)
        g = FlatIPGroup(dim, dtype, [0, 0, 0], 21_000)                   # plan: 3 x 7000 rows
        for a, b in ((0, 5000), (5000, 15_500), (15_500, n)):
            g.add_raw(raw[a:b], dtype)
        assert [g.shard(r).ntotal for r in range(3)] == [7000, 7000, n - 14_000]
        D, I = g.search(q, k)
        assert np.array_equal(I, Is) and np.array_equal(D, Ds)
        f32 = FlatIPIndex(dim, "f32", 0)                                 # 16-bit rows into fp32 storage
        f32.add_raw(raw, dtype)
        assert np.array_equal(f32.reconstruct_n(0, n), rows32)
        with pytest.raises(Exception):
            one.add_raw(raw.view(np.uint16) if dtype == "f16" else raw.view(np.float16), "bf16" if dtype == "f16" else "f16")
        for x in (ref, one, g, f32):
            x.close()


def test_group_redoes_flagged_queries_and_exchanges_again(shared_device):
    """A near-tie cluster that the tensor-core filter cannot certify sits in ONE shard: the group
    learns it from the status words K4 ships, the shard redoes the query, the exchange runs a
    second time, and the result equals the single index's (and the oracle's)."""
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.group import FlatIPGroup
    dim, n, k, dtype = 128, 30_000, 10, "f16"
    rng = np.random.default_rng(11)
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32").copy()
    queries = synthetic.synthetic_queries(40, dim)
    where = 20_000 + rng.choice(9_000, size=300, replace=False)        # all inside shard 2 of 3
    corpus[where] = 9.0 * queries[0][None, :] + 2e-5 * rng.standard_normal((300, dim)).astype(np.float32)
    corpus = synthetic.round_to_storage(corpus, dtype)
    one = FlatIPIndex(dim, dtype, 0)
    one.add(corpus)
    one.set_variant("tensor")
    Ds, Is = one.search(queries, k)
    assert one.stat("uncertified_queries") >= 1
    g = FlatIPGroup(dim, dtype, [0, 0, 0], n)
    g.add(corpus)
    g.set_variant("tensor")
    D, I = g.search(queries, k)
    assert g.stat("flagged_searches") == 1 and g.stat("exchange_launches") == 6
    assert g.shard(2).stat("uncertified_queries") >= 1 and g.shard(0).stat("uncertified_queries") == 0
    assert np.array_equal(I, Is) and np.array_equal(D, Ds)
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k + 1)
    res = flatip_oracle.compare_with_oracle(D, I, Dr[:, :k], Ir[:, :k], rtol=1e-3, kplus1_ref=Dr[:, k])
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    one.close()
    g.close()


def test_group_survives_a_rank_that_fails_before_the_exchange(shared_device):
    """One shard fails its local part (injected; in production: out of memory, a bad launch).  No
    rank may then launch the exchange kernel -- the others would wait 30 s for the missing rows and
    trap -- so the call must come back at once with the error, and because the exchange epochs stay
    in step the NEXT search of the same group is correct."""
    import time
    from rag_inference_pipeline_b200._native import FlatIPError
    from rag_inference_pipeline_b200.group import FlatIPGroup
    dim, n, nq, k, dtype = 128, 40_000, 20, 10, "f16"
    q = synthetic.synthetic_queries(nq, dim)
    corpus = synthetic.synthetic_corpus(0, n, dim, dtype)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    g = FlatIPGroup(dim, dtype, [0, 0, 0], n)
    # an add that fails on ONE shard leaves the group as it was (no shard keeps its share)
    g.set_option("debug_fail_rank", 2)
    with pytest.raises(FlatIPError, match="group unchanged"):
        g.add_synthetic(n, synthetic.CORPUS_SEED)
    assert g.ntotal == 0 and [g.shard(r).ntotal for r in range(3)] == [0, 0, 0]
    g.add_synthetic(n, synthetic.CORPUS_SEED)
    assert g.ntotal == n
    D, I = g.search(q, k)
    for failing in (1, 0, 2):
        g.set_option("debug_fail_rank", failing + 1)
        t0 = time.time()
        with pytest.raises(FlatIPError, match="injected failure"):
            g.search(q, k)
        assert time.time() - t0 < 5.0
        D2, I2 = g.search(q, k)
        assert np.array_equal(I2, I) and np.array_equal(D2, D)
    res = flatip_oracle.compare_with_oracle(D, I, Dr, Ir, rtol=SCORE_RTOL[dtype])
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    g.close()


def test_group_coalesces_concurrent_callers(shared_device):
    from rag_inference_pipeline_b200.group import FlatIPGroup
    dim, n, k = 64, 20_000, 5
    g = FlatIPGroup(dim, "f16", [0, 0], n)
    g.add_synthetic(n, synthetic.CORPUS_SEED)
    q = synthetic.synthetic_queries(64, dim)
    Dall, Iall = g.search(q, k)
    out = [None] * 8
    def work(t):
        out[t] = g.search(q[8 * t:8 * t + 8], k)
    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for t in range(8):
        assert np.array_equal(out[t][1], Iall[8 * t:8 * t + 8]) and np.array_equal(out[t][0], Dall[8 * t:8 * t + 8])
    assert g.stat("host_calls") == 9 and g.stat("host_batches") <= 9
    g.close()


def test_store_with_devices_config(shared_device, tmp_path):
    """B200FlatIPStore(config={"devices": [...]}) -- the component the registry hands to the
    reference's executor -- serves the same answers from several shards: synthetic and FIPB (bf16
    file, uploaded as 16 bits) corpora, reference error messages unchanged."""
    from rag_inference_pipeline_b200 import B200FlatIPStore, index_io
    dim, n, k = 128, 12_345, 10
    q = synthetic.synthetic_queries(32, dim)
    s1 = types.SimpleNamespace(faiss_index_path=f"synthetic:{n}?dim={dim}&dtype=f16", faiss_dim=dim)
    one = B200FlatIPStore(s1)
    one.load()
    many = B200FlatIPStore(s1, {"devices": [0, 0, 0]})
    many.load()
    assert many.is_loaded and many.index_size == n and "loaded" in repr(many)
    Da, Ia = one.search(q, k)
    Db, Ib = many.search(q, k)
    assert np.array_equal(Ia, Ib) and np.array_equal(Da, Db)
    with pytest.raises(ValueError, match="Embedding dimension mismatch"):
        many.search(np.zeros((2, dim + 1), np.float32), k)
    many.unload()
    one.unload()
    rows = synthetic.synthetic_corpus(0, n, dim, "bf16")
    path = tmp_path / "corpus.fipb"
    index_io.write_fipb(str(path), rows, "bf16", id_base=500)
    s2 = types.SimpleNamespace(faiss_index_path=str(path), faiss_dim=dim)
    a = B200FlatIPStore(s2)
    a.load()
    b = B200FlatIPStore(s2, {"devices": "0,0"})
    b.load()
    Da, Ia = a.search(q, k)
    Db, Ib = b.search(q, k)
    assert Ia.min() >= 500 and np.array_equal(Ia, Ib) and np.array_equal(Da, Db)
    Dr, Ir = flatip_oracle.flat_ip_search(q, rows, k, id_base=500)
    res = flatip_oracle.compare_with_oracle(Db, Ib, Dr, Ir, rtol=1e-3)
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    a.unload()
    b.unload()


def test_group_on_real_devices_equals_single_device():
    """>= 2 GPUs: a 2-device (and all-device) store equals the 1-device store bit for bit."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    from rag_inference_pipeline_b200 import B200FlatIPStore
    dim, n, k = 768, 400_000, 100
    q = synthetic.synthetic_queries(256, dim)
    s = types.SimpleNamespace(faiss_index_path=f"synthetic:{n}?dim={dim}&dtype=f16", faiss_dim=dim)
    one = B200FlatIPStore(s)
    one.load()
    Da, Ia = one.search(q, k)
    one.unload()
    for devs in ([0, 1], list(range(torch.cuda.device_count()))):
        st = B200FlatIPStore(s, {"devices": devs})
        st.load()
        for _ in range(3):
            Db, Ib = st.search(q, k)
            assert np.array_equal(Ia, Ib) and np.array_equal(Da, Db), devs
        st.unload()
    corpus = c_oracle.synthetic_corpus(0, n, dim, "f16")
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    res = flatip_oracle.compare_with_oracle(Da, Ia, Dr, Ir, rtol=1e-3)
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res


def test_async_search_and_status():
    """fip_search_device_async enqueues only; fip_search_status reports (and clears) whether any
    asynchronous search since the last call was flagged -- also when later searches were clean."""
    import torch
    from rag_inference_pipeline_b200 import FlatIPIndex
    dim, n, k = 128, 30_000, 10
    rng = np.random.default_rng(11)
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32").copy()
    queries = synthetic.synthetic_queries(40, dim)
    where = rng.choice(n, size=300, replace=False)
    corpus[where] = 9.0 * queries[0][None, :] + 2e-5 * rng.standard_normal((300, dim)).astype(np.float32)
    corpus = synthetic.round_to_storage(corpus, "f16")
    idx = FlatIPIndex(dim, "f16", 0)
    idx.add(corpus)
    idx.set_variant("tensor")
    idx.set_option("profile_events", 1)
    from rag_inference_pipeline_b200.index import torch_stream_handle
    s = torch_stream_handle()
    qd = torch.from_numpy(queries).cuda()
    clean = qd[1:]                       # without the query that sits on the near-tie cluster
    D = torch.empty((40, k), dtype=torch.float32, device="cuda")
    I = torch.empty((40, k), dtype=torch.int64, device="cuda")
    Dh, Ih = idx.search(queries, k)      # blocking reference (redoes query 0 itself)
    assert idx.search_status() == 0      # a blocking search leaves nothing behind
    for _ in range(3):
        idx.search_device_async(clean.data_ptr(), 39, k, D.data_ptr(), I.data_ptr(), s)
    assert idx.search_status() == 0
    assert idx.stat("scan_searches") == 3 and idx.stat("scan_time_ns") > 0
    assert np.array_equal(I[:39].cpu().numpy(), Ih[1:]) and np.array_equal(D[:39].cpu().numpy(), Dh[1:])
    idx.search_device_async(qd.data_ptr(), 40, k, D.data_ptr(), I.data_ptr(), s)       # flagged
    idx.search_device_async(clean.data_ptr(), 39, k, D.data_ptr(), I.data_ptr(), s)    # clean, later
    assert idx.search_status() & 2       # the flag of the earlier search is not lost
    assert idx.search_status() == 0      # ... and is reported once
    idx.search_device(qd.data_ptr(), 40, k, D.data_ptr(), I.data_ptr(), s)             # the redo
    assert np.array_equal(I.cpu().numpy(), Ih) and np.array_equal(D.cpu().numpy(), Dh)
    idx.close()


def test_exchange_kernel_larger_than_one_wave():
    """K4 at nq = 4096, k = 1024, 4 ranks on one device: 4 x 4096 logical tiles against a few hundred
    resident CTAs.  Every CTA publishes all its queries before it waits, so progress does not
    depend on the order in which the hardware dispatches CTAs."""
    import torch
    from rag_inference_pipeline_b200.exchange import PeerExchange
    g, nq, k = 4, 4096, 1024
    group = [PeerExchange(0, g, r, nq, k) for r in range(g)]
    PeerExchange.connect_local(group)
    streams = [torch.cuda.Stream() for _ in range(g)]
    rng = np.random.default_rng(3)
    d = -np.sort(-rng.standard_normal((g, nq, k)).astype(np.float32), axis=2)
    ids = (np.sort(rng.integers(0, 1 << 20, (g, nq, k)), axis=2) + (np.arange(g) << 20)[:, None, None]).astype(np.int64)
    Dr, Ir = flatip_oracle.merge_sorted_lists(d[:, :64], ids[:, :64], k)
    td = [torch.from_numpy(d[r]).cuda() for r in range(g)]
    ti = [torch.from_numpy(ids[r]).cuda() for r in range(g)]
    od = [torch.empty((nq, k), dtype=torch.float32, device="cuda") for _ in range(g)]
    oi = [torch.empty((nq, k), dtype=torch.int64, device="cuda") for _ in range(g)]
    status = torch.zeros((g,), dtype=torch.int32, device="cuda")
    flags = torch.tensor([0, 0, 1, 0, 0, 1, 0, 0], dtype=torch.int32, device="cuda")   # rank 1 overflow, rank 2 uncertified
    torch.cuda.synchronize()
    for epoch in range(2):
        for r in range(g):
            group[r].merge_status(td[r].data_ptr(), ti[r].data_ptr(), nq, k, od[r].data_ptr(), oi[r].data_ptr(),
                                  flags.data_ptr() + 8 * r, status.data_ptr() + 4 * r, streams[r].cuda_stream)
        torch.cuda.synchronize()
    assert status.cpu().tolist() == [3, 3, 3, 3]           # every rank learned both flags
    for r in range(g):
        assert np.array_equal(od[r][:64].cpu().numpy(), Dr) and np.array_equal(oi[r][:64].cpu().numpy(), Ir)
        assert torch.equal(od[r], od[0]) and torch.equal(oi[r], oi[0])
    for x in group:
        x.close()


@pytest.mark.parametrize("sharded", [False, True])
def test_in_batch_duplicates_are_searched_once(shared_device, sharded):
    """callers whose batches are coalesced into one launch often carry the same
    question.  Bit-identical queries (the reference's cache key is sha256 of the fp32 bytes,
    api.py:398) are searched once on the device and every caller still gets its own, bit-identical
    rows -- on one index and on the sharded group."""
    from rag_inference_pipeline_b200 import FlatIPIndex
    from rag_inference_pipeline_b200.group import FlatIPGroup
    dim, n, k = 128, 40_000, 20
    if sharded:
        idx = FlatIPGroup(dim, "f16", [0, 0, 0], n)
        idx.add_synthetic(n, synthetic.CORPUS_SEED)
    else:
        idx = FlatIPIndex(dim, "f16", 0)
        idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    base = synthetic.synthetic_queries(24, dim)
    Dref, Iref = idx.search(base, k)
    # the expected rows are themselves checked against the CPU oracle, not only against the CUDA path
    Dor, Ior = flatip_oracle.flat_ip_search(base, synthetic.synthetic_corpus(0, n, dim, "f16"), k)
    res = flatip_oracle.compare_with_oracle(Dref, Iref, Dor, Ior, rtol=SCORE_RTOL["f16"])
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    # one call with internal duplicates: only searched once each when dedupe = 2 ("always")
    mixed = base[[0, 5, 0, 7, 5, 5, 23, 0]]
    tweaked = mixed.copy()
    tweaked[3, 17] = np.nextafter(tweaked[3, 17], np.float32(2.0))     # one ulp apart: NOT a duplicate
    idx.set_option("dedupe", 2)
    q0, d0 = idx.stat("host_queries"), idx.stat("distinct_queries")
    D, I = idx.search(tweaked, k)
    assert idx.stat("host_queries") - q0 == 8 and idx.stat("distinct_queries") - d0 == 4
    for row, src in enumerate([0, 5, 0, None, 5, 5, 23, 0]):
        if src is not None:
            assert np.array_equal(I[row], Iref[src]) and np.array_equal(D[row], Dref[src])
    # default (dedupe = 1): concurrent callers with overlapping batches
    idx.set_option("dedupe", 1)
    q0, d0 = idx.stat("host_queries"), idx.stat("distinct_queries")
    out = [None] * 8
    gate = threading.Barrier(8)

    def work(t):
        gate.wait()
        out[t] = idx.search(base[2 * t:2 * t + 10], k)      # neighbours share 8 of their 10 queries

    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for t in range(8):
        assert np.array_equal(out[t][1], Iref[2 * t:2 * t + 10]) and np.array_equal(out[t][0], Dref[2 * t:2 * t + 10])
    assert idx.stat("host_queries") - q0 == 80
    if idx.stat("coalesced_calls") > 0:                       # (some calls did share a batch)
        assert idx.stat("distinct_queries") - d0 < 80
    idx.close()


def test_weighted_shards_and_calibration(shared_device):
    """Shard sizes proportional to measured per-device throughput (a search ends with the slowest
    GPU): same answers as the single index bit for bit, boundaries where the weights put them."""
    from rag_inference_pipeline_b200.group import FlatIPGroup, calibrate_devices
    from rag_inference_pipeline_b200.sharded import shard_range
    dim, n, nq, k = 128, 50_000, 70, 30
    rates = calibrate_devices([0], dim, "f16", batch=64, k=10, rows=200_000, seconds=0.3)
    assert len(rates) == 1 and rates[0] > 1e6            # rows per second of scan time
    w = [1.0, 1.3, 0.8]
    q = synthetic.synthetic_queries(nq, dim)
    Ds, Is = _single(dim, "f16", n, q, k)
    g = FlatIPGroup(dim, "f16", [0, 0, 0], n, weights=w)
    g.add_synthetic(n, synthetic.CORPUS_SEED)
    bounds = g.shard_bounds()
    assert bounds == [shard_range(n, r, 3, w)[0] for r in range(3)] + [n]
    assert [g.shard(r).ntotal for r in range(3)] == [bounds[r + 1] - bounds[r] for r in range(3)]
    D, I = g.search(q, k)
    assert np.array_equal(I, Is) and np.array_equal(D, Ds)
    Dor, Ior = flatip_oracle.flat_ip_search(q, synthetic.synthetic_corpus(0, n, dim, "f16"), k)
    res = flatip_oracle.compare_with_oracle(D, I, Dor, Ior, rtol=SCORE_RTOL["f16"])
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res
    g.close()
