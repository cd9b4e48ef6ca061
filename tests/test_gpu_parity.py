"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on
identical seeded inputs.  Acceptance rule (BASELINE.json north_star): ids bit-exact outside
near-tie windows; scores within 1e-5 relative (fp32 storage) / 1e-3 (fp16, bf16 storage)."""
import os
import threading

import numpy as np
import pytest

import c_oracle
import flatip_oracle
from flatip_oracle import SCORE_RTOL
from rag_inference_pipeline_b200 import synthetic

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _index(dim, dtype, **kw):
    from rag_inference_pipeline_b200 import FlatIPIndex
    return FlatIPIndex(dim, dtype, 0, **kw)


def _check(D, I, Dr, Ir, dtype, kplus1=None):
    res = flatip_oracle.compare_with_oracle(D, I, Dr, Ir, rtol=SCORE_RTOL[dtype], kplus1_ref=kplus1)
    assert res["score_ok"], res
    assert res["id_mismatch_outside_window"] == 0, res
    # windows must stay rare on random data, otherwise the rule is vacuous
    assert res["id_mismatch_total"] <= max(2, I.size // 200), res
    for row in D:  # best-first
        assert np.all(np.diff(row) <= 0)
    return res


def test_device_generator_matches_host_restatement():
    for dtype in synthetic.DTYPES:
        idx = _index(72, dtype)
        idx.add_synthetic(1500, 1234, 987654321)
        got = idx.reconstruct_n(0, 1500)
        assert np.array_equal(got, synthetic.synthetic_corpus(987654321, 1500, 72, dtype))
        idx.close()


def test_reference_fixture_inputs():
    """The reference's only real-index fixture (tests/test_retrieval_service.py:202-212,260-261),
    as a flat-IP search; expected values from the fp64 oracle (parity unpinned, see make_golden)."""
    g = np.load(os.path.join(GOLDEN, "reference_fixture_flatip.npz"))
    idx = _index(768, "f32")
    idx.add(g["corpus"])
    assert idx.ntotal == 10
    for k in (1, 3, 10, 12):
        D, I = idx.search(g["queries"], k)
        assert np.array_equal(I, g[f"I_k{k}"])
        np.testing.assert_allclose(D[I >= 0], g[f"D_k{k}"][I >= 0], rtol=1e-5)
        assert np.all(D[I < 0] == np.finfo(np.float32).min)
    idx.close()


@pytest.mark.parametrize("tag", ["f32_3000x64_q5_k10", "f16_5000x128_q33_k100", "bf16_2500x64_q8_k7"])
@pytest.mark.parametrize("variant", ["stream", "tensor"])
def test_golden_synthetic_cases(tag, variant):
    dtype, shape, q, k = tag.split("_")
    n, dim = (int(v) for v in shape.split("x"))
    nq, k = int(q[1:]), int(k[1:])
    g = np.load(os.path.join(GOLDEN, "synthetic_cases.npz"))
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    idx.set_variant(variant)
    D, I = idx.search(synthetic.synthetic_queries(nq, dim), k)
    assert idx.last_variant() == variant
    _check(D, I, g[f"D_{tag}"], g[f"I_{tag}"], dtype)
    idx.close()


CASES = [
    # variant, dtype, n, dim, nq, k
    ("stream", "f32", 20000, 768, 32, 10),      # config 2 shape, scaled down
    ("stream", "f32", 9999, 100, 1, 1),
    ("stream", "f16", 30011, 256, 3, 100),
    ("stream", "bf16", 12345, 64, 17, 50),
    ("tensor", "f16", 50000, 768, 64, 100),     # config 3 shape, scaled down
    ("tensor", "f16", 40000, 384, 130, 1024),   # config 5: large k, ragged query tile
    ("tensor", "bf16", 70001, 768, 200, 100),   # config 4: bf16, ragged corpus tile
    ("tensor", "bf16", 300, 64, 5, 10),         # corpus smaller than one tile
    ("tensor", "f16", 4096, 128, 128, 2048),    # k at its maximum, corpus == one dense phase
    ("tensor", "f32", 20000, 768, 32, 10),      # config 2 shape on the tf32 tensor path
    ("tensor", "f32", 33333, 96, 150, 100),     # fp32, dim % 32 == 0 only, two query tiles (CTA pair)
    ("tensor", "f16", 30000, 768, 1000, 100),   # 8 query tiles: CTA-pair kernel, ragged last tile
    ("tensor", "bf16", 25000, 128, 900, 50),    # 8 tiles (7 padded to 8), bf16 hi+lo terms, pair
    ("tensor", "f16", 20000, 256, 2048, 100),   # 16 query tiles (the sweep's B = 2048)
    ("tensor", "bf16", 15000, 128, 4096, 10),   # 32 query tiles (the sweep's B = 4096)
    ("tensor", "f16", 20000, 128, 127, 10),     # one tile, one row short
    ("tensor", "f16", 20000, 128, 129, 10),     # two tiles, the second holds one query
    ("tensor", "bf16", 20000, 128, 255, 10),    # two tiles (CTA pair), one row short
    ("tensor", "bf16", 20000, 128, 257, 10),    # three tiles: no pair
]

# Both sides of every switch of the AUTO rule (flatip_api.cu run_search): 16-bit storage streams
# only a single query, fp32 storage without the filter copy streams up to four.
AUTO_CASES = [
    # dtype, dim, nq, filter_copy, expected variant
    ("bf16", 512, 1, 1, "stream"), ("bf16", 512, 2, 1, "tensor"),
    ("f16", 512, 1, 1, "stream"), ("f16", 512, 2, 1, "tensor"),
    ("f32", 256, 4, 0, "stream"), ("f32", 256, 5, 0, "tensor"),
    ("f32", 256, 1, 1, "tensor"),
]


@pytest.mark.parametrize("dtype,dim,nq,filter_copy,expect", AUTO_CASES)
def test_auto_variant_switch_points_match_oracle(dtype, dim, nq, filter_copy, expect):
    n, k = 30000, 20
    corpus = c_oracle.synthetic_corpus(0, n, dim, dtype)
    queries = synthetic.synthetic_queries(nq, dim)
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k + 1)
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    if dtype == "f32":
        idx.set_option("filter_copy", filter_copy)
    D, I = idx.search(queries, k)
    assert idx.last_variant() == expect
    _check(D, I, Dr[:, :k], Ir[:, :k], dtype, Dr[:, k])
    idx.close()


@pytest.mark.parametrize("variant,dtype,n,dim,nq,k", CASES)
def test_parity_with_oracle(variant, dtype, n, dim, nq, k):
    corpus = c_oracle.synthetic_corpus(0, n, dim, dtype)
    queries = synthetic.synthetic_queries(nq, dim)
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, min(k + 1, max(n, 1)) if n > k else k)
    kplus1 = Dr[:, k] if Dr.shape[1] > k else None
    Dr, Ir = Dr[:, :k], Ir[:, :k]
    # second, independently written checker on the same case: the C port of faiss's algorithm
    # (sgemm blocks + heap / reservoir) must agree with the numpy fp64 oracle
    Dc, Ic = c_oracle.knn_inner_product(queries, corpus, k)
    _check(Dc, Ic, Dr, Ir, dtype, kplus1)
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    idx.set_variant(variant)
    D, I = idx.search(queries, k)
    assert idx.last_variant() == variant
    assert idx.stat("launches") > 0
    _check(D, I, Dr, Ir, dtype, kplus1)
    idx.close()


@pytest.mark.parametrize("scale,expect_copy", [(1.0, 1), (3.0e4, 0), (1.0e-3, 1)])
def test_fp32_storage_filters_on_fp16_copy(scale, expect_copy):
    """fp32 storage: the tensor-core scan reads an fp16 FILTER COPY of the corpus (half the bytes),
    survivors are rescored exactly from the fp32 rows: same bits as the tf32 filter and the exact
    streaming kernel; the copy follows adds and is skipped when a row leaves fp16's range."""
    n, dim, k = 30000, 256, 20
    corpus = (synthetic.synthetic_corpus(0, n + 5000, dim, "f32") * np.float32(scale)).astype(np.float32)
    queries = (synthetic.synthetic_queries(40, dim) * np.float32(scale)).astype(np.float32)
    idx = _index(dim, "f32")
    idx.add(corpus[:n])
    results = []
    for batch in (40, 1):
        q = queries[:batch]
        Dr, Ir = flatip_oracle.flat_ip_search(q, corpus[:n], k + 1)
        D, I = idx.search(q, k)
        # (without the copy a single fp32 query stays on the exact streaming kernel)
        assert idx.last_variant() == ("tensor" if expect_copy or batch > 4 else "stream")
        assert idx.stat("filter_f16") == expect_copy
        _check(D, I, Dr[:, :k], Ir[:, :k], "f32", Dr[:, k])
        results.append((D, I))
    idx.set_option("filter_copy", 0)
    D1, I1 = idx.search(queries, k)
    assert idx.stat("filter_f16") == 0
    idx.set_variant("stream")
    D2, I2 = idx.search(queries, k)
    assert idx.last_variant() == "stream"
    assert np.array_equal(I1, results[0][1]) and np.array_equal(D1, results[0][0])   # same rescoring kernel
    assert np.array_equal(I2, results[0][1])
    np.testing.assert_allclose(D2, results[0][0], rtol=2e-6, atol=1e-6 * scale * scale)
    idx.set_variant("auto")
    idx.set_option("filter_copy", 1)
    idx.add(corpus[n:])                                    # the copy is rebuilt for the grown corpus
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k + 1)
    D, I = idx.search(queries, k)
    assert idx.stat("filter_f16") == expect_copy
    _check(D, I, Dr[:, :k], Ir[:, :k], "f32", Dr[:, k])
    assert I.max() >= n                                    # (new rows do show up)
    idx.close()


def test_auto_variant_follows_arithmetic_intensity():
    idx = _index(512, "f16")                   # 1 KB rows
    idx.add_synthetic(5000, 1, 0)
    idx.search(synthetic.synthetic_queries(1, 512), 5)
    assert idx.last_variant() == "stream"      # 1 flop/byte: HBM-bound on CUDA cores
    idx.search(synthetic.synthetic_queries(2, 512), 5)
    assert idx.last_variant() == "tensor"      # 2 flop/byte: FMA + unpack fall to 0.9 of the roof, K1 holds it
    idx.search(synthetic.synthetic_queries(64, 512), 5)
    assert idx.last_variant() == "tensor"      # needs the tensor pipe to stay on the HBM roof
    idx.close()
    idx = _index(256, "f32")                   # 1 KB rows
    idx.add_synthetic(5000, 1, 0)
    idx.search(synthetic.synthetic_queries(64, 256), 5)
    assert idx.last_variant() == "tensor"      # fp32 storage: fp16 filter copy + exact fp32 rescoring
    assert idx.stat("filter_f16") == 1
    idx.search(synthetic.synthetic_queries(3, 256), 5)
    assert idx.last_variant() == "tensor"      # the copy halves the bytes: tensor path at any batch
    idx.set_option("filter_copy", 0)
    idx.search(synthetic.synthetic_queries(64, 256), 5)
    assert idx.last_variant() == "tensor" and idx.stat("filter_f16") == 0   # tf32 on the fp32 rows
    idx.search(synthetic.synthetic_queries(3, 256), 5)
    assert idx.last_variant() == "stream"      # 4-byte elements: CUDA cores keep up to batch 4
    idx.close()
    idx = _index(128, "f16")                   # 256-byte rows: half a warp per row -> tensor
    idx.add_synthetic(5000, 1, 0)
    idx.search(synthetic.synthetic_queries(1, 128), 5)
    assert idx.last_variant() == "tensor"
    idx.set_option("stream_max_batch", 4)      # explicit override
    idx.search(synthetic.synthetic_queries(4, 128), 5)
    assert idx.last_variant() == "stream"
    idx.close()
    idx = _index(40, "f16")                    # rows not a multiple of 128 bytes: no TMA tiles
    idx.add_synthetic(5000, 1, 0)
    idx.search(synthetic.synthetic_queries(64, 40), 5)
    assert idx.last_variant() == "stream"
    idx.close()


@pytest.mark.parametrize("dtype,dim", [("f16", 256), ("bf16", 128), ("f32", 64)])
def test_cta_pair_and_single_cta_kernels_agree(dtype, dim):
    """The cta_group::2 kernel and the single-CTA kernel must return identical results."""
    n, nq, k = 41000, 300, 20
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    idx.set_variant("tensor")
    q = synthetic.synthetic_queries(nq, dim)
    out = {}
    for mode in (1, 2):
        idx.set_option("cta_pair", mode)
        out[mode] = idx.search(q, k)
        assert idx.stat("cta_group") == mode
    np.testing.assert_array_equal(out[1][1], out[2][1])
    np.testing.assert_array_equal(out[1][0], out[2][0])
    Dr, Ir = flatip_oracle.flat_ip_search(q, c_oracle.synthetic_corpus(0, n, dim, dtype), k)
    _check(out[2][0], out[2][1], Dr, Ir, dtype)
    idx.close()


@pytest.mark.parametrize("variant,dtype", [("stream", "f32"), ("tensor", "f16")])
def test_edge_cases(variant, dtype):
    dim = 64
    idx = _index(dim, dtype)
    q = synthetic.synthetic_queries(4, dim)
    # empty corpus -> all padding, like faiss
    D, I = idx.search(q, 7)
    assert np.all(I == -1) and np.all(D == np.finfo(np.float32).min)
    # k > ntotal
    idx.add_synthetic(5, 3, 0)
    idx.set_variant(variant)
    D, I = idx.search(q, 9)
    corpus = synthetic.synthetic_corpus(0, 5, dim, dtype, seed=3)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, 9)
    _check(D, I, Dr, Ir, dtype)
    assert np.all(I[:, 5:] == -1)
    # zero queries -> empty result
    D0, I0 = idx.search(np.zeros((0, dim), np.float32), 3)
    assert D0.shape == (0, 3) and I0.shape == (0, 3)
    # the all-zero warm-up query of FAISSStore.load (faiss_store.py:103-107)
    Dz, Iz = idx.search(np.zeros((1, dim), np.float32), 1)
    assert Iz[0, 0] == 0 and Dz[0, 0] == 0.0    # all scores tie at 0 -> lowest id
    idx.close()


def test_bad_arguments_raise():
    from rag_inference_pipeline_b200._native import FlatIPError
    from rag_inference_pipeline_b200 import FlatIPIndex
    idx = _index(64, "f16")
    idx.add_synthetic(100, 1, 0)
    with pytest.raises(FlatIPError):
        idx.search(synthetic.synthetic_queries(2, 64), 0)
    with pytest.raises(FlatIPError):
        idx.search(synthetic.synthetic_queries(2, 64), 5000)
    with pytest.raises(ValueError):
        idx.search(synthetic.synthetic_queries(2, 32), 5)
    with pytest.raises(FlatIPError):
        FlatIPIndex(70, "f16", 0)      # rows must be 16-byte multiples
    idx.close()


@pytest.mark.parametrize("variant,dtype", [("stream", "f32"), ("tensor", "f16")])
def test_exact_duplicates_order_by_id(variant, dtype):
    dim, k = 64, 12
    base = synthetic.synthetic_corpus(0, 3000, dim, dtype)
    corpus = np.concatenate([base, base[:500], base[100:200]])
    queries = base[100:108] / np.linalg.norm(base[100:108], axis=1, keepdims=True)
    idx = _index(dim, dtype)
    idx.add(corpus)
    idx.set_variant(variant)
    D, I = idx.search(queries.astype(np.float32), k)
    Dr, Ir = flatip_oracle.flat_ip_search(queries.astype(np.float32), corpus, k)
    for qi in range(8):
        r = 100 + qi
        assert list(I[qi, :3]) == [r, 3000 + r, 3500 + qi]     # equal scores: ids ascending
        assert D[qi, 0] == D[qi, 1] == D[qi, 2]
    _check(D, I, Dr, Ir, dtype)
    idx.close()


@pytest.mark.parametrize("variant,dtype", [("stream", "f32"), ("tensor", "f16")])
def test_adversarial_order_overflow_fallback(variant, dtype):
    """Rows sorted so that every new row beats everything before it: the candidate lists overflow
    and the safe-mode redo must still return the exact answer."""
    dim, n, k = 64, 60000, 10
    rng = np.random.default_rng(5)
    direction = rng.standard_normal(dim).astype(np.float32)
    direction /= np.linalg.norm(direction)
    noise = 0.01 * rng.standard_normal((n, dim)).astype(np.float32)
    corpus = synthetic.round_to_storage(
        noise + np.linspace(0.0, 4.0, n, dtype=np.float32)[:, None] * direction[None, :], dtype)
    queries = np.stack([direction, -direction, synthetic.synthetic_queries(1, dim)[0]])
    idx = _index(dim, dtype)
    idx.add(corpus)
    idx.set_variant(variant)
    D, I = idx.search(queries, k)
    assert idx.stat("overflow_queries") >= 1
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k)
    _check(D, I, Dr, Ir, dtype)
    idx.close()


@pytest.mark.parametrize("dtype", ["f16", "bf16", "f32"])
def test_exactness_certificate_redoes_near_tie_clusters(dtype):
    """300 planted rows score within ~1e-4 of each other for query 0 -- closer than the error bound
    of the rounded tensor-core filter -- so that query cannot be certified and must be redone with
    a tighter variant; the other queries stay on the fast path.  Results are exact either way."""
    dim, n, k = 128, 30000, 10
    rng = np.random.default_rng(11)
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32").copy()
    queries = synthetic.synthetic_queries(40, dim)
    where = rng.choice(n, size=300, replace=False)
    corpus[where] = 9.0 * queries[0][None, :] + 2e-5 * rng.standard_normal((300, dim)).astype(np.float32)
    corpus = synthetic.round_to_storage(corpus, dtype)
    idx = _index(dim, dtype)
    idx.add(corpus)
    idx.set_variant("tensor")
    D, I = idx.search(queries, k)
    assert idx.stat("uncertified_queries") >= 1
    assert idx.stat("uncertified_queries") < 40          # ordinary queries certify on the fast path
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k + 1)
    _check(D, I, Dr[:, :k], Ir[:, :k], dtype, Dr[:, k])
    assert set(I[0].tolist()) <= set(where.tolist())
    # exact scores: the k best of the planted rows, recomputed in fp64
    ref0 = np.sort(corpus[where].astype(np.float64) @ queries[0].astype(np.float64))[::-1][:k]
    np.testing.assert_allclose(D[0], ref0, rtol=2e-6 if dtype == "f32" else 1e-5)
    # with the certificate off the flags are never raised
    idx.set_option("certify", 0)
    idx.search(queries, k)
    assert idx.stat("uncertified_queries") == 0
    idx.close()


@pytest.mark.parametrize("dtype", ["f32", "f16"])
def test_certificate_falls_back_on_norm_outlier(dtype):
    """One row with a huge norm inflates the (Cauchy-Schwarz) error bound so much that the exactness
    window of the tensor-core filter swallows the whole list: every query is flagged and redone
    (fp32 storage: straight on the exact streaming kernel).  Results stay exact."""
    dim, n, k = 64, 9000, 5
    corpus = synthetic.synthetic_corpus(0, n, dim, dtype).copy()
    corpus[1234] *= 4000.0
    corpus = synthetic.round_to_storage(corpus, dtype)
    queries = synthetic.synthetic_queries(6, dim)
    idx = _index(dim, dtype)
    idx.add(corpus)
    idx.set_variant("tensor")
    D, I = idx.search(queries, k)
    assert idx.stat("uncertified_queries") == 6
    Dr, Ir = flatip_oracle.flat_ip_search(queries, corpus, k)
    _check(D, I, Dr, Ir, dtype)
    idx.close()


def test_default_query_terms():
    for dtype in ("f16", "bf16", "f32"):
        idx = _index(128, dtype)
        idx.add_synthetic(20000, 3, 0)
        idx.search(synthetic.synthetic_queries(64, 128), 10)
        assert idx.last_variant() == "tensor" and idx.stat("query_split_used") == 1
        assert idx.stat("uncertified_queries") == 0
        idx.close()


def test_forced_safe_mode_matches():
    dim, n, k = 128, 50000, 20
    corpus = c_oracle.synthetic_corpus(0, n, dim, "f16")
    q = synthetic.synthetic_queries(9, dim)
    idx = _index(dim, "f16")
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    D1, I1 = idx.search(q, k)
    idx.set_option("force_safe_mode", 1)
    D2, I2 = idx.search(q, k)
    assert idx.stat("phases") > 3
    assert np.array_equal(I1, I2) and np.array_equal(D1, D2)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    _check(D1, I1, Dr, Ir, "f16")
    idx.close()


def test_results_are_deterministic_and_thread_safe():
    """search is called from a thread pool with overlapping batches
    (/root/reference/src/pipeline/services/retrieval/api.py:348-349) and feeds a sha256-keyed cache
    (api.py:398-421): same input -> same bits, from any thread."""
    dim, n, k = 128, 40000, 10
    idx = _index(dim, "f16")
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    qs = [synthetic.synthetic_queries(8 + 5 * t, dim, seed=100 + t) for t in range(6)]
    expected = [idx.search(q, k) for q in qs]
    results = [None] * len(qs)

    def work(t):
        for _ in range(3):
            results[t] = idx.search(qs[t], k)

    threads = [threading.Thread(target=work, args=(t,)) for t in range(len(qs))]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    for (D, I), (De, Ie) in zip(results, expected):
        assert np.array_equal(I, Ie) and np.array_equal(D, De)
    idx.close()


def test_id_base_and_incremental_add():
    dim, k = 64, 5
    corpus = synthetic.synthetic_corpus(0, 3000, dim, "f32")
    idx = _index(dim, "f32", id_base=1_000_000)
    idx.add(corpus[:1000])
    idx.add(corpus[1000:])
    assert idx.ntotal == 3000
    q = synthetic.synthetic_queries(4, dim)
    D, I = idx.search(q, k)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k, id_base=1_000_000)
    _check(D, I, Dr, Ir, "f32")
    idx.close()


def test_merge_kernel_matches_oracle():
    import torch
    from rag_inference_pipeline_b200 import merge_device

    rng = np.random.default_rng(3)
    g, nq, k = 8, 37, 100
    d = np.sort(rng.standard_normal((g, nq, k)).astype(np.float32), axis=2)[:, :, ::-1].copy()
    d[:, :, 40:60] = d[:, :, 40:41]                 # runs of equal scores inside a shard
    d[3] = d[2]                                     # and equal scores across shards
    ids = np.sort(rng.integers(0, 1000, (g, nq, k)), axis=2) + (np.arange(g) * 1000)[:, None, None]
    ids = ids.astype(np.int64)
    ids[5, :, 90:] = -1                             # a shard with fewer than k rows
    d[5, :, 90:] = np.finfo(np.float32).min
    Dr, Ir = flatip_oracle.merge_sorted_lists(d, ids, k)
    td, ti = torch.from_numpy(d).cuda(), torch.from_numpy(ids).cuda()
    od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    merge_device(0, td.data_ptr(), ti.data_ptr(), g, nq, k, k, od.data_ptr(), oi.data_ptr(),
                 torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(od.cpu().numpy(), Dr)
    # among equal scores the oracle orders by id; the kernel by (shard, position) == id here
    assert np.array_equal(oi.cpu().numpy(), Ir)


def test_fused_exchange_merge_kernel_in_process():
    """K4 with all ranks inside this process (several exchanges on one device, one stream each):
    every rank must end up with the oracle's merge of all ranks' lists, over several epochs with
    changing (nq, k) so that both buffer parities and layout changes are exercised."""
    import torch
    from rag_inference_pipeline_b200.exchange import PeerExchange

    g = 4
    group = [PeerExchange(0, g, r, 64, 128) for r in range(g)]
    PeerExchange.connect_local(group)
    streams = [torch.cuda.Stream() for _ in range(g)]
    rng = np.random.default_rng(5)
    for nq, k in [(37, 100), (64, 128), (1, 1), (5, 33), (37, 100)]:
        d = np.sort(rng.standard_normal((g, nq, k)).astype(np.float32), axis=2)[:, :, ::-1].copy()
        if k > 60:
            d[:, :, 40:60] = d[:, :, 40:41]
            d[3] = d[2]
        ids = (np.sort(rng.integers(0, 1000, (g, nq, k)), axis=2) + (np.arange(g) * 1000)[:, None, None]).astype(np.int64)
        if k > 10:
            ids[1, :, k - 5:] = -1
            d[1, :, k - 5:] = np.finfo(np.float32).min
        Dr, Ir = flatip_oracle.merge_sorted_lists(d, ids, k)
        td = [torch.from_numpy(d[r]).cuda() for r in range(g)]
        ti = [torch.from_numpy(ids[r]).cuda() for r in range(g)]
        od = [torch.empty((nq, k), dtype=torch.float32, device="cuda") for _ in range(g)]
        oi = [torch.empty((nq, k), dtype=torch.int64, device="cuda") for _ in range(g)]
        torch.cuda.synchronize()
        for r in range(g):   # rank r's kernel spins until the other ranks' kernels have published
            group[r].merge(td[r].data_ptr(), ti[r].data_ptr(), nq, k, od[r].data_ptr(), oi[r].data_ptr(),
                           streams[r].cuda_stream)
        torch.cuda.synchronize()
        for r in range(g):
            assert np.array_equal(od[r].cpu().numpy(), Dr), (r, nq, k)
            assert np.array_equal(oi[r].cpu().numpy(), Ir), (r, nq, k)
    with pytest.raises(Exception):
        group[0].merge(td[0].data_ptr(), ti[0].data_ptr(), 65, 10, od[0].data_ptr(), oi[0].data_ptr(), 0)
    for x in group:
        x.close()


def test_concurrent_callers_are_coalesced_into_one_batch():
    """The retrieval service calls search() from a thread pool with several batches in flight
    (/root/reference/src/pipeline/services/retrieval/api.py:348-349).  Calls that arrive while a
    search is running are served together; every caller still gets exactly its own rows."""
    import threading

    dim, n, k, nthreads, rounds, nq = 256, 3_000_000, 10, 8, 12, 8
    idx = _index(dim, "f16")
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    qs = [synthetic.synthetic_queries(nq, dim, seed=100 + t) for t in range(nthreads)]
    idx.set_option("coalesce", 0)
    serial = [idx.search(q, k) for q in qs]
    assert idx.stat("coalesced_calls") == 0
    # what the coalesced calls are compared with is itself checked against the CPU oracle
    Dr, Ir = _c_oracle_chunked(np.concatenate(qs), 0, n, dim, "f16", k + 1, 1_000_000)
    _check(np.concatenate([s[0] for s in serial]), np.concatenate([s[1] for s in serial]),
           Dr[:, :k], Ir[:, :k], "f16", Dr[:, k])
    idx.set_option("coalesce", 1)
    errors = []
    start = threading.Barrier(nthreads)

    def worker(t):
        try:
            start.wait()
            for _ in range(rounds):
                kk = k if t % 4 else k + 3          # one caller in four asks for a different k
                D, I = idx.search(qs[t], kk)
                assert np.array_equal(I[:, :k], serial[t][1]) and np.array_equal(D[:, :k], serial[t][0])
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(nthreads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    calls, batches = idx.stat("host_calls"), idx.stat("host_batches")
    assert calls == nthreads + nthreads * rounds
    assert idx.stat("coalesced_calls") > 0 and batches < calls, (calls, batches)
    idx.close()


@pytest.mark.parametrize("variant,dtype,nq,k", [("tensor", "f16", 70, 100), ("tensor", "bf16", 33, 10), ("tensor", "f16", 200, 1024),
                                                ("tensor", "f32", 16, 300), ("stream", "f16", 2, 100),
                                                ("stream", "f32", 5, 1)])
def test_live_thresholds_one_filter_launch_same_answer(variant, dtype, nq, k):
    """Live thresholds (a filter launch raises its own thresholds from the score histogram) must
    give bit-identical results to the fixed-threshold geometric phases, in fewer phases."""
    dim, n = 128, 1_500_000
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    idx.set_variant(variant)
    q = synthetic.synthetic_queries(nq, dim)
    idx.set_option("live_thresholds", 0)
    D0, I0 = idx.search(q, k)
    phases0 = idx.stat("phases")
    idx.set_option("live_thresholds", 1)
    D1, I1 = idx.search(q, k)
    assert idx.stat("overflow_queries") == 0
    assert idx.stat("phases") == 2 and phases0 >= 2, (idx.stat("phases"), phases0)
    assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    corpus = synthetic.synthetic_corpus(0, n, dim, dtype)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    _check(D1, I1, Dr, Ir, dtype)
    idx.close()


def test_packed_exchange_buffer_merge():
    """The packed [distances | pad | ids] per-rank buffer of sharded.py, gathered as [g, total]
    bytes, merged by the strided K3 entry point == the oracle merge."""
    import torch
    from rag_inference_pipeline_b200 import merge_strided_device
    from rag_inference_pipeline_b200.sharded import pack_layout

    rng = np.random.default_rng(11)
    g, nq, k = 4, 5, 33                               # nq * k odd: the ids need the padding
    d = np.sort(rng.standard_normal((g, nq, k)).astype(np.float32), axis=2)[:, :, ::-1].copy()
    ids = (np.sort(rng.integers(0, 500, (g, nq, k)), axis=2) + (np.arange(g) * 500)[:, None, None]).astype(np.int64)
    off_i, total = pack_layout(nq, k)
    packed = np.zeros((g, total), dtype=np.uint8)
    for r in range(g):
        packed[r, :nq * k * 4] = d[r].view(np.uint8).ravel()
        packed[r, off_i:off_i + nq * k * 8] = ids[r].view(np.uint8).ravel()
    Dr, Ir = flatip_oracle.merge_sorted_lists(d, ids, k)
    tp = torch.from_numpy(packed).cuda()
    od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    merge_strided_device(0, tp.data_ptr(), tp.data_ptr() + off_i, total // 4, total // 8, g, nq, k, k,
                         od.data_ptr(), oi.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(od.cpu().numpy(), Dr)
    assert np.array_equal(oi.cpu().numpy(), Ir)


def test_sharded_searcher_single_rank_matches_oracle():
    import torch
    from rag_inference_pipeline_b200 import ShardedFlatIPSearcher

    dim, n, nq, k = 128, 30_001, 9, 7
    s = ShardedFlatIPSearcher(dim, "f16", n, rank=0, world_size=1, device=0)
    s.load_synthetic(synthetic.CORPUS_SEED)
    q = synthetic.synthetic_queries(nq, dim)
    D, I = s.search_tensor(torch.from_numpy(q).cuda(), k)
    assert s.check() == 0          # search_tensor only enqueues: the verdict comes from check()
    corpus = synthetic.synthetic_corpus(0, n, dim, "f16")
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    _check(D.cpu().numpy(), I.cpu().numpy(), Dr, Ir, "f16")
    s.close()


def test_sharded_search_equals_single_index():
    """Row-shard emulation on one GPU: g shard indexes searched one after the other, results
    stacked like an all-gather leaves them, merged by K3 == one index over the whole corpus."""
    import torch
    from rag_inference_pipeline_b200 import merge_device
    from rag_inference_pipeline_b200.sharded import shard_range

    dim, n, nq, k, g = 128, 100_003, 48, 100, 4
    q = synthetic.synthetic_queries(nq, dim)
    whole = _index(dim, "f16")
    whole.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    Dw, Iw = whole.search(q, k)
    whole.close()
    ds, is_ = [], []
    for r in range(g):
        b, e = shard_range(n, r, g)
        sh = _index(dim, "f16", id_base=b)
        sh.add_synthetic(e - b, synthetic.CORPUS_SEED, b)
        d, i = sh.search(q, k)
        ds.append(d)
        is_.append(i)
        sh.close()
    td = torch.from_numpy(np.stack(ds)).cuda()
    ti = torch.from_numpy(np.stack(is_)).cuda()
    od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    oi = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    merge_device(0, td.data_ptr(), ti.data_ptr(), g, nq, k, k, od.data_ptr(), oi.data_ptr(),
                 torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(oi.cpu().numpy(), Iw)
    assert np.array_equal(od.cpu().numpy(), Dw)


def test_store_dropin_on_gpu(tmp_path):
    """B200FlatIPStore end to end: faiss flat file -> load (with warm-up) -> search -> unload."""
    import types
    from rag_inference_pipeline_b200 import B200FlatIPStore, index_io

    dim, n, k = 768, 2000, 10
    corpus = synthetic.synthetic_corpus(0, n, dim, "f32")
    path = tmp_path / "faiss_index.bin"
    index_io.write_faiss_flat(str(path), corpus)
    settings = types.SimpleNamespace(faiss_index_path=str(path), faiss_dim=dim, faiss_use_mmap=False)
    store = B200FlatIPStore(settings)
    store.load()
    store.load()  # second call is a no-op, like the reference
    assert store.is_loaded and store.index_size == n
    q = synthetic.synthetic_queries(32, dim)
    D, I = store.search(q.astype(np.float64), k)     # any float dtype is accepted
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k)
    _check(D, I, Dr, Ir, "f32")
    assert isinstance(I[0].tolist()[0], int) and isinstance(D[0].tolist()[0], float)
    assert store._index.stat("filter_f16") == 1      # fp32 file: filtered on the fp16 copy, exact results
    store.unload()
    assert not store.is_loaded
    lean = B200FlatIPStore(settings, {"filter_copy": False})   # keep only the fp32 rows
    lean.load()
    D1, I1 = lean.search(q, k)
    assert lean._index.stat("filter_f16") == 0
    assert np.array_equal(I1, I) and np.array_equal(D1, D)
    lean.unload()
    # fp16 storage + synthetic URI
    s2 = types.SimpleNamespace(faiss_index_path=f"synthetic:{n}?dim={dim}&dtype=f16", faiss_dim=dim)
    st2 = B200FlatIPStore(s2)
    st2.load()
    D2, I2 = st2.search(q, k)
    Dr2, Ir2 = flatip_oracle.flat_ip_search(q, synthetic.synthetic_corpus(0, n, dim, "f16"), k)
    _check(D2, I2, Dr2, Ir2, "f16")
    # embedding hand-off fast path: torch CUDA tensor in (fp16, as an embedder would emit),
    # torch CUDA tensors out, same answer as the host path on the same (rounded) queries
    import torch
    q16 = torch.from_numpy(q).cuda().half()
    D3, I3 = st2.search_device(q16, k)
    assert D3.is_cuda and I3.dtype == torch.int64 and tuple(D3.shape) == (32, k)
    D4, I4 = st2.search(q16.float().cpu().numpy(), k)
    assert np.array_equal(I3.cpu().numpy(), I4) and np.array_equal(D3.cpu().numpy(), D4)
    with pytest.raises(ValueError, match="Embedding dimension mismatch"):
        st2.search_device(torch.zeros((2, dim + 8), device="cuda"), k)
    with pytest.raises(ValueError, match="CUDA tensor"):
        st2.search_device(torch.zeros((2, dim)), k)
    st2.unload()
    with pytest.raises(RuntimeError, match="not loaded"):
        st2.search_device(q16, k)


def test_large_scale_properties():
    """10 M x 768 bf16, batch 1024, k = 100 (BASELINE config 4 at full size): too big for the
    oracle, so check size-independent properties -- rows sorted, returned scores reproduce on the
    host from regenerated rows, and no sampled corpus block holds a row that should have displaced
    the k-th result."""
    dim, n, nq, k = 768, 10_000_000, 1024, 100
    idx = _index(dim, "bf16", capacity_rows=n)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    q = synthetic.synthetic_queries(nq, dim)
    D, I = idx.search(q, k)
    assert idx.last_variant() == "tensor" and idx.stat("overflow_queries") == 0
    assert np.all(np.diff(D, axis=1) <= 0) and I.min() >= 0 and I.max() < n
    for row in I[::97]:
        assert len(set(row.tolist())) == k
    # recompute a sample of returned scores from regenerated rows (fp64)
    for qi in range(0, nq, 61):
        rows = np.stack([c_oracle.synthetic_corpus(int(r), 1, dim, "bf16")[0] for r in I[qi, ::9]])
        ref = rows.astype(np.float64) @ q[qi].astype(np.float64)
        np.testing.assert_allclose(D[qi, ::9], ref, rtol=1e-3, atol=1e-3)
    # sampled blocks: nothing outside the result may beat the k-th score by more than the tolerance
    for b0 in (0, 4_999_936, n - 4096):
        blk = c_oracle.synthetic_corpus(b0, 4096, dim, "bf16")
        s = q.astype(np.float32) @ blk.T
        for qi in range(0, nq, 17):
            better = np.nonzero(s[qi] > D[qi, -1] + 1e-3 * max(1.0, abs(D[qi, -1])))[0] + b0
            assert set(better.tolist()) <= set(I[qi].tolist())
    idx.close()


# ---------------------------------------------------------------------------------------------
# The BASELINE configurations AT THEIR REAL SIZES.  The checker is the C port
# of faiss's algorithm (fp32 sgemm blocks + heap / reservoir) run over row chunks and merged with
# the numpy oracle's k-way merge; both sides accumulate exact products of the stored values in
# fp32, so scores must agree to ~1e-5 and the near-tie windows of the id rule are tiny.
# ---------------------------------------------------------------------------------------------
FULL_RTOL = 2e-5


def _c_oracle_chunked(q, row0, n, dim, dtype, k, chunk_rows):
    c_oracle.set_threads(os.cpu_count() or 1)
    Ds, Is = [], []
    for r0 in range(0, n, chunk_rows):
        m = min(chunk_rows, n - r0)
        x = c_oracle.synthetic_corpus(row0 + r0, m, dim, dtype)
        D, I = c_oracle.knn_inner_product(q, x, k, id_base=row0 + r0)
        Ds.append(D)
        Is.append(I)
        del x
    return flatip_oracle.merge_sorted_lists(np.stack(Ds), np.stack(Is), k)


def _check_full(D, I, Dr, Ir, k):
    """Dr / Ir hold k + 1 columns (the (k+1)-th closes the last near-tie window)."""
    res = flatip_oracle.compare_with_oracle(D, I, Dr[:, :k], Ir[:, :k], rtol=FULL_RTOL,
                                            kplus1_ref=Dr[:, k])
    assert res["score_ok"], res
    assert res["id_mismatch_outside_window"] == 0, res
    # Both sides sum exact products in fp32 in different orders, so neighbours closer than ~1e-6
    # may swap (0.2 % of the positions at k = 1024 of 1e7 rows).  Swaps must stay swaps: wherever
    # the k-th and (k+1)-th oracle scores are clearly apart, the two id SETS of a row are equal.
    assert res["id_mismatch_total"] <= max(2, I.size // 200), res
    tol = FULL_RTOL * np.maximum(1.0, np.abs(Dr[:, k - 1].astype(np.float64)))
    clear = (Dr[:, k - 1].astype(np.float64) - Dr[:, k].astype(np.float64)) > 2 * tol
    assert clear.mean() > 0.3      # (at k = 1024 of 1e7 rows about half of the boundaries are that close)
    for r in np.nonzero(clear)[0]:
        assert set(I[r].tolist()) == set(Ir[r, :k].tolist()), r
    assert np.all(np.diff(D, axis=1) <= 0)
    return res


def test_cfg2_full_size_all_variants():
    """BASELINE configs[1] in full: 1 M x 768 fp32, batch 32, k = 10, against the fp64 oracle
    (tolerance 1e-5 as north_star states for fp32), on every route the index can take: tensor-core
    filter on the fp16 copy (default), tf32 filter on the fp32 rows (filter_copy = 0, the
    memory-neutral configuration) and the exact streaming kernel K2."""
    n, dim, nq, k = 1_000_000, 768, 32, 10
    corpus = c_oracle.synthetic_corpus(0, n, dim, "f32")
    q = synthetic.synthetic_queries(nq, dim)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k + 1)
    Dc, Ic = c_oracle.knn_inner_product(q, corpus, k + 1)          # the two checkers agree
    _check_full(Dc[:, :k], Ic[:, :k], Dr, Ir, k)
    del corpus
    idx = _index(dim, "f32", capacity_rows=n)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    got = {}
    for name, opts in (("filter16", {}), ("tf32", {"filter_copy": 0}), ("stream", {"variant": "stream"})):
        idx.set_variant(opts.get("variant", "auto"))
        idx.set_option("filter_copy", opts.get("filter_copy", 1))
        D, I = idx.search(q, k)
        assert idx.last_variant() == ("stream" if name == "stream" else "tensor")
        assert idx.stat("filter_f16") == (1 if name == "filter16" else 0)
        res = flatip_oracle.compare_with_oracle(D, I, Dr[:, :k], Ir[:, :k], rtol=1e-5, kplus1_ref=Dr[:, k])
        assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, (name, res)
        assert res["id_mismatch_total"] <= 2, (name, res)
        got[name] = (D, I)
    assert np.array_equal(got["filter16"][1], got["tf32"][1]) and np.array_equal(got["filter16"][0], got["tf32"][0])
    assert np.array_equal(got["filter16"][1], got["stream"][1])
    idx.close()


def test_cfg3_shard_full_size():
    """One 12.5 M x 768 fp16 shard of the headline configuration (BASELINE configs[2]: shard 3 of
    8, global rows [37.5 M, 50 M)), batch 1024, k = 100: ONE filter launch over 1.25e7 rows with
    live thresholds, exactly as the bench runs it, against the C port over the same rows."""
    dim, n, nq, k, row0 = 768, 12_500_000, 1024, 100, 37_500_000
    q = synthetic.synthetic_queries(nq, dim)
    idx = _index(dim, "f16", capacity_rows=n, id_base=row0)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, row0)
    D, I = idx.search(q, k)
    assert idx.last_variant() == "tensor" and idx.stat("phases") == 2
    assert idx.stat("overflow_queries") == 0
    assert I.min() >= row0 and I.max() < row0 + n
    idx.close()
    Dr, Ir = _c_oracle_chunked(q, row0, n, dim, "f16", k + 1, 2_500_000)
    _check_full(D, I, Dr, Ir, k)


def test_cfg4_full_size():
    """BASELINE configs[3] in full at the largest batch the sweep quotes a roofline fraction for:
    10 M x 768 bf16, batch 1024, k = 100, against the C port over the same rows."""
    dim, n, nq, k = 768, 10_000_000, 1024, 100
    q = synthetic.synthetic_queries(nq, dim)
    idx = _index(dim, "bf16", capacity_rows=n)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    D, I = idx.search(q, k)
    assert idx.last_variant() == "tensor" and idx.stat("overflow_queries") == 0
    idx.close()
    Dr, Ir = _c_oracle_chunked(q, 0, n, dim, "bf16", k + 1, 2_500_000)
    _check_full(D, I, Dr, Ir, k)


def test_cfg5_full_size():
    """BASELINE configs[4] in full: 10 M x 384 fp16, batch 256, k = 1024 (the large-k selection
    stress) against the C port (reservoir path, k >= 100)."""
    dim, n, nq, k = 384, 10_000_000, 256, 1024
    q = synthetic.synthetic_queries(nq, dim)
    idx = _index(dim, "f16", capacity_rows=n)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    D, I = idx.search(q, k)
    assert idx.last_variant() == "tensor" and idx.stat("overflow_queries") == 0
    idx.close()
    Dr, Ir = _c_oracle_chunked(q, 0, n, dim, "f16", k + 1, 5_000_000)
    _check_full(D, I, Dr, Ir, k)


@pytest.mark.parametrize("dtype,n,dim,nq,k", [("f16", 60000, 768, 100, 100), ("bf16", 30000, 128, 7, 10),
                                              ("f32", 40000, 96, 40, 20), ("f16", 50000, 384, 130, 1024)])
def test_fused_rescore_equals_separate_kernel(dtype, n, dim, nq, k):
    """The pass after the last scan finishes the query in one kernel (exactness window, exact
    rescoring of what is inside, final sort: two launches and two passes over the lists less);
    window select + stand-alone rescoring kernel + final select must give the same bits."""
    idx = _index(dim, dtype)
    idx.add_synthetic(n, synthetic.CORPUS_SEED, 0)
    idx.set_variant("tensor")
    q = synthetic.synthetic_queries(nq, dim)
    D1, I1 = idx.search(q, k)
    l1 = idx.stat("launches")
    idx.set_option("fused_rescore", 0)
    D0, I0 = idx.search(q, k)
    assert idx.stat("launches") == l1 + 2 and idx.stat("overflow_queries") == 0
    assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    corpus = c_oracle.synthetic_corpus(0, n, dim, dtype)
    Dr, Ir = flatip_oracle.flat_ip_search(q, corpus, k + 1)
    _check(D1, I1, Dr[:, :k], Ir[:, :k], dtype, Dr[:, k])
    idx.close()
