"""CPU tests: the oracle (numpy fp64 + the C restatement of faiss's algorithm) against the golden
fixtures, and the synthetic generator restatements against each other."""
import os

import numpy as np
import pytest

import c_oracle
import flatip_oracle
from rag_inference_pipeline_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_reference_fixture_inputs_and_results():
    g = np.load(os.path.join(GOLDEN, "reference_fixture_flatip.npz"))
    # inputs are the reference's own (tests/test_retrieval_service.py:202-212,260-261)
    corpus = np.random.default_rng(42).random((10, 768)).astype("float32")
    queries = np.random.default_rng(42).random((2, 768)).astype("float32")
    assert np.array_equal(g["corpus"], corpus)
    assert np.array_equal(g["queries"], queries)
    for k in (1, 3, 10, 12):
        D, I = flatip_oracle.flat_ip_search(queries, corpus, k)
        assert np.array_equal(I, g[f"I_k{k}"])
        np.testing.assert_allclose(D, g[f"D_k{k}"], rtol=0, atol=0)
        Dc, Ic = c_oracle.knn_inner_product(queries, corpus, k)
        assert np.array_equal(Ic, I)
        np.testing.assert_allclose(Dc, D, rtol=1e-5)
    # query 0 equals corpus row 0 (same rng stream): it must be its own best match
    assert g["I_k1"][0, 0] == 0
    # k > ntotal pads with -1 / -FLT_MAX like faiss
    assert np.all(g["I_k12"][:, 10:] == -1)
    assert np.all(g["D_k12"][:, 10:] == np.finfo(np.float32).min)


@pytest.mark.parametrize("tag", ["f32_3000x64_q5_k10", "f16_5000x128_q33_k100", "bf16_2500x64_q8_k7"])
def test_synthetic_golden_cases(tag):
    g = np.load(os.path.join(GOLDEN, "synthetic_cases.npz"))
    dtype, shape, q, k = tag.split("_")
    n, dim = (int(v) for v in shape.split("x"))
    nq, k = int(q[1:]), int(k[1:])
    corpus = synthetic.synthetic_corpus(0, n, dim, dtype)
    assert np.array_equal(corpus[:4, :8], g[f"sample_{tag}"])
    queries = synthetic.synthetic_queries(nq, dim)
    D, I = flatip_oracle.flat_ip_search(queries, corpus, k)
    assert np.array_equal(I, g[f"I_{tag}"])
    assert np.array_equal(D, g[f"D_{tag}"])
    # C restatement of faiss's algorithm: both code paths (seq for nq < 20, blas above) and both
    # result handlers (heap for k < 100, reservoir from 100)
    Dc, Ic = c_oracle.knn_inner_product(queries, corpus, k)
    res = flatip_oracle.compare_with_oracle(Dc, Ic, D, I, rtol=1e-5)
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res


@pytest.mark.parametrize("use_blas", [True, False])
@pytest.mark.parametrize("nq,k", [(1, 1), (19, 99), (20, 100), (45, 250)])
def test_c_restatement_paths(nq, k, use_blas):
    rng = np.random.default_rng(nq * 1000 + k)
    corpus = rng.standard_normal((4000, 96)).astype(np.float32)
    queries = rng.standard_normal((nq, 96)).astype(np.float32)
    D, I = flatip_oracle.flat_ip_search(queries, corpus, k)
    Dc, Ic = c_oracle.knn_inner_product(queries, corpus, k, use_blas=use_blas)
    res = flatip_oracle.compare_with_oracle(Dc, Ic, D, I, rtol=1e-5)
    assert res["score_ok"] and res["id_mismatch_outside_window"] == 0, res


def test_tie_break_score_desc_id_asc():
    rng = np.random.default_rng(7)
    base = rng.standard_normal((50, 32)).astype(np.float32)
    corpus = np.concatenate([base, base[:20], base[:10]])  # exact duplicates at higher ids
    queries = base[:4] / np.linalg.norm(base[:4], axis=1, keepdims=True)
    D, I = flatip_oracle.flat_ip_search(queries, corpus, 9)
    for q in range(4):
        assert list(I[q, :3]) == [q, 50 + q, 70 + q]  # same score three times: ids ascending
        assert D[q, 0] == D[q, 1] == D[q, 2]
    Dc, Ic = c_oracle.knn_inner_product(queries, corpus, 9)
    assert np.array_equal(Ic[:, :3], I[:, :3])


def test_empty_and_small_corpus():
    q = synthetic.synthetic_queries(3, 16)
    D, I = flatip_oracle.flat_ip_search(q, np.zeros((0, 16), np.float32), 4)
    assert np.all(I == -1) and np.all(D == flatip_oracle.LOWEST_F32)
    Dc, Ic = c_oracle.knn_inner_product(q, np.zeros((0, 16), np.float32), 4)
    assert np.all(Ic == -1) and np.all(Dc == flatip_oracle.LOWEST_F32)


def test_synthetic_generators_agree_bitwise():
    for dtype in synthetic.DTYPES:
        a = synthetic.synthetic_corpus(123456789012, 257, 70, dtype)
        b = c_oracle.synthetic_corpus(123456789012, 257, 70, dtype)
        assert np.array_equal(a, b)
    x = c_oracle.synthetic_corpus(0, 20000, 64, "f32")
    assert abs(float(x.mean())) < 0.01 and abs(float(x.std()) - 1.0) < 0.01
    q = synthetic.synthetic_queries(5, 64)
    np.testing.assert_allclose(np.linalg.norm(q, axis=1), 1.0, rtol=1e-6)


def test_bf16_rounding_is_rne():
    x = np.array([1.0, 1.00390625, 1.005859375, -2.0078125, 3.0e-39, np.inf], dtype=np.float32)
    bits = synthetic.round_to_bf16_bits(x)
    back = synthetic.bf16_bits_to_f32(bits)
    assert back[0] == 1.0
    assert back[1] == 1.0            # exactly half way between 1.0 and 1.0078125 -> even
    assert back[2] == 1.0078125
    assert back[5] == np.inf


def test_merge_oracle_matches_global_search():
    corpus = synthetic.synthetic_corpus(0, 3000, 32, "f16")
    q = synthetic.synthetic_queries(6, 32)
    k = 20
    D, I = flatip_oracle.flat_ip_search(q, corpus, k)
    parts_d, parts_i = [], []
    for r in range(3):
        d, i = flatip_oracle.flat_ip_search(q, corpus[r * 1000:(r + 1) * 1000], k, id_base=r * 1000)
        parts_d.append(d)
        parts_i.append(i)
    Dm, Im = flatip_oracle.merge_sorted_lists(np.stack(parts_d), np.stack(parts_i), k)
    assert np.array_equal(Im, I) and np.array_equal(Dm, D)
