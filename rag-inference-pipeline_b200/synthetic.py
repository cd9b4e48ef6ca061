"""Counter-based synthetic corpus / query generator (host restatement).

The device generator (``csrc/synthetic.cuh``) and this numpy restatement produce the SAME bits
for any (seed, row, column), so any corpus row can be regenerated anywhere: on the GPU that owns
the shard, on another rank, or on the host for a spot check.  That is what lets the 100 M-row
benchmark corpus be created on device (no 150 GB host transfer) and still be verified on the host
by sampling rows.

Recipe (integer-only until the last multiply, so CPU and GPU agree bit for bit):

* Philox4x32-10 keyed by the 64-bit seed, counter = (row_lo, row_hi, col // 2, stream);
  one call yields four 32-bit words = two elements.
* element = (sum of the four 16-bit halves of two words - 131070) * SCALE, i.e. an Irwin-Hall(4)
  variate with zero mean and unit variance (bell-shaped, support +-3.46 sigma).  It stands in for
  the reference's ``rng.standard_normal`` corpus (``scripts/create_test_docs.py:87-96``).
* storage rounding: fp32 kept as is; fp16 / bf16 by round-to-nearest-even.

Queries use the same generator with another seed and are L2-normalised in fp32, like the
reference's embedder output (``src/pipeline/components/embedding.py:132,160``).
"""

from __future__ import annotations

import numpy as np

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)

# 1 / sqrt(4 * (65536**2 - 1) / 12): the std-dev of a sum of four uniform 16-bit integers.
IRWIN_HALL_SCALE = np.float32(1.0 / 37837.22723876)
IRWIN_HALL_MEAN = 131070  # 4 * 65535 / 2

CORPUS_SEED = 1234
QUERY_SEED = 4321

DTYPES = ("f32", "f16", "bf16")


def philox4x32_10(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10.  c0..c3: uint64 arrays holding 32-bit counters."""
    c0 = c0.astype(np.uint64)
    c1 = c1.astype(np.uint64)
    c2 = c2.astype(np.uint64)
    c3 = c3.astype(np.uint64)
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(10):
        p0 = PHILOX_M0 * c0
        p1 = PHILOX_M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + PHILOX_W0) & 0xFFFFFFFF
        k1 = (k1 + PHILOX_W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def _halves_sum(a, b):
    m16 = np.uint64(0xFFFF)
    s16 = np.uint64(16)
    return ((a & m16) + (a >> s16) + (b & m16) + (b >> s16)).astype(np.int64)


def synthetic_rows_f32(row0: int, nrows: int, dim: int, seed: int, stream: int = 0) -> np.ndarray:
    """fp32 rows [row0, row0+nrows) of the synthetic matrix with `dim` columns."""
    if nrows <= 0:
        return np.zeros((0, dim), dtype=np.float32)
    npairs = (dim + 1) // 2
    rows = np.arange(row0, row0 + nrows, dtype=np.uint64)[:, None]
    pairs = np.arange(npairs, dtype=np.uint64)[None, :]
    c0 = np.broadcast_to(rows & _MASK32, (nrows, npairs))
    c1 = np.broadcast_to(rows >> np.uint64(32), (nrows, npairs))
    c2 = np.broadcast_to(pairs, (nrows, npairs))
    c3 = np.full((nrows, npairs), stream, dtype=np.uint64)
    r0, r1, r2, r3 = philox4x32_10(c0, c1, c2, c3, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    e0 = (_halves_sum(r0, r1) - IRWIN_HALL_MEAN).astype(np.float32) * IRWIN_HALL_SCALE
    e1 = (_halves_sum(r2, r3) - IRWIN_HALL_MEAN).astype(np.float32) * IRWIN_HALL_SCALE
    out = np.empty((nrows, npairs * 2), dtype=np.float32)
    out[:, 0::2] = e0
    out[:, 1::2] = e1
    return np.ascontiguousarray(out[:, :dim])


def round_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit pattern (uint16), round-to-nearest-even (matches __float2bfloat16_rn)."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    lsb = (u >> np.uint32(16)) & np.uint32(1)
    rounded = u + np.uint32(0x7FFF) + lsb
    out = (rounded >> np.uint32(16)).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        out = np.where(nan, np.uint16(0x7FC0), out)
    return out


def bf16_bits_to_f32(bits: np.ndarray) -> np.ndarray:
    return (bits.astype(np.uint32) << np.uint32(16)).view(np.float32)


def round_to_storage(x: np.ndarray, dtype: str) -> np.ndarray:
    """Round fp32 values to the storage dtype and return them up-cast to fp32 again.

    This is "what the corpus really holds": the oracle sees exactly the values the GPU stores.
    """
    if dtype == "f32":
        return np.ascontiguousarray(x, dtype=np.float32)
    if dtype == "f16":
        return np.ascontiguousarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)
    if dtype == "bf16":
        return bf16_bits_to_f32(round_to_bf16_bits(x))
    raise ValueError(f"unknown storage dtype {dtype!r}; expected one of {DTYPES}")


def synthetic_corpus(row0: int, nrows: int, dim: int, dtype: str = "f32",
                     seed: int = CORPUS_SEED) -> np.ndarray:
    """Corpus rows as the GPU stores them (rounded to `dtype`), returned as fp32."""
    return round_to_storage(synthetic_rows_f32(row0, nrows, dim, seed), dtype)


def synthetic_queries(nq: int, dim: int, seed: int = QUERY_SEED) -> np.ndarray:
    """L2-normalised fp32 queries [nq, dim]."""
    q = synthetic_rows_f32(0, nq, dim, seed, stream=1)
    norm = np.sqrt((q.astype(np.float64) ** 2).sum(axis=1, keepdims=True))
    norm[norm == 0] = 1.0
    return (q / norm).astype(np.float32)
