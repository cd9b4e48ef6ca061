// K6: tensor-core tile kernel for HEAVILY PROBED inverted lists of the IVF-Flat (L2) operator.
//
// When many queries of a batch probe the same list, scoring them against it is a dense
// [queries x list rows x dim] contraction: it belongs on the tcgen05 tensor cores, not on FFMA2
// (ivf.cuh's ivf_scan_kernel tops out at ~18 TFMA/s).  The host plan (ivf_api.cuh) gathers, per
// list, the queries that probe it into groups of up to 128 rows of a prepared-query matrix and
// emits items (256-row tile of the list x one 128-query group).  This kernel is K1's single-CTA
// tf32 pipeline (tc_scan.cuh: TMA producer warp, one MMA-issuing thread, TMEM double buffer,
// epilogue warps with one query per TMEM lane) driven by that item list, with the IVF epilogue:
//     score = 2 * <q, x>_tf32 - ||x||^2      (per-row norm added as a bias, broadcast loads)
// filtered against the query's LIVE threshold and appended to the same candidate lists the FFMA2
// kernel feeds.  tf32 scores are approximate: the operator's exactness window (ivf_eps_kernel,
// widened by the tf32 term when this kernel runs) keeps every candidate that could belong to the
// exact top-k, and ivf_rescore_kernel ranks the survivors by sum (q - x)^2 in fp32.
#pragma once

#include "ivf.cuh"
#include "tc_scan.cuh"

namespace fip {
namespace ivf {

struct TcItem {
    uint32_t row0;      // first list position of the tile (sorted storage)
    uint32_t row_end;   // end of the list segment (tile covers [row0, min(row0 + 256, row_end)))
    uint32_t qrow0;     // first row of the 128-query group in the prepared-query matrix
    uint32_t list0;     // first position of the list (TileQuery::skip counts rows from here)
    uint32_t qrows;     // prepared rows of the group: its query count rounded up to 32 (<= 128)
};
struct TcIvfArgs {
    const TcItem* items;
    int n_items;
    const int32_t* qmap;     // [prepared rows] query number, -1 = padding
    const uint32_t* qskip;   // [prepared rows] rows of the list this query already scored in phase 0
    const float* xnorm;      // [N]
    int dim;
    ScanArgs lists;          // thr / cnt / cand / cap / hist / hbase / k_thr / win_* (common.cuh)
};

constexpr int TCI_EPI_WARPS = 16;  // four per TMEM lane quarter: 2 chunks of 32 columns each (the epilogue,
                                   // ~5 instructions per score with the norm bias, must keep up with kind::f16)
constexpr int TCI_THREADS = (tc::EPI_WARP0 + TCI_EPI_WARPS) * 32;
constexpr int TCI_CHUNKS = tc::BN / 32 / (TCI_EPI_WARPS / 4);

// One thread = one query of the group; chunks c0 .. c0 + TCI_CHUNKS - 1 of 32 accumulator columns.
// Same two-pass scheme as tc::epilogue_filter (pass 1 builds per-chunk masks, ONE atomicAdd
// reserves the slots, pass 2 re-reads the chunks that had survivors), with the norm bias applied.
__device__ __forceinline__ void tci_load_chunk(const TcIvfArgs& a, uint32_t taddr, int c, uint32_t row_base,
                                               uint32_t row_end, uint32_t row_min, float (&v)[32]) {
    tc::tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
    if (row_base < row_min) {   // this query pre-scanned the head of the list in phase 0 (rare)
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (row_base + j < row_min) v[j] = kNegInf;
    }
    // every lane reads the same norm: one broadcast transaction per load (lists start at arbitrary
    // positions, so the norms cannot be fetched as aligned vectors); rows past the list are dropped
    if (row_base + 32u <= row_end) {   // warp-uniform: the chunk lies inside the list (the common case)
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaf(2.0f, v[j], -__ldg(a.xnorm + row_base + j));
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            v[j] = row_base + j < row_end ? fmaf(2.0f, v[j], -__ldg(a.xnorm + row_base + j)) : kNegInf;
    }
}

__device__ __forceinline__ void tci_epilogue(const TcIvfArgs& a, uint32_t taddr, int q, float thr, uint32_t hb,
                                             uint32_t tile_row0, uint32_t row_end, uint32_t row_min, int c0) {
    static_assert(TCI_CHUNKS <= 4, "masks are packed into two 64-bit registers");
    uint64_t m01 = 0ull, m23 = 0ull;
    uint32_t n = 0;
#pragma unroll 1
    for (int ci = 0; ci < TCI_CHUNKS; ++ci) {
        const int c = c0 + ci;
        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
        if (row_base >= row_end) break;  // warp-uniform
        float v[32];
        tci_load_chunk(a, taddr, c, row_base, row_end, row_min, v);
        uint32_t mask = 0u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            float mx = fmaxf(fmaxf(v[8 * g], v[8 * g + 1]), fmaxf(v[8 * g + 2], v[8 * g + 3]));
            mx = fmaxf(mx, fmaxf(fmaxf(v[8 * g + 4], v[8 * g + 5]), fmaxf(v[8 * g + 6], v[8 * g + 7])));
            if (mx > thr) {
#pragma unroll
                for (int j = 0; j < 8; ++j) tc::mask_or_if_gt(mask, v[8 * g + j], thr, 1u << (8 * g + j));
            }
        }
        const uint64_t shifted = (uint64_t)mask << (32 * (ci & 1));
        if (ci & 2) m23 |= shifted; else m01 |= shifted;
        n += (uint32_t)__popc(mask);
    }
    uint32_t slot = 0u;
    bool fits = false;
    if (n) {   // n != 0 implies an active query (inactive lanes carry thr = +inf)
        slot = atomicAdd(a.lists.cnt + q, n);
        fits = slot + n <= a.lists.cap;   // else: overflow, detected by the select pass -> safe mode
    }
    if (!__any_sync(0xFFFFFFFFu, n != 0u)) return;
    uint64_t* list = a.lists.cand + (size_t)(q < 0 ? 0 : q) * a.lists.cap + slot;
    uint32_t* hrow = (a.lists.hist && q >= 0) ? a.lists.hist + (size_t)q * kHistBins : nullptr;
#pragma unroll 1
    for (int ci = 0; ci < TCI_CHUNKS; ++ci) {
        const uint32_t cmask = (uint32_t)(((ci & 2) ? m23 : m01) >> (32 * (ci & 1)));
        if (!__any_sync(0xFFFFFFFFu, cmask != 0u)) continue;  // warp-uniform
        const uint32_t mask = fits ? cmask : 0u;
        const int c = c0 + ci;
        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
        float v[32];
        tci_load_chunk(a, taddr, c, row_base, row_end, row_min, v);
        const uint32_t inv_row0 = 0xFFFFFFFFu - row_base;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if ((mask >> (8 * g)) & 0xFFu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t bits = __float_as_uint(v[8 * g + j]);
                    const uint32_t take = mask & (1u << (8 * g + j));
                    tc::list_append_if(list, inv_row0 - (uint32_t)(8 * g + j), bits, take);
                    if (take && hrow) atomicAdd(hrow + hist_bucket(bits, hb), 1u);
                }
            }
        }
    }
    if (a.lists.hist)
        refresh_thresholds_warp(a.lists, fits && ((slot + n) / kRefreshEvery != slot / kRefreshEvery), q,
                                (int)(threadIdx.x & 31));
}

// tmap_q: prepared queries [sum of qrows, dim] fp32 (read as tf32), [32 floats x 32 rows]
// SWIZZLE_128B boxes: a group occupies qrows (a multiple of 32, <= 128) rows and only those are
// loaded; the rest of the 128-row A tile keeps stale shared memory, whose (garbage) accumulator
// lanes belong to no query.  tmap_x: sorted list storage [N, dim] fp32, make_tmap's
// [32 floats x 128 rows] boxes.
// kFmt = 2: both maps fp32, kind::tf32; kFmt = 0: both maps fp16 (the index's fp16 FILTER COPY of the
// lists and fp16-rounded prepared queries), kind::f16 -- half the bytes and twice the tensor rate;
// the fp32 rows are only touched by the exact rescoring of the survivors.
template <int kFmt>
__global__ void __launch_bounds__(TCI_THREADS, 1)
tc_ivf_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x, TcIvfArgs a) {
    using C = tc::Cfg<kFmt, 1, 1>;
    constexpr int STAGES = C::STAGES;
    constexpr int STAGE_BYTES = C::STAGE_BYTES;
    constexpr int B_OFF = tc::A_BYTES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = base + STAGES * STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
    auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - tc::smem_u32(smem_raw)));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) { tc::tma_prefetch_desc(&tmap_q); tc::tma_prefetch_desc(&tmap_x); }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(full_bar(s), 1); tc::mbar_init(empty_bar(s), 1); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(tfull_bar(b), 1); tc::mbar_init(tempty_bar(b), TCI_EPI_WARPS); }
        tc::fence_barrier_init();
    }
    if (warp == 2) tc::tmem_alloc<1>(tmem_slot, tc::TMEM_COLS);
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    const int nkb = a.dim / C::BK;   // 128 bytes of K per stage: 32 floats or 64 halves
    // static round-robin over the items (adjacent items = adjacent row tiles of one list: L2 locality)
    const int first = (int)blockIdx.x, step = (int)gridDim.x;

    if (warp == 0) {
        if (lane == 0) {  // ===================== TMA producer
            int s = 0;
            uint32_t ph = 0;
            for (int t = first; t < a.n_items; t += step) {
                const TcItem it = a.items[t];
                for (int kb = 0; kb < nkb; ++kb) {
                    tc::mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t sa = base + s * STAGE_BYTES;
                    tc::mbar_arrive_expect_tx(full_bar(s), tc::B_BYTES + it.qrows * tc::BK_BYTES);
                    for (uint32_t r = 0; r < it.qrows; r += 32)
                        tc::tma_load_2d(sa + r * tc::BK_BYTES, &tmap_q, full_bar(s), kb * C::BK, (int)(it.qrow0 + r));
                    tc::tma_load_2d(sa + B_OFF, &tmap_x, full_bar(s), kb * C::BK, (int)it.row0);
                    tc::tma_load_2d(sa + B_OFF + tc::BOX_ROWS * tc::BK_BYTES, &tmap_x, full_bar(s), kb * C::BK,
                                    (int)it.row0 + tc::BOX_ROWS);
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===================== MMA issuer
            constexpr uint32_t idesc = tc::make_idesc(kFmt, kFmt, tc::BM, tc::BN);
            constexpr int kKind = kFmt == 2 ? 1 : 0;
            int s = 0, n = 0;
            uint32_t ph = 0;
            for (int t = first; t < a.n_items; t += step, ++n) {
                const int buf = n & 1;
                const uint32_t aph = (uint32_t)(n >> 1) & 1u;
                tc::mbar_wait(tempty_bar(buf), aph ^ 1u);
                tc::tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)buf * tc::BN;
                for (int kb = 0; kb < nkb; ++kb) {
                    tc::mbar_wait(full_bar(s), ph);
                    tc::tcgen05_fence_after();
                    const uint32_t sa = base + s * STAGE_BYTES;
                    const uint64_t adesc = tc::make_sw128_desc(sa), bdesc = tc::make_sw128_desc(sa + B_OFF);
#pragma unroll
                    for (int k = 0; k < C::MMAS_PER_STAGE; ++k)
                        tc::umma<1, kKind>(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (uint32_t)((kb | k) != 0));
                    tc::umma_commit<1>(empty_bar(s));
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
                tc::umma_commit<1>(tfull_bar(buf));
            }
        }
    } else if (warp >= tc::EPI_WARP0) {  // ========== epilogue
        const int w = (warp - tc::EPI_WARP0) & 3;
        const int c0 = ((warp - tc::EPI_WARP0) >> 2) * TCI_CHUNKS;
        int n = 0;
        for (int t = first; t < a.n_items; t += step, ++n) {
            const TcItem it = a.items[t];
            const int buf = n & 1;
            const uint32_t aph = (uint32_t)(n >> 1) & 1u;
            const uint32_t qr = (uint32_t)(w * 32 + lane);
            const int q = qr < it.qrows ? a.qmap[it.qrow0 + qr] : -1;
            const uint32_t row_min = it.list0 + (qr < it.qrows ? a.qskip[it.qrow0 + qr] : 0u);
            const float thr = q >= 0 ? __ldcg(a.lists.thr + q) : __builtin_huge_valf();
            const uint32_t hb = (q >= 0 && a.lists.hist) ? __ldcg(a.lists.hbase + q) : 0u;
            tc::mbar_wait(tfull_bar(buf), aph);
            tc::tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(w * 32) << 16) + (uint32_t)buf * tc::BN;
            tci_epilogue(a, taddr, q, thr, hb, it.row0, it.row_end, row_min, c0);
            tc::tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(tempty_bar(buf));
        }
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc::tcgen05_fence_after();
        tc::tmem_dealloc<1>(tmem_base, tc::TMEM_COLS);
    }
}

// Prepared queries of the tensor-core groups: row r = q32[qmap[r]] (zeros for padding rows), as fp32
// or rounded to fp16 (RNE).
template <typename T>
__global__ void gather_queries_kernel(const float* __restrict__ q32, const int32_t* __restrict__ qmap,
                                      int64_t rows, int dim, T* __restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;   // four elements each
    const int vec = dim / 4;
    if (i >= rows * vec) return;
    const int64_t r = i / vec;
    const int v = (int)(i - r * vec);
    const int q = qmap[r];
    float4 val = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q >= 0) val = __ldg(reinterpret_cast<const float4*>(q32 + (size_t)q * dim) + v);
    if constexpr (sizeof(T) == 4) {
        reinterpret_cast<float4*>(out)[i] = val;
    } else {
        const __half2 lo = __floats2half2_rn(val.x, val.y), hi = __floats2half2_rn(val.z, val.w);
        uint2 packed;
        packed.x = *reinterpret_cast<const uint32_t*>(&lo);
        packed.y = *reinterpret_cast<const uint32_t*>(&hi);
        reinterpret_cast<uint2*>(out)[i] = packed;
    }
}

// fp32 rows -> the fp16 filter copy (RNE), 8 elements per thread.
__global__ void rows_to_f16_kernel(const float* __restrict__ in, __half* __restrict__ out, int64_t n8) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 a = __ldcs(reinterpret_cast<const float4*>(in) + 2 * i);
        const float4 b = __ldcs(reinterpret_cast<const float4*>(in) + 2 * i + 1);
        const __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w);
        const __half2 h2 = __floats2half2_rn(b.x, b.y), h3 = __floats2half2_rn(b.z, b.w);
        uint4 o;
        o.x = *reinterpret_cast<const uint32_t*>(&h0); o.y = *reinterpret_cast<const uint32_t*>(&h1);
        o.z = *reinterpret_cast<const uint32_t*>(&h2); o.w = *reinterpret_cast<const uint32_t*>(&h3);
        reinterpret_cast<uint4*>(out)[i] = o;
    }
}

}  // namespace ivf
}  // namespace fip
