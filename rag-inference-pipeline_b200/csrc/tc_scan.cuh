// K1: tensor-core scan.  S = Q . X^T tile by tile on the 5th-gen tensor cores (tcgen05.mma, TMEM
// accumulators, TMA-staged operands) with the running-threshold filter fused into the epilogue, so
// the [B, N] score matrix never exists in HBM.
//
// Replaces faiss's exhaustive_inner_product_blas (sgemm on 4096 x 1024 blocks + block result
// handler) behind IndexFlatIP.search (/root/reference/src/pipeline/components/faiss_store.py:152).
//
// Tiling: queries are the UMMA M operand (128 rows per CTA = the 128 TMEM lanes), corpus rows the N
// operand (256 columns), K = 128 bytes (one SWIZZLE_128B atom: 64 fp16/bf16 or 32 fp32 elements)
// per pipeline stage.  Putting queries on the lanes makes the fused selection thread-private:
// `tcgen05.ld.32x32b` gives every epilogue thread ONE query and a run of corpus-row scores, so the
// thread compares against that query's threshold held in a register and only the rare survivors
// touch memory.
//
// kCtas = 2 (CTA pair, cta_group::2): two CTAs on one TPC share a corpus tile.  Each loads its own
// 128-query A tile and HALF of the 256-row B tile; one thread of the leader CTA issues
// M = 256 x N = 256 MMAs that read both halves and write each CTA's 128 x 256 accumulator into that
// CTA's TMEM.  Per tile and CTA the L2 -> smem traffic drops from A + B to A + B/2 and a stage
// shrinks from 48 KB to 32 KB, so the ring is 6 deep instead of 4.
//
// Warp roles (640 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane, leader
// CTA only), warp 2 = TMEM allocator, warps 4..19 = epilogue (warp % 4 selects the TMEM
// lane quarter, the four warps of a quarter split the 256 columns).  Pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty double buffer
// (MMA <-> epilogue), persistent static tile schedule (corpus tile major, query tile minor, so the
// query tiles of one corpus tile run back to back and re-read it from L2; HBM sees it once).
#pragma once

#include <type_traits>

#include "common.cuh"
#include "synthetic.cuh"
#include "tc_ptx.cuh"

namespace fip {
namespace tc {

constexpr int BM = 128;          // queries per CTA tile (UMMA M per CTA)
constexpr int BN = 256;          // corpus rows per tile (UMMA N)
constexpr int BK_BYTES = 128;    // K bytes per stage = one SWIZZLE_128B atom
constexpr int BOX_ROWS = 128;    // rows per TMA box (both operands)
constexpr int A_BYTES = BM * BK_BYTES;         // one query operand tile (per split term)
constexpr int B_BYTES = BN * BK_BYTES;         // full corpus tile
constexpr int TMEM_COLS = 512;   // two 256-column fp32 accumulators
constexpr int EPI_WARP0 = 4;
#ifndef FIP_EPI_WARPS
#define FIP_EPI_WARPS 16   // 8 -> 16: the per-tile latency chain (TMEM reads, list reservation) halves; cfg5 2.43 -> 2.24 ms
#endif
constexpr int EPI_WARPS = FIP_EPI_WARPS;   // EPI_WARPS / 4 warps per TMEM lane quarter split the columns
constexpr int NUM_THREADS = (EPI_WARP0 + EPI_WARPS) * 32;

// kFmt: 0 = fp16, 1 = bf16 (kind::f16), 2 = fp32 storage read as tf32 (kind::tf32).
// kSplit = number of terms the fp32 queries are split into (q = q0 + q1): 1 for fp16 / fp32
// storage, 2 for bf16 storage (bf16 hi + lo, 2^-17).  tcgen05.mma needs A and B in the SAME format
// (mixing fp16 x bf16 is an illegal instruction on sm_100a), hence bf16 queries for bf16 corpora.
// The filter runs on these rounded products; survivors are rescored exactly (select.cuh).
// kStage = survivor slots per epilogue thread and tile in the shared-memory staging area of the
// DEFERRED epilogue (0 = off: survivors are stored straight from TMEM in a second pass).
template <int kFmt, int kSplit, int kCtas, int kStage = 0> struct Cfg {
    static constexpr int ELEM_BYTES = kFmt == 2 ? 4 : 2;
    static constexpr int BK = BK_BYTES / ELEM_BYTES;       // elements per stage
    static constexpr int MMAS_PER_STAGE = 4;               // 32 bytes of K per instruction
    static constexpr int B_LOCAL_BYTES = B_BYTES / kCtas;  // this CTA's share of the corpus tile
    static constexpr int STAGE_BYTES = kSplit * A_BYTES + B_LOCAL_BYTES;
    // warps x slots x lanes x 8 B
    static constexpr int STAGING_BYTES = EPI_WARPS * kStage * 32 * 8;
    static constexpr int RING_BYTES = 225 * 1024 - STAGING_BYTES;
    static constexpr int STAGES = RING_BYTES / STAGE_BYTES > 8 ? 8 : RING_BYTES / STAGE_BYTES;
    // operand ring + barriers + staging (the epilogue otherwise works out of TMEM and registers)
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 + STAGING_BYTES;
};

// Epilogue of one thread (= one query) over its EPI_CHUNKS chunks of 32 accumulator columns starting
// at chunk c0: filter against `thr`, two passes over TMEM, no shared memory.
//   pass 1  every chunk is read once; each group of 8 scores is guarded by its max, so "nothing
//           passes" (the common case: ~1e-4 of the scores survive in the main phase) costs ~0.75
//           instructions per score.  A group with a survivor records which columns passed in the
//           chunk's 32-bit mask (registers only).
//   reserve ONE atomicAdd per thread and tile reserves the list slots of all its survivors (the
//           round trip to L2 is paid once per tile, not once per chunk, and overlaps pass 2's TMEM
//           reads).
//   pass 2  chunks in which some lane of the warp had a survivor are read from TMEM again (the
//           accumulator is still live) and the survivors are stored straight from registers as raw
//           keys with predicated stores.
constexpr int EPI_THREADS = EPI_WARPS * 32;
constexpr int EPI_CHUNKS = BN / 32 / (EPI_WARPS / 4);        // chunks per epilogue warp and tile

// The chunk loops are NOT unrolled: one copy of each pass keeps the kernel inside the instruction
// cache (a fully unrolled version lost 2x on `no_inst` fetch stalls); the per-chunk masks live in
// two 64-bit registers addressed with shifts.
__device__ __forceinline__ void epilogue_filter(const ScanArgs& a, uint32_t taddr, int q, float thr,
                                                uint32_t hb, uint32_t tile_row0, int c0) {
    static_assert(EPI_CHUNKS <= 4, "masks are packed into two 64-bit registers");
    uint64_t m01 = 0ull, m23 = 0ull;   // masks of chunks 0,1 and 2,3
    uint32_t n = 0;
#pragma unroll 1
    for (int ci = 0; ci < EPI_CHUNKS; ++ci) {
        const int c = c0 + ci;
        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
        if (row_base >= a.row_end) break;  // warp-uniform
        float v[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
        if (row_base + 32 > a.row_end) {  // ragged last chunk (warp-uniform)
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (row_base + j >= a.row_end) v[j] = kNegInf;
        }
        uint32_t mask = 0u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            float mx = fmaxf(fmaxf(v[8 * g], v[8 * g + 1]), fmaxf(v[8 * g + 2], v[8 * g + 3]));
            mx = fmaxf(mx, fmaxf(fmaxf(v[8 * g + 4], v[8 * g + 5]), fmaxf(v[8 * g + 6], v[8 * g + 7])));
            if (mx > thr) {
#pragma unroll
                for (int j = 0; j < 8; ++j) mask_or_if_gt(mask, v[8 * g + j], thr, 1u << (8 * g + j));
            }
        }
        const uint64_t shifted = (uint64_t)mask << (32 * (ci & 1));
        if (ci & 2) m23 |= shifted; else m01 |= shifted;
        n += (uint32_t)__popc(mask);
    }
    // n != 0 implies an active query (inactive lanes carry thr = +inf).  A reservation that does
    // not fit writes nothing: cnt > cap makes the select kernel flag the query as overflowed and
    // the host redoes it in safe mode.
    uint32_t slot = 0u;
    bool fits = false;
    if (n) {
        slot = atomicAdd(a.cnt + q, n);
        fits = slot + n <= a.cap;
    }
    if (!__any_sync(0xFFFFFFFFu, n != 0u)) return;
    uint64_t* list = a.cand + (size_t)q * a.cap + slot;
    uint32_t* hrow = a.hist ? a.hist + (size_t)q * kHistBins : nullptr;
#pragma unroll 1
    for (int ci = 0; ci < EPI_CHUNKS; ++ci) {
        const uint32_t cmask = (uint32_t)(((ci & 2) ? m23 : m01) >> (32 * (ci & 1)));
        if (!__any_sync(0xFFFFFFFFu, cmask != 0u)) continue;  // warp-uniform
        const uint32_t mask = fits ? cmask : 0u;
        const int c = c0 + ci;
        float v[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
        const uint32_t inv_row0 = 0xFFFFFFFFu - (tile_row0 + (uint32_t)c * 32);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if ((mask >> (8 * g)) & 0xFFu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t bits = __float_as_uint(v[8 * g + j]);
                    const uint32_t take = mask & (1u << (8 * g + j));
                    list_append_if(list, inv_row0 - (uint32_t)(8 * g + j), bits, take);
                    if (take && hrow) atomicAdd(hrow + hist_bucket(bits, hb), 1u);
                }
            }
        }
    }
    // live threshold: a list that just grew past a multiple of kRefreshEvery entries is re-read
    if (a.hist)
        refresh_thresholds_warp(a, fits && ((slot + n) / kRefreshEvery != slot / kRefreshEvery), q,
                                (int)(threadIdx.x & 31));
}

// ---------------------------------------------------------------------------------------------
// DEFERRED epilogue (kStage > 0; chosen for large k, where almost every tile has survivors).
// With k = 1024 about half of all (warp, tile) pairs hold a survivor, so in the two-pass form
// every tile pays  pass 1 -> atomicAdd round trip to L2 -> pass 2 (TMEM re-read + stores)  BEFORE
// its accumulator is handed back: the tensor pipe waits (70 % active in the k = 1024 profile,
// profiles/r02_ncu_full_cfg5_before_deferred_epilogue.csv).  Here pass 1 also copies the survivors'
// keys into a small per-thread staging area in shared memory; the accumulator is released right
// after pass 1, and only then does the warp reserve list slots (the atomicAdd round trip) and
// copy the staged keys to the candidate list (epilogue_flush) -- off the tensor pipe's critical
// path, in time the epilogue warps otherwise spend waiting for the next accumulator.
// A tile in which some thread of the warp has more than kStage survivors (the first tiles after
// the dense phase) takes the two-pass route for that warp.
struct Pending {          // a tile's survivors, staged but not yet in the candidate list
    uint32_t n = 0;       // this thread's survivors
    int q = 0;
    uint32_t hb = 0;
    uint32_t stg = 0;     // shared-memory address of this thread's slot 0
};

__device__ __forceinline__ void stage_append_if(uint32_t& addr, uint32_t lo, uint32_t hi, uint32_t take) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %3, 0;\n\t"
        "@p st.shared.v2.u32 [%0], {%1, %2};\n\t"
        "@p add.u32 %0, %0, 256;\n\t"      // next slot of the same lane: 32 lanes x 8 bytes further
        "}"
        : "+r"(addr)
        : "r"(lo), "r"(hi), "r"(take)
        : "memory");
}

// Writes a pending tile's staged keys to its query's candidate list (all 32 lanes must call).
template <int kStage>
__device__ __forceinline__ void epilogue_flush(const ScanArgs& a, const Pending& p) {
    if (!__any_sync(0xFFFFFFFFu, p.n != 0u)) return;
    uint32_t slot = 0u;
    if (p.n) slot = atomicAdd(a.cnt + p.q, p.n);
    const bool fits = p.n != 0u && slot + p.n <= a.cap;   // else: overflow, flagged by the select pass
    uint64_t* list = a.cand + (size_t)p.q * a.cap + slot;
    uint32_t* hrow = a.hist ? a.hist + (size_t)p.q * kHistBins : nullptr;
#pragma unroll
    for (int s = 0; s < kStage; ++s) {
        if (fits && (uint32_t)s < p.n) {
            uint32_t lo, hi;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(p.stg + 256u * s));
            asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(list + s), "r"(lo), "r"(hi) : "memory");
            if (hrow) atomicAdd(hrow + hist_bucket(hi, p.hb), 1u);
        }
    }
    if (a.hist)
        refresh_thresholds_warp(a, fits && ((slot + p.n) / kRefreshEvery != slot / kRefreshEvery), p.q,
                                (int)(threadIdx.x & 31));
}

// Pass 1 + staging.  Returns what is left pending (n == 0 when nothing survived, or when the warp
// took the two-pass route and already stored everything).
template <int kStage>
__device__ __forceinline__ Pending epilogue_filter_deferred(const ScanArgs& a, uint32_t taddr, int q, float thr,
                                                            uint32_t hb, uint32_t tile_row0, int c0,
                                                            uint32_t stg_addr) {
    static_assert(EPI_CHUNKS <= 4, "masks are packed into two 64-bit registers");
    uint64_t m01 = 0ull, m23 = 0ull;
    uint32_t n = 0;
    uint32_t stg = stg_addr;
    bool ovf = false;
#pragma unroll 1
    for (int ci = 0; ci < EPI_CHUNKS; ++ci) {
        const int c = c0 + ci;
        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
        if (row_base >= a.row_end) break;  // warp-uniform
        float v[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
        if (row_base + 32 > a.row_end) {  // ragged last chunk (warp-uniform)
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (row_base + j >= a.row_end) v[j] = kNegInf;
        }
        uint32_t mask = 0u;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            float mx = fmaxf(fmaxf(v[8 * g], v[8 * g + 1]), fmaxf(v[8 * g + 2], v[8 * g + 3]));
            mx = fmaxf(mx, fmaxf(fmaxf(v[8 * g + 4], v[8 * g + 5]), fmaxf(v[8 * g + 6], v[8 * g + 7])));
            if (mx > thr) {
#pragma unroll
                for (int j = 0; j < 8; ++j) mask_or_if_gt(mask, v[8 * g + j], thr, 1u << (8 * g + j));
            }
        }
        if (mask) {
            const uint32_t cnt = (uint32_t)__popc(mask);
            if (!ovf && n + cnt <= (uint32_t)kStage) {
                const uint32_t inv_row0 = 0xFFFFFFFFu - row_base;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    if ((mask >> (8 * g)) & 0xFFu) {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            stage_append_if(stg, inv_row0 - (uint32_t)(8 * g + j), __float_as_uint(v[8 * g + j]),
                                            mask & (1u << (8 * g + j)));
                    }
                }
            } else {
                ovf = true;
            }
            n += cnt;
        }
        const uint64_t shifted = (uint64_t)mask << (32 * (ci & 1));
        if (ci & 2) m23 |= shifted; else m01 |= shifted;
    }
    Pending p;
    p.q = q; p.hb = hb; p.stg = stg_addr;
    if (!__any_sync(0xFFFFFFFFu, ovf)) {
        p.n = n;      // slots are reserved and the keys copied after the accumulator's release
        return p;
    }
    // two-pass route for this warp and tile (some thread has more survivors than staging slots)
    uint32_t slot = 0u;
    if (n) slot = atomicAdd(a.cnt + q, n);
    const bool fits = n != 0u && slot + n <= a.cap;
    uint64_t* list = a.cand + (size_t)q * a.cap + slot;
    uint32_t* hrow = a.hist ? a.hist + (size_t)q * kHistBins : nullptr;
#pragma unroll 1
    for (int ci = 0; ci < EPI_CHUNKS; ++ci) {
        const uint32_t cmask = (uint32_t)(((ci & 2) ? m23 : m01) >> (32 * (ci & 1)));
        if (!__any_sync(0xFFFFFFFFu, cmask != 0u)) continue;  // warp-uniform
        const uint32_t mask = fits ? cmask : 0u;
        const int c = c0 + ci;
        float v[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
        const uint32_t inv_row0 = 0xFFFFFFFFu - (tile_row0 + (uint32_t)c * 32);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            if ((mask >> (8 * g)) & 0xFFu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t bits = __float_as_uint(v[8 * g + j]);
                    const uint32_t take = mask & (1u << (8 * g + j));
                    list_append_if(list, inv_row0 - (uint32_t)(8 * g + j), bits, take);
                    if (take && hrow) atomicAdd(hrow + hist_bucket(bits, hb), 1u);
                }
            }
        }
    }
    if (a.hist)
        refresh_thresholds_warp(a, fits && ((slot + n) / kRefreshEvery != slot / kRefreshEvery), q,
                                (int)(threadIdx.x & 31));
    p.n = 0u;
    return p;
}

// Phase 0: every score of the tile is written at slot = row - row_begin (no threshold yet).
__device__ __forceinline__ void epilogue_dense(const ScanArgs& a, uint32_t taddr, int q, bool active,
                                               uint32_t tile_row0, int c0, int c1) {
#pragma unroll 1
    for (int c = c0; c < c1; ++c) {
        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
        if (row_base >= a.row_end) break;
        float v[32];
        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
        uint64_t* dst = a.cand + (size_t)q * a.cap + (row_base - a.row_begin);
        const uint32_t inv_row = 0xFFFFFFFFu - row_base;
#pragma unroll
        for (int j = 0; j < 32; ++j)
            st_global_v2_pred(dst + j, inv_row - (uint32_t)j, __float_as_uint(v[j]),
                              active && row_base + j < a.row_end);
    }
}

// tmap_q covers the prepared query matrix [kSplit * nq_pad, dim] (term t of query q at row
// t * nq_pad + q); tmap_x the corpus shard [n_rows, dim].  Both use [BK x 128]-element boxes.
// Launch with a (2,1,1) cluster when kCtas == 2; gridDim.x must then be even.
template <int kFmt, int kSplit, int kCtas, int kStage = 0>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_scan_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
               ScanArgs a, int n_qtiles) {
    using C = Cfg<kFmt, kSplit, kCtas, kStage>;
    constexpr int STAGES = C::STAGES;
    constexpr int STAGE_BYTES = C::STAGE_BYTES;
    constexpr int B_OFF = kSplit * A_BYTES;
    constexpr int B_BOXES = C::B_LOCAL_BYTES / (BOX_ROWS * BK_BYTES);  // 2 (single CTA) or 1 (pair)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = base + STAGES * STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
    auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    const uint32_t stg_base = bar_base + 256u;   // deferred epilogue: [EPI_WARPS][kStage][32 lanes] keys
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t cta_rank = kCtas == 2 ? cluster_ctarank() : 0u;
    const bool leader = cta_rank == 0;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        // tempty: one arrival per epilogue warp of every CTA of the group (lives in the leader)
        for (int b = 0; b < 2; ++b) { mbar_init(tfull_bar(b), 1); mbar_init(tempty_bar(b), EPI_WARPS * kCtas); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<kCtas>(tmem_slot, TMEM_COLS);
    tcgen05_fence_before();
    if (kCtas == 2) cluster_sync_all(); else __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    const int nkb = (a.dim * C::ELEM_BYTES) / BK_BYTES;
    const int n_ctiles = (int)((a.row_end - a.row_begin + BN - 1) / BN);
    // A "group tile" is one corpus tile x kCtas adjacent query tiles (one per CTA of the group).
    const int n_qgroups = n_qtiles / kCtas;
    const int total_tiles = n_ctiles * n_qgroups;
    const int group = (int)blockIdx.x / kCtas, n_groups = (int)gridDim.x / kCtas;
    // Static schedule: group g owns the CONTIGUOUS range [t_begin, t_end) of the (corpus tile major,
    // query group minor) order, so it walks all query tiles of one corpus tile back to back.
    const int t_begin = (int)(((long long)total_tiles * group) / n_groups);
    const int t_end = (int)(((long long)total_tiles * (group + 1)) / n_groups);

    if (warp == 0) {
        if (lane == 0) {  // ===================== TMA producer (every CTA loads its own operands)
            const uint32_t full0 = kCtas == 2 ? mapa_shared(full_bar(0), 0) : full_bar(0);
            int s = 0;
            uint32_t ph = 0;
            for (int t = t_begin; t < t_end; ++t) {
                const int ct = t / n_qgroups, qg = t - ct * n_qgroups;
                const int qt = qg * kCtas + (int)cta_rank;
                const int row0 = (int)a.row_begin + ct * BN + (int)cta_rank * C::B_LOCAL_BYTES / BK_BYTES;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t sa = base + s * STAGE_BYTES;
                    if (kCtas == 1) {
                        mbar_arrive_expect_tx(full_bar(s), STAGE_BYTES);
#pragma unroll
                        for (int part = 0; part < kSplit; ++part)
                            tma_load_2d(sa + part * A_BYTES, &tmap_q, full_bar(s), kb * C::BK,
                                        part * n_qtiles * BM + qt * BM);
#pragma unroll
                        for (int b = 0; b < B_BOXES; ++b)
                            tma_load_2d(sa + B_OFF + b * BOX_ROWS * BK_BYTES, &tmap_x, full_bar(s),
                                        kb * C::BK, row0 + b * BOX_ROWS);
                    } else {
                        // the leader's barrier collects the bytes of BOTH CTAs' loads
                        if (leader) mbar_arrive_expect_tx(full_bar(s), 2 * STAGE_BYTES);
                        const uint32_t fb = full0 + 8u * s;
#pragma unroll
                        for (int part = 0; part < kSplit; ++part)
                            tma_load_2d_pair(sa + part * A_BYTES, &tmap_q, fb, kb * C::BK,
                                             part * n_qtiles * BM + qt * BM);
                        tma_load_2d_pair(sa + B_OFF, &tmap_x, fb, kb * C::BK, row0);
                    }
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && leader) {  // ===================== MMA issuer
            constexpr uint32_t idesc = make_idesc(kFmt, kFmt, BM * kCtas, BN);
            constexpr int kKind = kFmt == 2 ? 1 : 0;
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int t = t_begin; t < t_end; ++t, ++it) {
                const int buf = it & 1;
                const uint32_t aph = (uint32_t)(it >> 1) & 1u;
                mbar_wait<kCtas == 2>(tempty_bar(buf), aph ^ 1u);  // epilogues drained this accumulator
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)buf * BN;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(full_bar(s), ph);
                    tcgen05_fence_after();
                    const uint32_t sa = base + s * STAGE_BYTES;
                    const uint64_t bdesc = make_sw128_desc(sa + B_OFF);
#pragma unroll
                    for (int part = 0; part < kSplit; ++part) {
                        const uint64_t adesc = make_sw128_desc(sa + part * A_BYTES);
#pragma unroll
                        for (int k = 0; k < C::MMAS_PER_STAGE; ++k) {
                            // advance 32 bytes inside the swizzle atom: +2 (16-byte units)
                            umma<kCtas, kKind>(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc,
                                               (uint32_t)((kb | k | part) != 0));
                        }
                    }
                    umma_commit<kCtas>(empty_bar(s));  // frees the smem stage (in both CTAs)
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
                umma_commit<kCtas>(tfull_bar(buf));  // accumulator complete -> epilogue(s)
            }
        }
    } else if (warp >= EPI_WARP0) {  // ========== epilogue: TMEM -> registers -> threshold filter
        // TMEM lane quarter = warp % 4 (hardware rule); the two warps of a quarter split the columns
        const int w = (warp - EPI_WARP0) & 3;
        const int c0 = ((warp - EPI_WARP0) >> 2) * EPI_CHUNKS;
        const uint32_t tempty0 = kCtas == 2 ? mapa_shared(tempty_bar(0), 0) : tempty_bar(0);
        // Threshold and histogram origin of a thread's query come from L2 (thresholds are raised
        // live by whoever appends candidates) and are fetched ONE TILE AHEAD, so the round trip is
        // never on the epilogue's critical path.  Tile coordinates advance incrementally (no
        // integer division per tile).
        auto query_at = [&](int qg_) { return (qg_ * kCtas + (int)cta_rank) * BM + w * 32 + lane; };
        auto load_thr = [&](int qq) { return qq < a.nq ? __ldcg(a.thr + qq) : __builtin_huge_valf(); };
        auto load_hb = [&](int qq) { return (qq < a.nq && a.hist) ? __ldcg(a.hbase + qq) : 0u; };
        int ct = t_begin / n_qgroups, qg = t_begin - ct * n_qgroups;
        float thr_next = 0.f;
        uint32_t hb_next = 0u;
        if (t_begin < t_end) { thr_next = load_thr(query_at(qg)); hb_next = load_hb(query_at(qg)); }
        const uint32_t stg_warp = stg_base + (uint32_t)(warp - EPI_WARP0) * (kStage * 32 * 8) + (uint32_t)lane * 8u;
        int it = 0;
        for (int t = t_begin; t < t_end; ++t, ++it) {
            const int buf = it & 1;
            const uint32_t aph = (uint32_t)(it >> 1) & 1u;
            const int q = query_at(qg);
            const bool active = q < a.nq;
            const float thr = thr_next;
            const uint32_t hb = hb_next;
            const uint32_t tile_row0 = a.row_begin + (uint32_t)ct * BN;
            if (++qg == n_qgroups) { qg = 0; ++ct; }
            if (t + 1 < t_end) { thr_next = load_thr(query_at(qg)); hb_next = load_hb(query_at(qg)); }
            mbar_wait(tfull_bar(buf), aph);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(w * 32) << 16) + (uint32_t)buf * BN;
            Pending cur;
            if (a.dense) epilogue_dense(a, taddr, q, active, tile_row0, c0, c0 + EPI_CHUNKS);
            else if (kStage > 0)
                cur = epilogue_filter_deferred<kStage>(a, taddr, q, thr, hb, tile_row0, c0, stg_warp);
            else epilogue_filter(a, taddr, q, thr, hb, tile_row0, c0);
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (kCtas == 2) mbar_arrive_cluster(tempty0 + 8u * buf);
                else mbar_arrive(tempty_bar(buf));
            }
            // the accumulator is back with the MMA warp: reserve list slots and write out what this
            // tile staged
            if (kStage > 0) epilogue_flush<(kStage > 0 ? kStage : 1)>(a, cur);
        }
    }

    tcgen05_fence_before();
    if (kCtas == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        tmem_dealloc<kCtas>(tmem_base, TMEM_COLS);
    }
}

// Queries fp32 [nq, dim] -> kSplit terms [kSplit][nq_pad][dim] in the MMA operand type T, with
// q ~= sum of the terms.  fp16: a per-query power-of-two scale first moves the largest component
// into [0.5, 1) so the fp16 range is never an issue (exact to undo: out_scale[q] = 2^e); bf16 and
// fp32 (tf32 operand: the tensor core ignores the low 13 mantissa bits) keep scale 1.
// Rows >= nq are zero.  One warp per query row.
//
// Also emits the query's certificate radius eps[q]: a rigorous bound on |approx - exact| score for
// ANY stored row, in unscaled score units:
//     eps = ( ||q - sum(terms)|| + rel * ||q|| ) * max_r ||x_r||
// where `rel` covers what software cannot see: fp32 accumulation inside the tensor core
// (dim * 2^-23) and, for tf32, the truncation of BOTH operands to 10 mantissa bits (2^-9 + 2^-20).
template <typename T, int kSplit>
__global__ void prepare_queries_kernel(const float* __restrict__ q32, T* __restrict__ qt,
                                       float* __restrict__ out_scale, float* __restrict__ out_eps,
                                       const float* __restrict__ max_norm2, float rel, int nq,
                                       int nq_pad, int dim, ListInit init) {
    // fused list / flag reset (32 threads per padded query row >= 2 + 2 nq elements)
    init_lists_thread(init, (int)(blockIdx.x * blockDim.x + threadIdx.x));
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= nq_pad) return;
    if (row >= nq) {
        for (int t = 0; t < kSplit; ++t)
            for (int c = lane; c < dim; c += 32)
                qt[((size_t)t * nq_pad + row) * dim + c] = to_storage<T>(0.f);
        return;
    }
    const float* src = q32 + (size_t)row * dim;
    float up = 1.0f, down = 1.0f;
    if (std::is_same<T, __half>::value) {
        float mx = 0.f;
        for (int c = lane; c < dim; c += 32) mx = fmaxf(mx, fabsf(src[c]));
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, off));
        int e = 0;
        if (mx > 0.f && mx < __builtin_huge_valf()) (void)frexpf(mx, &e);  // mx = m * 2^e
        e = max(-100, min(100, e));
        up = ldexpf(1.0f, -e);
        down = ldexpf(1.0f, e);
    }
    float res2 = 0.f, q2 = 0.f;
    for (int c = lane; c < dim; c += 32) {
        float rem = src[c] * up;
        q2 = fmaf(rem, rem, q2);
#pragma unroll
        for (int t = 0; t < kSplit; ++t) {
            T h = to_storage<T>(rem);
            qt[((size_t)t * nq_pad + row) * dim + c] = h;
            rem -= from_storage<T>(h);
        }
        res2 = fmaf(rem, rem, res2);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        res2 += __shfl_xor_sync(0xFFFFFFFFu, res2, off);
        q2 += __shfl_xor_sync(0xFFFFFFFFu, q2, off);
    }
    if (lane == 0) {
        out_scale[row] = down;
        // 1.01: slack for the fp32 rounding of the norms themselves
        out_eps[row] = 1.01f * (sqrtf(res2) + rel * sqrtf(q2)) * down * sqrtf(*max_norm2);
    }
}

}  // namespace tc
}  // namespace fip
