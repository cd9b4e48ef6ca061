// K1: tensor-core scan.  S = Q . X^T tile by tile on the 5th-gen tensor cores (tcgen05.mma, TMEM
// accumulators, TMA-staged operands) with the running-threshold filter fused into the epilogue, so
// the [B, N] score matrix never exists in HBM.
//
// Replaces faiss's exhaustive_inner_product_blas (sgemm on 4096 x 1024 blocks + block result
// handler) behind IndexFlatIP.search (/root/reference/src/pipeline/components/faiss_store.py:152).
//
// Tiling: queries are the UMMA M operand (128 rows = the 128 TMEM lanes), corpus rows the N
// operand (256 columns), K = 64 elements (one 128-byte swizzle atom) per pipeline stage.  Putting
// queries on the lanes makes the fused selection thread-private: `tcgen05.ld.32x32b` gives every
// epilogue thread ONE query and a run of corpus-row scores, so the thread compares against that
// query's threshold held in a register and only the rare survivors touch memory.
//
// Warp roles (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane),
// warp 2 = TMEM allocator, warps 4..7 = epilogue (warp % 4 selects the TMEM lane quarter).
// Pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty double buffer (MMA <-> epilogue),
// persistent static tile schedule (tile = corpus tile major, query tile minor, so CTAs running
// concurrently share corpus tiles through L2 and HBM reads the corpus once).
#pragma once

#include <cuda.h>

#include <type_traits>

#include "common.cuh"
#include "synthetic.cuh"

namespace fip {
namespace tc {

constexpr int BM = 128;   // queries per tile  (UMMA M)
constexpr int BN = 256;   // corpus rows per tile (UMMA N)
constexpr int BK = 64;    // K elements per stage: 64 * 2 B = 128 B = one SWIZZLE_128B atom
constexpr int UMMA_K = 16;
constexpr int A_BYTES = BM * BK * 2;   // one query operand tile (per split term)
constexpr int B_BYTES = BN * BK * 2;
constexpr int TMEM_COLS = 512;  // two 256-column fp32 accumulators
constexpr int NUM_THREADS = 256;
constexpr int EPI_WARP0 = 4;
// kSplit = number of 16-bit terms the fp32 queries are split into (q = q0 + q1 ...): 1 for fp16
// storage (fp16 queries, 2^-11 rounding, repaired by the exact rescoring pass), 2 for bf16 storage
// (bf16 hi + lo, 2^-17).  tcgen05.mma kind::f16 needs A and B in the SAME 16-bit format
// (mixing fp16 x bf16 is an illegal instruction on sm_100a), hence bf16 queries for bf16 corpora.
template <int kSplit> struct Cfg {
    static constexpr int STAGE_BYTES = kSplit * A_BYTES + B_BYTES;
    static constexpr int STAGES = kSplit == 1 ? 4 : 3;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
    return ok != 0;
}
// Blocking wait with a watchdog: a pipeline bug traps (CUDA error on the host) instead of hanging
// the GPU.  No legitimate wait in this kernel lasts anywhere near 2^32 cycles (> 2 s).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > (1ll << 32)) {
            printf("flatip_b200: mbarrier wait timed out (block %d thread %d bar 0x%x parity %u)\n",
                   (int)blockIdx.x, (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)map) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}
// tcgen05.commit: arrive on `bar` once every MMA issued so far by this thread has completed.
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, 128 x BN x 16, fp32 accumulate.
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                         uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor for a K-major operand tile laid out by TMA with SWIZZLE_128B:
// rows of 128 bytes, 8-row groups 1024 bytes apart (SBO); LBO is unused for swizzled K-major.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);  // start address       bits [0,14)
    d |= (uint64_t)1 << 16;                        // leading byte offset bits [16,30) (ignored)
    d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset  bits [32,46)
    d |= (uint64_t)1 << 46;                        // descriptor version 1 (sm_100)
    d |= (uint64_t)2 << 61;                        // layout type SWIZZLE_128B
    return d;
}
// Instruction descriptor, kind::f16: fp32 accumulate, both operands K-major.
__host__ __device__ constexpr uint32_t make_idesc(int fmt_a, int fmt_b, int m, int n) {
    return (1u << 4) | ((uint32_t)fmt_a << 7) | ((uint32_t)fmt_b << 10) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// kFmt: 0 = fp16, 1 = bf16 (UMMA F16F32Format) for both operands.  tmap_q covers the prepared query
// matrix [kSplit * nq_pad, dim] (term t of query q at row t * nq_pad + q).
template <int kFmt, int kSplit>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_scan_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_x,
               ScanArgs a, int n_qtiles) {
    constexpr int STAGES = Cfg<kSplit>::STAGES;
    constexpr int STAGE_BYTES = Cfg<kSplit>::STAGE_BYTES;
    constexpr int B_OFF = kSplit * A_BYTES;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = base + STAGES * STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
    auto tfull_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + b); };
    auto tempty_bar = [&](int b) { return bar_base + 8u * (2 * STAGES + 2 + b); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);
    volatile uint32_t* tmem_slot_ptr =
        reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_q);
        tma_prefetch_desc(&tmap_x);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(tfull_bar(b), 1); mbar_init(tempty_bar(b), 128); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    const int nkb = a.dim / BK;
    const int n_ctiles = (int)((a.row_end - a.row_begin + BN - 1) / BN);
    const int total_tiles = n_ctiles * n_qtiles;
    // Static schedule: CTA b owns the CONTIGUOUS tile range [t_begin, t_end) of the (corpus tile
    // major, query tile minor) order, so it walks all query tiles of one corpus tile back to back:
    // the corpus tile is fetched from HBM once and re-read from L2 for the other query tiles.
    const int t_begin = (int)(((long long)total_tiles * blockIdx.x) / gridDim.x);
    const int t_end = (int)(((long long)total_tiles * (blockIdx.x + 1)) / gridDim.x);

    if (warp == 0) {
        if (lane == 0) {  // ===================== TMA producer
            int s = 0;
            uint32_t ph = 0;
            for (int t = t_begin; t < t_end; ++t) {
                const int ct = t / n_qtiles, qt = t - ct * n_qtiles;
                const int row0 = (int)a.row_begin + ct * BN;
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    mbar_arrive_expect_tx(full_bar(s), STAGE_BYTES);
                    const uint32_t sa = base + s * STAGE_BYTES;
#pragma unroll
                    for (int part = 0; part < kSplit; ++part)
                        tma_load_2d(sa + part * A_BYTES, &tmap_q, full_bar(s), kb * BK,
                                    part * n_qtiles * BM + qt * BM);
                    tma_load_2d(sa + B_OFF, &tmap_x, full_bar(s), kb * BK, row0);
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {  // ===================== MMA issuer
            constexpr uint32_t idesc = make_idesc(kFmt, kFmt, BM, BN);
            int s = 0, it = 0;
            uint32_t ph = 0;
            for (int t = t_begin; t < t_end; ++t, ++it) {
                const int buf = it & 1;
                const uint32_t aph = (uint32_t)(it >> 1) & 1u;
                mbar_wait(tempty_bar(buf), aph ^ 1u);  // epilogue has drained this accumulator
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
                        for (int k = 0; k < BK / UMMA_K; ++k) {
                            // advance 16 elements = 32 bytes inside the swizzle atom: +2 (16-byte units)
                            umma_f16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc,
                                     (uint32_t)((kb | k | part) != 0));
                        }
                    }
                    umma_commit(empty_bar(s));  // frees the smem stage when these MMAs retire
                    if (++s == STAGES) { s = 0; ph ^= 1u; }
                }
                umma_commit(tfull_bar(buf));  // accumulator complete -> epilogue
            }
        }
    } else if (warp >= EPI_WARP0) {  // ========== epilogue: TMEM -> registers -> threshold filter
        const int w = warp - EPI_WARP0;  // == warp % 4: TMEM lanes [32w, 32w+32)
        int it = 0;
        for (int t = t_begin; t < t_end; ++t, ++it) {
            const int ct = t / n_qtiles, qt = t - ct * n_qtiles;
            const int buf = it & 1;
            const uint32_t aph = (uint32_t)(it >> 1) & 1u;
            const int q = qt * BM + w * 32 + lane;
            const bool active = q < a.nq;
            const float thr = active ? a.thr[q] : __builtin_huge_valf();
            const uint32_t tile_row0 = a.row_begin + (uint32_t)ct * BN;
            mbar_wait(tfull_bar(buf), aph);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(w * 32) << 16) + (uint32_t)buf * BN;
            if (a.dense) {
#pragma unroll 1
                for (int c = 0; c < BN / 32; ++c) {
                    float v[32];
                    tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
                    const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
                    if (row_base >= a.row_end) continue;  // warp-uniform
                    const bool full_chunk = row_base + 32 <= a.row_end;
                    if (active) {
                        uint64_t* dst = a.cand + (size_t)q * a.cap + (row_base - a.row_begin);
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (full_chunk || row_base + j < a.row_end)
                                dst[j] = make_key(v[j], row_base + j);
                    }
                }
            } else {
                // Pass 1: count this thread's survivors in the tile (registers only).
                uint32_t npass = 0;
#pragma unroll 1
                for (int c = 0; c < BN / 32; ++c) {
                    float v[32];
                    tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
                    const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
                    if (row_base >= a.row_end) continue;  // warp-uniform
                    const bool full_chunk = row_base + 32 <= a.row_end;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        npass += (v[j] > thr && (full_chunk || row_base + j < a.row_end)) ? 1u : 0u;
                }
                // Pass 2 (only when some thread of the warp has survivors): ONE atomic per thread
                // reserves its slots, then the accumulator is re-read and the survivors written.
                // (One returning atomic per survivor would serialise ~1 us round trips.)
                if (__any_sync(0xFFFFFFFFu, npass != 0)) {
                    uint32_t slot = npass ? atomicAdd(a.cnt + q, npass) : 0u;
                    uint64_t* dst = a.cand + (size_t)q * a.cap;
#pragma unroll 1
                    for (int c = 0; c < BN / 32; ++c) {
                        float v[32];
                        tmem_ld_32x32b_x32(taddr + (uint32_t)c * 32, v);
                        const uint32_t row_base = tile_row0 + (uint32_t)c * 32;
                        if (row_base >= a.row_end) continue;  // warp-uniform
                        const bool full_chunk = row_base + 32 <= a.row_end;
                        float mx = v[0];
#pragma unroll
                        for (int j = 1; j < 32; ++j) mx = fmaxf(mx, v[j]);
                        if (mx > thr) {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (v[j] > thr && (full_chunk || row_base + j < a.row_end)) {
                                    if (slot < a.cap) dst[slot] = make_key(v[j], row_base + j);
                                    ++slot;
                                }
                        }
                    }
                }
            }
            tcgen05_fence_before();
            mbar_arrive(tempty_bar(buf));
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// Queries fp32 [nq, dim] -> kSplit 16-bit terms [kSplit][nq_pad][dim] with q ~= sum of the terms.
// fp16: a per-query power-of-two scale first moves the largest component into [0.5, 1) so the
// fp16 range is never an issue (exact to undo: out_scale[q] = 2^e); bf16 keeps scale 1.
// Rows >= nq are zero.  One warp per query row.
template <typename T16, int kSplit>
__global__ void prepare_queries_kernel(const float* __restrict__ q32, T16* __restrict__ q16,
                                       float* __restrict__ out_scale, int nq, int nq_pad, int dim) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= nq_pad) return;
    if (row >= nq) {
        for (int t = 0; t < kSplit; ++t)
            for (int c = lane; c < dim; c += 32)
                q16[((size_t)t * nq_pad + row) * dim + c] = to_storage<T16>(0.f);
        return;
    }
    const float* src = q32 + (size_t)row * dim;
    float up = 1.0f, down = 1.0f;
    if (sizeof(T16) == 2 && std::is_same<T16, __half>::value) {
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
    for (int c = lane; c < dim; c += 32) {
        float rem = src[c] * up;
#pragma unroll
        for (int t = 0; t < kSplit; ++t) {
            T16 h = to_storage<T16>(rem);
            q16[((size_t)t * nq_pad + row) * dim + c] = h;
            rem -= from_storage<T16>(h);
        }
    }
    if (lane == 0) out_scale[row] = down;
}

}  // namespace tc
}  // namespace fip
