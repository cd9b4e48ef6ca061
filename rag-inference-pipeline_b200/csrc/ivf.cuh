// K5: IVF-Flat (L2) operator -- the index type the reference's own builder writes
// (/root/reference/scripts/create_test_docs.py:83-104: IndexIVFFlat(IndexFlatL2(768), 768, 4096,
// METRIC_L2), nprobe = 64) and FAISSStore.load configures (faiss_store.py:84-93, `nprobe`), behind
// the same `search(queries, k) -> (distances, ids)` boundary (faiss_store.py:152).  SURVEY 8(f4).
//
// Replaces faiss's IndexIVF::search_preassigned + IVFFlatScannerL2 (per query: the nprobe nearest
// centroids, then a scan of those inverted lists with a k-heap) with a LIST-MAJOR scan: the
// (query, list) pairs are inverted into (list -> queries probing it), so an inverted list is read
// from HBM once per group of up to QB queries instead of once per query.
//
// Layout in HBM: the vectors are stored sorted by list (`x[pos]`, fp32 [N, dim]; list l owns
// positions [off[l], off[l+1]) in add order), with ids[pos] (int64) and xnorm[pos] = ||x||^2.
// Scores are `2 <q, x> - ||x||^2` (= ||q||^2 - ||q - x||^2: larger is nearer), exact fp32 FMA, so
// the flat path's candidate lists, live thresholds and radix select (common.cuh, select.cuh) are
// reused unchanged; `ivf_finalize_kernel` turns (score, pos) into (squared L2 distance, id).
#pragma once

#include "common.cuh"
#include "tc_ptx.cuh"

namespace fip {
namespace ivf {

using tc::fence_barrier_init;
using tc::mbar_arrive_expect_tx;
using tc::mbar_init;
using tc::mbar_wait;
using tc::smem_u32;
using tc::tma_load_2d;
using tc::tma_prefetch_desc;

// ------------------------------------------------------------------------------------------
// Squared norms of the rows of a row-major fp32 matrix.  One warp per row.
__global__ void row_norm2_kernel(const float* __restrict__ x, int64_t n, int dim, float* __restrict__ out) {
    const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    const float* p = x + (size_t)row * dim;
    float s = 0.f;
    for (int c = lane; c < dim; c += 32) s = fmaf(p[c], p[c], s);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, off);
    if (lane == 0) out[row] = s;
}

// ------------------------------------------------------------------------------------------
// Coarse quantizer: score[q][c] = 2 <x_q, c> - ||c||^2 for EVERY centroid, written as candidate
// keys at keys[q * cap + c] (the layout select_topk_kernel reads after a dense phase).  Replaces
// the IndexFlatL2 quantizer search of IndexIVF::search / add (faiss coarse assignment).
// Classic 64 x 64 x 32 register-tiled fp32 GEMM (exact FMA chain in k order per output; nlist is
// small, this is a few percent of a search): 256 threads, 4 x 4 outputs per thread.  The next K
// slab is fetched into registers while the current one is multiplied: a small batch launches fewer
// CTAs than there are SMs, so nothing else hides the global-load latency of each step.
constexpr int CO_T = 64, CO_K = 32;
__global__ void __launch_bounds__(256)
coarse_scores_kernel(const float* __restrict__ xq, int nq, const float* __restrict__ cent, int nlist,
                     const float* __restrict__ cnorm, int dim, uint64_t* __restrict__ keys, uint32_t cap) {
    __shared__ float s_a[CO_K][CO_T + 4];   // [k][query]
    __shared__ float s_b[CO_K][CO_T + 4];   // [k][centroid]
    const int q0 = blockIdx.y * CO_T, c0 = blockIdx.x * CO_T;
    const int tid = threadIdx.x;
    const int tq = (tid >> 4) * 4, tc_ = (tid & 15) * 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    // loader mapping: thread loads two float4 along k for row (tid / 4), at k offsets (tid % 4) * 4 and + 16
    const int lr = tid >> 2, lk = (tid & 3) * 4;
    const bool qa = q0 + lr < nq, cb = c0 + lr < nlist;
    const float* pa = xq + (size_t)(qa ? q0 + lr : 0) * dim + lk;
    const float* pb = cent + (size_t)(cb ? c0 + lr : 0) * dim + lk;
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 va[2], vb[2];
    auto fetch = [&](int k0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool in = k0 + lk + 16 * h < dim;   // dim % 4 == 0: a float4 is inside or outside as a whole
            va[h] = (qa && in) ? __ldg(reinterpret_cast<const float4*>(pa + k0 + 16 * h)) : zero;
            vb[h] = (cb && in) ? __ldg(reinterpret_cast<const float4*>(pb + k0 + 16 * h)) : zero;
        }
    };
    fetch(0);
    for (int k0 = 0; k0 < dim; k0 += CO_K) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int kk = lk + 16 * h;
            s_a[kk + 0][lr] = va[h].x; s_a[kk + 1][lr] = va[h].y; s_a[kk + 2][lr] = va[h].z; s_a[kk + 3][lr] = va[h].w;
            s_b[kk + 0][lr] = vb[h].x; s_b[kk + 1][lr] = vb[h].y; s_b[kk + 2][lr] = vb[h].z; s_b[kk + 3][lr] = vb[h].w;
        }
        __syncthreads();
        if (k0 + CO_K < dim) fetch(k0 + CO_K);   // in flight during the multiply below
#pragma unroll
        for (int k = 0; k < CO_K; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&s_a[k][tq]);
            const float4 b = *reinterpret_cast<const float4*>(&s_b[k][tc_]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = q0 + tq + i;
        if (q >= nq) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = c0 + tc_ + j;
            if (c < nlist) keys[(size_t)q * cap + c] = make_raw_key(2.0f * acc[i][j] - cnorm[c], (uint32_t)c);
        }
    }
}

// Nearest centroid of every row from the dense key matrix (add / k-means assignment): the largest
// key = highest score, lowest centroid number on exact ties.  One warp per row.
__global__ void row_argmax_kernel(const uint64_t* __restrict__ keys, int n, int nlist, uint32_t cap,
                                  int32_t* __restrict__ assign) {
    const int row = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= n) return;
    uint64_t best = 0ull;
    for (int c = lane; c < nlist; c += 32) best = max(best, raw_to_key(keys[(size_t)row * cap + c]));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) best = max(best, __shfl_xor_sync(0xFFFFFFFFu, best, off));
    if (lane == 0) assign[row] = (int32_t)key_row(best);
}

// Sorted-by-list storage: dst[pos] = src[perm[pos]] where the source is the virtual concatenation
// [src_a (n_a rows) ; src_b]; ids come from src_ids (already concatenated), squared norms are
// computed on the way.  One warp per destination row.
__global__ void gather_sorted2_kernel(const float* __restrict__ src_a, const float* __restrict__ src_b,
                                      int64_t n_a, const int64_t* __restrict__ src_ids,
                                      const int64_t* __restrict__ perm, int64_t n, int dim,
                                      float* __restrict__ dst, int64_t* __restrict__ dst_ids,
                                      float* __restrict__ dst_norm) {
    const int64_t pos = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (pos >= n) return;
    const int64_t s = perm[pos];
    const float* row = s < n_a ? src_a + (size_t)s * dim : src_b + (size_t)(s - n_a) * dim;
    const float4* a = reinterpret_cast<const float4*>(row);
    float4* b = reinterpret_cast<float4*>(dst + (size_t)pos * dim);
    float acc = 0.f;
    for (int v = lane; v < dim / 4; v += 32) {
        const float4 t = a[v];
        b[v] = t;
        acc = fmaf(t.x, t.x, acc); acc = fmaf(t.y, t.y, acc); acc = fmaf(t.z, t.z, acc); acc = fmaf(t.w, t.w, acc);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
    if (lane == 0) { dst_ids[pos] = src_ids[s]; dst_norm[pos] = acc; }
}

// k-means update (Lloyd step of faiss's Clustering::train): centroid c = mean of the training rows
// perm[off[c] .. off[c+1]) summed in that (deterministic) order.  Empty clusters keep their old
// centroid (the host re-seeds them, ivf_api.cuh).  grid = (nlist, ceil(dim / 256)).
__global__ void centroid_mean_kernel(const float* __restrict__ x, const int64_t* __restrict__ perm,
                                     const int64_t* __restrict__ off, int dim, float* __restrict__ cent) {
    const int c = blockIdx.x;
    const int col = blockIdx.y * blockDim.x + threadIdx.x;
    if (col >= dim) return;
    const int64_t b = off[c], e = off[c + 1];
    if (e == b) return;
    float s = 0.f;
    for (int64_t i = b; i < e; ++i) s += x[(size_t)perm[i] * dim + col];
    cent[(size_t)c * dim + col] = s / (float)(e - b);
}

// ------------------------------------------------------------------------------------------
// List scan.  A tile = up to SC_ROWS consecutive positions of ONE inverted list x up to QB of the
// queries probing that list (host plan, ivf_api.cuh).
struct Tile {
    uint32_t row0;    // first position (sorted storage)
    uint32_t nrows;   // <= SC_ROWS
    uint32_t qoff;    // first entry of the tile's queries in tile_q
    uint32_t nq;      // <= QB
    uint32_t rel0;    // row0 - first position of the list
};
struct TileQuery {
    int32_t q;        // query number
    uint32_t skip;    // rows [0, skip) of this list were already scored for q in the dense phase
};
struct ScanIvfArgs {
    const Tile* tiles;
    int n_tiles;
    const TileQuery* tile_q;
    const float* q32;          // [nq_total, dim]
    const float* xnorm;        // [N]
    int dim;
    int dense;                 // 1: phase 0 (no threshold, every score appended)
    ScanArgs lists;            // thr / cnt / cand / cap / hist / hbase / k_thr (common.cuh)
};

constexpr int SC_ROWS = 512;                 // rows per tile: up to 4 per consumer thread
constexpr int SC_CONSUMERS = 128;            // warps 0..3 compute, warp 4 is the TMA producer; 2 CTAs per SM
constexpr int SC_THREADS = SC_CONSUMERS + 32;
constexpr int SC_BOX_ROWS = 128;             // TMA box: 64 bytes x 128 rows, SWIZZLE_64B (= one row slot per thread)
constexpr int SC_SLOTS = SC_ROWS / SC_BOX_ROWS;
constexpr int SC_CHUNK = 16;                 // floats of K per pipeline stage (64 bytes of every row)
constexpr int SC_BOX_BYTES = SC_BOX_ROWS * SC_CHUNK * 4;   // 8 KB
constexpr int SC_STAGE_BYTES = SC_SLOTS * SC_BOX_BYTES;    // 32 KB
constexpr int SC_MAX_STAGES = 8;
// Two CTAs per SM: (228 KB - 2 x 1 KB system reserve) / 2 per CTA, everything in dynamic shared
// memory (no static arrays): [512-byte alignment slack | ring | queries | tile state + barriers].
constexpr size_t SC_SMEM_LIMIT = (228 * 1024 - 2 * 1024) / 2;
constexpr size_t SC_AUX_BYTES = 384;   // s_qid[16], s_skip[16], s_thr[16], full[8], empty[8]
inline int scan_stages(int qb, int dim) {
    const long long room = (long long)SC_SMEM_LIMIT - 512 - (long long)SC_AUX_BYTES - (long long)qb * dim * 4;
    return (int)std::max<long long>(0, std::min<long long>(SC_MAX_STAGES, room / SC_STAGE_BYTES));
}
inline size_t scan_smem_bytes(int qb, int dim) {
    return (size_t)scan_stages(qb, dim) * SC_STAGE_BYTES + (size_t)qb * dim * 4 + 512 + SC_AUX_BYTES;
}

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
// Packed fp32 pairs (Blackwell FFMA2: two IEEE fp32 FMAs per lane and instruction).
__device__ __forceinline__ void lds_2x64(uint32_t addr, uint64_t& a, uint64_t& b) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr) : "memory");
}
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ void fma2(uint64_t& d, uint64_t a, uint64_t b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
}

// One tile for one consumer thread: R row slots (rows tid + 128 h) x QB queries.  The tile's
// queries sit TRANSPOSED in shared memory (s_q[k][j]), so one warp-wide broadcast load yields the
// k-th component of four queries as two packed pairs and the products run on FFMA2 (the scalar x
// operand is broadcast by the instruction: `FFMA2 R, Rx.F32, Rq.F32x2, R`).  Shared-memory
// wavefronts per FFMA2 = 2 / QB + 1 / R (a 128-bit broadcast load costs 2 wavefronts, a per-lane
// one 4): 0.375 at QB = 16, R = 4, below the 0.5 at which the LSU instead of the FMA pipe would
// bind (ncu: R = 2 ran at 67 % LSU / 57 % FMA).  Every (row, query) score is one fp32 FMA chain in
// k order: bit-identical for any tiling or grouping.
template <int QB, int R>
__device__ __forceinline__ void consume_tile(const ScanIvfArgs& a, const Tile& t, uint32_t base, uint32_t sq_addr,
                                             uint64_t* s_full, uint64_t* s_empty, const int* s_qid,
                                             const uint32_t* s_skip, const float* s_thr, int stages, int nkc,
                                             int tid, int lane, int& s, uint32_t& ph) {
    const uint32_t r_off = (uint32_t)tid * 64u;          // 64-byte row slots, 8 KB per box
    const uint32_t sw = (uint32_t)((tid >> 1) & 3);      // SWIZZLE_64B: granule ^= (row / 2) % 4
    uint64_t acc[R][QB / 2];   // [row slot][query pair]
#pragma unroll
    for (int h = 0; h < R; ++h)
#pragma unroll
        for (int p = 0; p < QB / 2; ++p) acc[h][p] = 0ull;
    for (int kc = 0; kc < nkc; ++kc) {
        mbar_wait(smem_u32(&s_full[s]), ph);
        const uint32_t st = base + s * SC_STAGE_BYTES + r_off;
#pragma unroll 2
        for (int g = 0; g < SC_CHUNK / 4; ++g) {   // unrolled by two only: four row-slot variants share the i-cache
            const uint32_t go = ((uint32_t)g ^ sw) * 16u;
            float xv[R][4];
#pragma unroll
            for (int h = 0; h < R; ++h) {
                const float4 x = lds_f4(st + h * SC_BOX_BYTES + go);
                xv[h][0] = x.x; xv[h][1] = x.y; xv[h][2] = x.z; xv[h][3] = x.w;
            }
            const uint32_t qa = sq_addr + (uint32_t)((kc * SC_CHUNK + g * 4) * QB) * 4u;
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
                for (int jj = 0; jj < QB / 4; ++jj) {
                    uint64_t q01, q23;
                    lds_2x64(qa + (uint32_t)((kk * QB + 4 * jj) * 4), q01, q23);
#pragma unroll
                    for (int h = 0; h < R; ++h) {
                        const uint64_t px = pack2(xv[h][kk], xv[h][kk]);
                        fma2(acc[h][2 * jj], px, q01);
                        fma2(acc[h][2 * jj + 1], px, q23);
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(smem_u32(&s_empty[s]));   // this warp is done with stage s
        if (++s == stages) { s = 0; ph ^= 1u; }
    }
    // epilogue: score = 2 <q, x> - ||x||^2, filter, append
    bool need = false;
    int need_q = 0;
#pragma unroll
    for (int h = 0; h < R; ++h) {
        const uint32_t r = (uint32_t)tid + (uint32_t)(SC_BOX_ROWS * h);
        if (r < t.nrows) {
            const uint32_t pos = t.row0 + r;
            const uint32_t rel = t.rel0 + r;
            const float xn = __ldg(a.xnorm + pos);
#pragma unroll
            for (int p = 0; p < QB / 2; ++p) {
                float d2[2];
                unpack2(acc[h][p], d2[0], d2[1]);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int j = 2 * p + e;
                    const float sc = fmaf(2.0f, d2[e], -xn);
                    if (sc > s_thr[j] && rel >= s_skip[j]) {   // unused slots carry +inf, dense phases -inf
                        const int q = s_qid[j];
                        if (a.dense) {
                            const uint32_t slot = atomicAdd(a.lists.cnt + q, 1u);
                            if (slot < a.lists.cap) a.lists.cand[(size_t)q * a.lists.cap + slot] = make_raw_key(sc, pos);
                        } else if (append_candidate(a.lists, q, sc, pos)) {
                            need = true; need_q = q;
                        }
                    }
                }
            }
        }
    }
    __syncwarp();
    if (a.lists.hist && !a.dense) refresh_thresholds_warp(a.lists, need, need_q, lane);
}

// Persistent CTAs (two per SM), static round-robin tile schedule, warp-specialised: warp 4 streams
// the 16-float K chunks of ALL of the CTA's tiles through a ring of `stages` stages (TMA, full /
// empty mbarriers), running ahead across tile boundaries, so HBM stays busy during epilogues and
// query staging; warps 0..3 consume (no CTA-wide barrier in the K loop: a warp releases a stage
// with one mbarrier arrival).  x rows come in through a SWIZZLE_64B tensor map, one 128-row box
// per row slot (lane i reads row i: the swizzle spreads the granules of 8 consecutive rows over all
// banks, so the 128-bit loads are conflict-free without padding); only the boxes that hold rows of
// the tile are loaded, and the inner loop is instantiated per number of live row slots, so a short
// tile costs neither bandwidth nor FMAs for its missing rows.
template <int QB>
__global__ void __launch_bounds__(SC_THREADS, 2)
ivf_scan_kernel(const __grid_constant__ CUtensorMap tmap_x, ScanIvfArgs a, int stages) {
    static_assert(QB % 4 == 0, "queries are loaded four at a time");
    static_assert(QB <= 16, "tile state is sized for 16 queries");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 511u) & ~511u;   // SWIZZLE_64B pattern repeats every 512 bytes
    const uint32_t sq_addr = base + (uint32_t)stages * SC_STAGE_BYTES;
    uint8_t* gen = smem_raw + (sq_addr - smem_u32(smem_raw));
    float* s_q = reinterpret_cast<float*>(gen);                  // [dim][QB]
    uint8_t* aux = gen + (size_t)QB * a.dim * 4;
    uint64_t* s_full = reinterpret_cast<uint64_t*>(aux);         // [8]
    uint64_t* s_empty = s_full + SC_MAX_STAGES;                  // [8]
    int* s_qid = reinterpret_cast<int*>(aux + 128);              // [16]
    uint32_t* s_skip = reinterpret_cast<uint32_t*>(aux + 192);   // [16]
    float* s_thr = reinterpret_cast<float*>(aux + 256);          // [16]
    const int tid = threadIdx.x, lane = tid & 31;
    const int nkc = a.dim / SC_CHUNK;
    const int my_tiles = a.n_tiles > (int)blockIdx.x ? (a.n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    if (tid == 0) {
        tma_prefetch_desc(&tmap_x);
        for (int s = 0; s < stages; ++s) {
            mbar_init(smem_u32(&s_full[s]), 1);
            mbar_init(smem_u32(&s_empty[s]), SC_CONSUMERS / 32);
        }
        fence_barrier_init();
    }
    __syncthreads();
    if (tid >= SC_CONSUMERS) {   // ===================== TMA producer (one lane)
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int ti = 0; ti < my_tiles; ++ti) {
                const Tile t = a.tiles[(size_t)blockIdx.x + (size_t)ti * gridDim.x];
                const int slots = (int)(t.nrows + SC_BOX_ROWS - 1) / SC_BOX_ROWS;
                for (int kc = 0; kc < nkc; ++kc) {
                    mbar_wait(smem_u32(&s_empty[s]), ph ^ 1u);
                    const uint32_t bar = smem_u32(&s_full[s]);
                    mbar_arrive_expect_tx(bar, (uint32_t)slots * SC_BOX_BYTES);
                    for (int b = 0; b < slots; ++b)   // rows past the list's end are loaded and ignored
                        tma_load_2d(base + s * SC_STAGE_BYTES + b * SC_BOX_BYTES, &tmap_x, bar,
                                    kc * SC_CHUNK, (int)t.row0 + b * SC_BOX_ROWS);
                    if (++s == stages) { s = 0; ph ^= 1u; }
                }
            }
        }
        return;
    }
    int s = 0;
    uint32_t ph = 0;
    for (int ti = 0; ti < my_tiles; ++ti) {
        const Tile t = a.tiles[(size_t)blockIdx.x + (size_t)ti * gridDim.x];
        asm volatile("bar.sync 1, 128;" ::: "memory");   // previous tile's epilogue and K loop are done
        if (tid < QB) {
            TileQuery tq{-1, 0u};
            if (tid < (int)t.nq) tq = a.tile_q[t.qoff + tid];
            s_qid[tid] = tq.q;
            s_skip[tid] = tq.skip;
            s_thr[tid] = tq.q < 0 ? __builtin_huge_valf() : (a.dense ? kNegInf : __ldcg(a.lists.thr + tq.q));
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // stage the tile's queries transposed: s_q[k][j]; loads are issued eight at a time so the
        // L2 round trips overlap (lanes walk the queries: <= 4-way store conflicts; slots of unused
        // queries keep stale values: their scores meet a +inf threshold)
        {
            const int vec = a.dim / 4, total = QB * vec;
            constexpr int BATCH = 8;
            for (int i0 = tid; i0 < total; i0 += SC_CONSUMERS * BATCH) {
                float4 val[BATCH];
#pragma unroll
                for (int u = 0; u < BATCH; ++u) {
                    const int i = i0 + u * SC_CONSUMERS;
                    const int j = i % QB, v = i / QB;
                    val[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (i < total && j < (int)t.nq)
                        val[u] = __ldg(reinterpret_cast<const float4*>(a.q32 + (size_t)s_qid[j] * a.dim) + v);
                }
#pragma unroll
                for (int u = 0; u < BATCH; ++u) {
                    const int i = i0 + u * SC_CONSUMERS;
                    if (i < total) {
                        const int j = i % QB, v = i / QB;
                        float* dst = s_q + (size_t)(4 * v) * QB + j;
                        dst[0] = val[u].x; dst[QB] = val[u].y; dst[2 * QB] = val[u].z; dst[3 * QB] = val[u].w;
                    }
                }
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        switch ((t.nrows + SC_BOX_ROWS - 1) / SC_BOX_ROWS) {
            case 1: consume_tile<QB, 1>(a, t, base, sq_addr, s_full, s_empty, s_qid, s_skip, s_thr, stages, nkc, tid, lane, s, ph); break;
            case 2: consume_tile<QB, 2>(a, t, base, sq_addr, s_full, s_empty, s_qid, s_skip, s_thr, stages, nkc, tid, lane, s, ph); break;
            case 3: consume_tile<QB, 3>(a, t, base, sq_addr, s_full, s_empty, s_qid, s_skip, s_thr, stages, nkc, tid, lane, s, ph); break;
            default: consume_tile<QB, 4>(a, t, base, sq_addr, s_full, s_empty, s_qid, s_skip, s_thr, stages, nkc, tid, lane, s, ph); break;
        }
    }
}

// Largest squared row norm (for the error bound below).  Norms are non-negative, so their bit
// patterns order like unsigned integers.
__global__ void max_norm_kernel(const float* __restrict__ xnorm, int64_t n, float* __restrict__ out) {
    float m = 0.f;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = fmaxf(m, xnorm[i]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xFFFFFFFFu, m, off));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned int*>(out), __float_as_uint(m));
}

// Exactness window of the scan's scores.  The scan ranks by s = 2 <q, x> - ||x||^2 evaluated in
// fp32 (one FMA chain over dim terms, then one rounding; ||x||^2 itself summed in fp32), which
// suffers cancellation when the distance is small next to the norms.  A rigorous bound on
// |s_fp32 - s_exact| for ANY stored row (u = 2^-24, standard gamma_n bounds, Cauchy-Schwarz):
//     eps = 1.01 u (dim + 1) (2 ||q|| X + X^2),   X = max_r ||x_r||.
// The selects keep every candidate within 2 eps of the k-th best (select.cuh, "exactness window"),
// ivf_rescore_kernel then evaluates sum (q - x)^2 directly (what faiss's IVFFlatScannerL2 computes,
// no cancellation) and the final select ranks by that.
// With `mode` 1 (some lists go through the tensor-core kernel on tf32, tc_ivf.cuh) the bound also
// covers the truncation of both operands to tf32 (2^-9 + 2^-20 relative to ||q|| ||x||) and the
// tensor core's fp32 accumulation, on the doubled inner product.  `mode` 2 (fp16 filter copy and
// fp16-rounded queries): both operands rounded to nearest, 2^-10 + 2^-22 relative, plus the
// absolute error of components that fall into fp16's subnormal range (at most 2^-25 each:
// sqrt(dim) 2^-25 (||q|| + ||x||) on the inner product); overflow is excluded by the caller
// (max ||x||^2 and every ||q||^2 below 4e9, i.e. every component below 65504).
__global__ void ivf_eps_kernel(const float* __restrict__ qnorm, const float* __restrict__ x2max, int nq,
                               int dim, int mode, float* __restrict__ eps, float* __restrict__ ones) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float X2 = *x2max;
    float e = 1.01f * 5.9604645e-8f * (float)(dim + 1) * (2.0f * sqrtf(qnorm[q] * X2) + X2);
    if (mode == 1) {
        const float rel = (float)dim * 1.1920929e-7f + 1.0f / 512 + 1.0f / 1048576;
        e = fmaxf(e, 1.01f * (2.0f * rel * sqrtf(qnorm[q] * X2) + 5.9604645e-8f * (float)(dim + 1) * X2));
    } else if (mode == 2) {
        const float rel = (float)dim * 1.1920929e-7f + 1.0f / 1024 + 1.0f / 4194304;
        const float under = sqrtf((float)dim) * 2.9802322e-8f * (sqrtf(qnorm[q]) + sqrtf(X2)) + (float)dim * 8.9e-16f;
        e = fmaxf(e, 1.01f * (2.0f * rel * sqrtf(qnorm[q] * X2) + 2.0f * under +
                              5.9604645e-8f * (float)(dim + 1) * X2));
    }
    eps[q] = e;
    ones[q] = 1.0f;
}

// Exact rescoring of the surviving candidates: key <- (-(sum (q - x)^2), pos).  One warp per
// (query, list slot).
__global__ void ivf_rescore_kernel(uint64_t* __restrict__ cand, const uint32_t* __restrict__ cnt,
                                   uint32_t cap, int nq, int slots, const float* __restrict__ x,
                                   const float* __restrict__ q32, int dim) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t w = warp; w < (int64_t)nq * slots; w += nwarps) {
        const int q = (int)(w / slots), slot = (int)(w - (int64_t)q * slots);
        if (slot >= (int)min(cnt[q], cap)) continue;  // warp-uniform
        const uint32_t pos = key_row(cand[(size_t)q * cap + slot]);
        const float4* xv = reinterpret_cast<const float4*>(x + (size_t)pos * dim);
        const float4* qv = reinterpret_cast<const float4*>(q32 + (size_t)q * dim);
        float acc = 0.f;
        for (int v = lane; v < dim / 4; v += 32) {
            const float4 a = xv[v], b = qv[v];
            const float d0 = b.x - a.x, d1 = b.y - a.y, d2 = b.z - a.z, d3 = b.w - a.w;
            acc = fmaf(d0, d0, acc); acc = fmaf(d1, d1, acc); acc = fmaf(d2, d2, acc); acc = fmaf(d3, d3, acc);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
        if (lane == 0) cand[(size_t)q * cap + slot] = make_raw_key(-acc, pos);
    }
}

// (score, pos) -> (squared L2 distance, id).  faiss pads missing results with id -1 and the
// CMax-heap neutral FLT_MAX.  One thread per output slot.
__global__ void ivf_finalize_kernel(const float* __restrict__ sc, const int64_t* __restrict__ pos,
                                    const int64_t* __restrict__ ids,
                                    int nq, int k, float* __restrict__ out_d, int64_t* __restrict__ out_i) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)nq * k) return;
    const int64_t p = pos[i];
    if (p < 0) { out_d[i] = FLT_MAX; out_i[i] = -1; return; }
    out_d[i] = -sc[i];   // the final select ranked by -(sum (q - x)^2)
    out_i[i] = ids[p];
}

}  // namespace ivf
}  // namespace fip
