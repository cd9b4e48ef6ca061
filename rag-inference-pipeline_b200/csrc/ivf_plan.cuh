// IVF-Flat search plan built ON THE DEVICE (normal mode): inverts the probe table [nq, nprobe] into
// the list-major work lists the scan kernels consume -- phase-0 prefix tiles, phase-1 FFMA2 tiles
// (ivf.cuh) and tensor-core items with their prepared-query map (tc_ivf.cuh).  Same plan as the host
// planner (build_plan in ivf_api.cuh, kept for safe mode and as the CPU-testable statement of the
// rules), except that the order of the queries INSIDE a list's bucket is whatever order the atomics
// produced: grouping differs from run to run, results do not (every pair is still scored exactly
// once and the final ranking is on exactly rescored distances).
//
// The host reads back ONE small struct (PlanCounts) between the counting and the emitting kernels,
// to size the work lists and the launches; the probe table itself never leaves the GPU.
//
//   plan_query_kernel   one thread per query: phase-0 prefix lengths (take), bucket counts per list
//   plan_lists_kernel   one block: per-list shapes and the five exclusive scans over the lists
//   plan_scatter_kernel one thread per (query, probe): bucket fill
//   plan_tiles0_kernel  one thread per query: phase-0 tiles
//   plan_emit_kernel    one block per list: FFMA2 tiles or tensor items + query map
#pragma once

#include "ivf.cuh"
#include "tc_ivf.cuh"

namespace fip {
namespace ivf {

struct PlanCounts {
    uint32_t n_tiles0;    // phase-0 single-query tiles
    uint32_t n_t16;       // phase-1 tiles of more than 4 queries
    uint32_t n_t4;        // phase-1 tiles of up to 4 queries
    uint32_t n_items;     // tensor-core items
    uint32_t qrows;       // prepared-query rows (groups padded to 32)
    uint32_t n_pairs1;    // (query, list) pairs of phase 1 = bucket entries
    uint32_t q_over;      // 1: some query has ||q||^2 >= qlimit (no fp16 filter for this batch)
    uint32_t pad;
    unsigned long long rows, pairs, tc_pairs, distinct_rows;
    unsigned long long tiles_total;   // 64-bit sum of all tile / item counts (overflow guard)
};
static_assert(sizeof(PlanCounts) == 72, "read back as one block");

constexpr uint32_t PLAN_NO_PAIR = 0xFFFFFFFFu;
constexpr int PLAN_BASES = 5;   // bucket start, prepared-row base, item base, t16 base, t4 base

struct ListShape { uint32_t v[PLAN_BASES]; uint32_t groups; };

// Work of one list probed by m queries (phase 1).  Mirrors build_plan's grouping rules.
__device__ __forceinline__ ListShape list_shape(uint32_t m, uint32_t len, int tc_min, int qb_big) {
    ListShape s{};
    s.v[0] = m;
    if (m == 0 || len == 0) return s;
    if (tc_min > 0 && m >= (uint32_t)tc_min) {
        const uint32_t groups = (m + 127u) / 128u;
        const uint32_t last = m - (groups - 1u) * 128u;
        s.groups = groups;
        s.v[1] = (groups - 1u) * 128u + (last + 31u) / 32u * 32u;
        s.v[2] = groups * ((len + (uint32_t)tc::BN - 1u) / (uint32_t)tc::BN);
        return s;
    }
    const uint32_t tpg = (len + SC_ROWS - 1u) / SC_ROWS;
    uint32_t n_big = 0, n_small = 0;
    if (qb_big >= 16) {
        const uint32_t rem = m % 16u;
        n_big = m / 16u + (rem > 4u ? 1u : 0u);
        n_small = rem >= 1u && rem <= 4u ? 1u : 0u;
    } else {
        n_small = (m + 3u) / 4u;
    }
    s.v[3] = n_big * tpg;
    s.v[4] = n_small * tpg;
    s.groups = n_big + n_small;
    return s;
}

// probes [nq, nprobe] nearest first (-1 = none).  take[q, j] = rows of list probes[q, j] that q
// scores in phase 0 (PLAN_NO_PAIR: no such list, or an empty one).  One WARP per query: the lanes
// take 32 consecutive probes per round and the running prefix of the list lengths (saturated at
// target0, so it fits 32 bits) comes from a warp scan -- a single thread walking the probe
// sequence paid one dependent L2 round trip per probe.
__global__ void plan_query_kernel(const int64_t* __restrict__ probes, int nq, int nprobe, int nlist,
                                  const int64_t* __restrict__ off, uint32_t target0,
                                  uint32_t* __restrict__ take, uint32_t* __restrict__ lcount,
                                  uint32_t* __restrict__ probed, uint32_t* __restrict__ tile0base,
                                  const float* __restrict__ qnorm, float qlimit,
                                  PlanCounts* __restrict__ counts) {
    const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;   // warp-uniform
    if (lane == 0 && !(qnorm[q] < qlimit)) counts->q_over = 1u;
    uint32_t cum = 0, ntile0 = 0;   // cum: rows taken by the probes of earlier rounds (<= target0), warp-uniform
    for (int j0 = 0; j0 < nprobe; j0 += 32) {
        const int j = j0 + lane;
        int64_t l = -1;
        uint32_t len = 0;
        if (j < nprobe) {
            l = probes[(size_t)q * nprobe + j];
            if (l >= 0 && l < nlist) len = (uint32_t)(off[l + 1] - off[l]);
        }
        const uint32_t capped = min(len, target0);
        uint32_t incl = capped;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += y;
        }
        const uint32_t before = min(target0, cum + (incl - capped));   // rows taken before this probe
        uint32_t t = PLAN_NO_PAIR;
        if (len > 0) {
            probed[l] = 1u;
            t = min(len, target0 - before);
            ntile0 += (t + SC_ROWS - 1u) / SC_ROWS;
            if (t < len) atomicAdd(&lcount[l], 1u);
        }
        if (j < nprobe) take[(size_t)q * nprobe + j] = t;
        cum = min(target0, cum + __shfl_sync(0xFFFFFFFFu, incl, 31));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ntile0 += __shfl_xor_sync(0xFFFFFFFFu, ntile0, o);
    if (lane == 0) {
        tile0base[q] = ntile0 ? atomicAdd(&counts->n_tiles0, ntile0) : 0u;
        if (cum) {
            atomicAdd(&counts->rows, (unsigned long long)cum);
            atomicAdd(&counts->pairs, (unsigned long long)cum);
        }
    }
}

constexpr int PLAN_LIST_THREADS = 1024;

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_warp[lane];   // PLAN_LIST_THREADS / 32 == 32 warps
        uint32_t wx = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, wx, o);
            if (lane >= o) wx += y;
        }
        s_warp[lane] = wx - w;
        if (lane == 31) s_warp[32] = wx;
    }
    __syncthreads();
    const uint32_t r = x - v + s_warp[warp];
    *total = s_warp[32];
    __syncthreads();
    return r;
}

// bases: [PLAN_BASES][nlist + 1] exclusive scans over the lists (entry nlist = total).
__global__ void __launch_bounds__(PLAN_LIST_THREADS)
plan_lists_kernel(const uint32_t* __restrict__ lcount, const uint32_t* __restrict__ probed,
                  const int64_t* __restrict__ off, int nlist, int tc_min, int qb_big,
                  uint32_t* __restrict__ bases, PlanCounts* __restrict__ counts) {
    __shared__ uint32_t s_warp[33];
    const int tid = threadIdx.x;
    const int per = (nlist + PLAN_LIST_THREADS - 1) / PLAN_LIST_THREADS;
    const int l0 = min(tid * per, nlist), l1 = min(l0 + per, nlist);
    unsigned long long rows = 0, pairs = 0, tc_pairs = 0, distinct = 0, tiles_total = 0;
    for (int l = l0; l < l1; ++l) {
        const uint32_t len = (uint32_t)(off[l + 1] - off[l]);
        if (probed[l]) distinct += len;
        const uint32_t m = lcount[l];
        const ListShape s = list_shape(m, len, tc_min, qb_big);
        rows += (unsigned long long)len * s.groups;
        pairs += (unsigned long long)len * (len ? m : 0u);
        if (s.v[2]) tc_pairs += (unsigned long long)len * m;
        tiles_total += (unsigned long long)s.v[2] + s.v[3] + s.v[4];
    }
    // shapes of this thread's lists stay in registers across the five scans when there are few of
    // them (nlist <= 4096), else they are recomputed
    constexpr int kCache = 4;
    ListShape sh[kCache];
    const bool cached = per <= kCache;
    if (cached) {
#pragma unroll
        for (int i = 0; i < kCache; ++i) {
            const int l = l0 + i;
            sh[i] = l < l1 ? list_shape(lcount[l], (uint32_t)(off[l + 1] - off[l]), tc_min, qb_big) : ListShape{};
        }
    }
#pragma unroll
    for (int w = 0; w < PLAN_BASES; ++w) {
        uint32_t sum = 0;
        if (cached) {
#pragma unroll
            for (int i = 0; i < kCache; ++i) sum += sh[i].v[w];
        } else {
            for (int l = l0; l < l1; ++l)
                sum += list_shape(lcount[l], (uint32_t)(off[l + 1] - off[l]), tc_min, qb_big).v[w];
        }
        uint32_t total;
        uint32_t b = block_excl_scan(sum, s_warp, &total);
        uint32_t* dst = bases + (size_t)w * (nlist + 1);
        if (cached) {
#pragma unroll
            for (int i = 0; i < kCache; ++i) {
                if (l0 + i < l1) dst[l0 + i] = b;
                b += sh[i].v[w];
            }
        } else {
            for (int l = l0; l < l1; ++l) {
                dst[l] = b;
                b += list_shape(lcount[l], (uint32_t)(off[l + 1] - off[l]), tc_min, qb_big).v[w];
            }
        }
        if (tid == 0) {
            dst[nlist] = total;
            if (w == 0) counts->n_pairs1 = total;
            if (w == 1) counts->qrows = total;
            if (w == 2) counts->n_items = total;
            if (w == 3) counts->n_t16 = total;
            if (w == 4) counts->n_t4 = total;
        }
    }
    if (rows) atomicAdd(&counts->rows, rows);
    if (pairs) atomicAdd(&counts->pairs, pairs);
    if (tc_pairs) atomicAdd(&counts->tc_pairs, tc_pairs);
    if (distinct) atomicAdd(&counts->distinct_rows, distinct);
    if (tiles_total) atomicAdd(&counts->tiles_total, tiles_total);
}

// Bucket fill: lq[bucket start of the list + arrival order] = {query, rows already scored}.
__global__ void plan_scatter_kernel(const int64_t* __restrict__ probes, const uint32_t* __restrict__ take,
                                    int64_t n_pairs, int nprobe, const int64_t* __restrict__ off,
                                    const uint32_t* __restrict__ lstart, uint32_t* __restrict__ cur,
                                    TileQuery* __restrict__ lq) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    const uint32_t t = take[i];
    if (t == PLAN_NO_PAIR) return;
    const int64_t l = probes[i];
    const uint32_t len = (uint32_t)(off[l + 1] - off[l]);
    if (t >= len) return;   // the whole list was scored in phase 0
    const uint32_t slot = lstart[l] + atomicAdd(&cur[l], 1u);
    lq[slot] = TileQuery{(int32_t)(i / nprobe), t};
}

// Phase-0 tiles of one query (its first `take` rows of each probed list) and its own tile_q entry
// at tq0_base + q.
__global__ void plan_tiles0_kernel(const int64_t* __restrict__ probes, const uint32_t* __restrict__ take,
                                   int nq, int nprobe, const int64_t* __restrict__ off,
                                   const uint32_t* __restrict__ tile0base, uint32_t tq0_base,
                                   Tile* __restrict__ tiles0, TileQuery* __restrict__ tile_q) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    tile_q[tq0_base + q] = TileQuery{q, 0u};
    uint32_t w = tile0base[q];
    for (int j = 0; j < nprobe; ++j) {
        const uint32_t t = take[(size_t)q * nprobe + j];
        if (t == PLAN_NO_PAIR) continue;
        if (t == 0) break;   // prefix lengths are positive up to the target, then zero
        const uint32_t row0 = (uint32_t)off[probes[(size_t)q * nprobe + j]];
        for (uint32_t r = 0; r < t; r += SC_ROWS)
            tiles0[w++] = Tile{row0 + r, min((uint32_t)SC_ROWS, t - r), tq0_base + (uint32_t)q, 1u, r};
    }
}

// One block per list: tensor items + prepared-query map, or FFMA2 tiles over the bucket itself
// (tile_q == lq: a tile's queries are a contiguous slice of the list's bucket).
__global__ void __launch_bounds__(128)
plan_emit_kernel(const uint32_t* __restrict__ bases, const int64_t* __restrict__ off, int nlist,
                 int tc_min, int qb_big, const TileQuery* __restrict__ lq, Tile* __restrict__ t16,
                 Tile* __restrict__ t4, TcItem* __restrict__ items, int32_t* __restrict__ qmap,
                 uint32_t* __restrict__ qskip) {
    const int l = blockIdx.x;
    const uint32_t* b0 = bases;
    const uint32_t lstart = b0[l], m = b0[l + 1] - lstart;
    if (m == 0) return;
    const uint32_t list0 = (uint32_t)off[l], len = (uint32_t)(off[l + 1] - off[l]);
    if (len == 0) return;
    const ListShape s = list_shape(m, len, tc_min, qb_big);
    const int tid = threadIdx.x, nt = blockDim.x;
    if (s.v[2]) {
        const uint32_t rowbase = bases[(size_t)1 * (nlist + 1) + l];
        const uint32_t itembase = bases[(size_t)2 * (nlist + 1) + l];
        const uint32_t groups = s.groups, qrows = s.v[1];
        for (uint32_t i = tid; i < qrows; i += nt) {   // groups of 128 are dense, only the last one is padded
            const bool real = i < m;
            qmap[rowbase + i] = real ? lq[lstart + i].q : -1;
            qskip[rowbase + i] = real ? lq[lstart + i].skip : 0u;
        }
        const uint32_t last_rows = qrows - (groups - 1u) * 128u;
        // row-tile major: the groups of one 256-row tile are adjacent items, i.e. run on different
        // SMs at the same time: the tile comes from HBM once, then from L2
        for (uint32_t t = tid; t < s.v[2]; t += nt) {
            const uint32_t rt = t / groups, g = t - rt * groups;
            items[itembase + t] = TcItem{list0 + rt * (uint32_t)tc::BN, list0 + len, rowbase + g * 128u, list0,
                                         g + 1u < groups ? 128u : last_rows};
        }
        return;
    }
    const uint32_t tpg = (len + SC_ROWS - 1u) / SC_ROWS;
    const uint32_t n_big = s.v[3] / tpg, n_small = s.v[4] / tpg;
    const uint32_t t16base = bases[(size_t)3 * (nlist + 1) + l];
    const uint32_t t4base = bases[(size_t)4 * (nlist + 1) + l];
    for (uint32_t t = tid; t < s.v[3]; t += nt) {
        const uint32_t g = t / tpg, rt = t - g * tpg;
        const uint32_t g0 = g * 16u;
        t16[t16base + t] = Tile{list0 + rt * SC_ROWS, min((uint32_t)SC_ROWS, len - rt * SC_ROWS), lstart + g0,
                                min(16u, m - g0), rt * SC_ROWS};
    }
    const uint32_t small0 = qb_big >= 16 ? n_big * 16u : 0u;
    for (uint32_t t = tid; t < s.v[4]; t += nt) {
        const uint32_t g = t / tpg, rt = t - g * tpg;
        const uint32_t g0 = small0 + g * 4u;
        t4[t4base + t] = Tile{list0 + rt * SC_ROWS, min((uint32_t)SC_ROWS, len - rt * SC_ROWS), lstart + g0,
                              min(4u, m - g0), rt * SC_ROWS};
    }
    (void)n_small;
}

}  // namespace ivf
}  // namespace fip
