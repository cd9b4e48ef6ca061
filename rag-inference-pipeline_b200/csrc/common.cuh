// Shared definitions for the B200 flat inner-product search kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cfloat>
#include <cstdio>
#include <cstring>
#include <string>

namespace fip {

enum Dtype : int { kF32 = 0, kF16 = 1, kBF16 = 2 };

__host__ __device__ inline int dtype_size(int dt) { return dt == kF32 ? 4 : 2; }

// ---------------------------------------------------------------------------------------------
// Candidate keys.  A candidate is (score, local row) packed into 64 bits so that one unsigned
// compare implements the fixed tie-break "score descending, then id ascending":
//   key = orderable(score) << 32 | (0xFFFFFFFF - row)          larger key == better candidate
// key 0 is the "empty slot" sentinel: every real score (even -inf) has orderable(score) > 0.
// ---------------------------------------------------------------------------------------------
__host__ __device__ inline uint32_t orderable_f32(float s) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(s);
#else
    uint32_t u;
    memcpy(&u, &s, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float from_orderable_f32(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float s;
    memcpy(&s, &u, 4);
    return s;
#endif
}
__host__ __device__ inline uint64_t make_key(float score, uint32_t row) {
    return ((uint64_t)orderable_f32(score) << 32) | (uint64_t)(0xFFFFFFFFu - row);
}
__host__ __device__ inline float key_score(uint64_t key) {
    return from_orderable_f32((uint32_t)(key >> 32));
}
__host__ __device__ inline uint32_t key_row(uint64_t key) {
    return 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFu);
}

// Lists in GLOBAL memory hold RAW keys (score bits << 32 | 0xFFFFFFFF - row): the scan epilogues
// store them without any ALU work per survivor.  The select kernel converts raw <-> orderable when
// it stages a list in shared memory.  The row half is identical in both forms.
__host__ __device__ inline uint64_t make_raw_key(float score, uint32_t row) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(score);
#else
    uint32_t u;
    memcpy(&u, &score, 4);
#endif
    return ((uint64_t)u << 32) | (uint64_t)(0xFFFFFFFFu - row);
}
__host__ __device__ inline uint64_t raw_to_key(uint64_t raw) {
    uint32_t u = (uint32_t)(raw >> 32);
    uint32_t o = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ((uint64_t)o << 32) | (raw & 0xFFFFFFFFull);
}
__host__ __device__ inline uint64_t key_to_raw(uint64_t key) {
    uint32_t o = (uint32_t)(key >> 32);
    uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
    return ((uint64_t)u << 32) | (key & 0xFFFFFFFFull);
}

constexpr float kNegInf = -__builtin_huge_valf();
constexpr float kLowestF32 = -FLT_MAX;  // faiss pads unfilled result slots with -FLT_MAX / id -1

// Per-search device workspace handed to the scan kernels (all pointers are device pointers).
struct ScanArgs {
    const void* corpus;        // [n_rows, dim] storage dtype, row-major
    int64_t n_rows;            // rows stored on this device
    int dim;
    int dtype;
    int nq;                    // number of real queries
    uint32_t row_begin;        // phase row range [row_begin, row_end)
    uint32_t row_end;
    int dense;                 // 1: phase 0, every score is written at slot = row - row_begin
    float* thr;                // [nq] running threshold per query (strict > passes); raised live
    uint32_t* cnt;             // [nq] candidates currently in each list
    uint64_t* cand;            // [nq, cap] candidate keys
    uint32_t cap;
    // Live thresholds (null hist = off).  Every candidate in a list is also counted in
    // hist[q][bucket], bucket = min(255, (orderable(score) - hbase[q]) >> 16), hbase[q] = the
    // orderable bits of the threshold the select pass published.  Whenever a list grows past a
    // multiple of kRefreshEvery entries the appending warp walks the histogram from the top: if
    // k_thr candidates sit in buckets >= b, the k_thr-th best score so far is >= the lower edge of
    // bucket b, so everything below that edge (minus the exactness window 2 eps / scale of the
    // tensor-core filter) can be dropped from now on.  Counts only grow and thresholds are only
    // raised, so a stale read of either is merely conservative.  This is what lets ONE filter
    // launch cover the whole shard: survivors grow like k log(N / m0) instead of k N / m0.
    uint32_t* hist;            // [nq, 256]
    const uint32_t* hbase;     // [nq]
    const float* win_eps;      // [nq] or null (exact scores)
    const float* win_scale;    // [nq] or null
    int k_thr;
};
constexpr uint32_t kRefreshEvery = 128;   // list entries between two threshold refreshes
constexpr int kHistBins = 256;
constexpr int kHistShift = 16;            // 2^16 ulps per bucket = 2^-7 relative score resolution

__device__ __forceinline__ uint32_t hist_bucket(uint32_t score_bits, uint32_t hbase) {
    const uint32_t o = (score_bits & 0x80000000u) ? ~score_bits : (score_bits | 0x80000000u);
    const uint32_t d = (o - hbase) >> kHistShift;   // o > hbase for every candidate (score > threshold)
    return d < (uint32_t)(kHistBins - 1) ? d : (uint32_t)(kHistBins - 1);
}
// Monotone raise of a float in global memory (works across signs: positive floats order like
// signed ints, negative floats inversely like unsigned ints).
__device__ __forceinline__ void raise_threshold(float* addr, float v) {
    if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
// New threshold implied by "at least k_thr candidates in buckets >= b" (b >= 1).
__device__ __forceinline__ float threshold_from_bucket(const ScanArgs& a, int q, uint32_t b) {
    const uint32_t base = a.hbase[q];
    const uint32_t add = b << kHistShift;
    uint32_t edge = base + add;
    if (edge < base) edge = 0xFFFFFFFFu;          // saturate
    // edge - 1: a score EQUAL to the bucket edge still passes the strict compare (ties are broken
    // by id, and rows scanned concurrently may carry lower ids than the counted candidates)
    float t = from_orderable_f32(edge - 1u);
    if (a.win_eps) t -= 2.0f * a.win_eps[q] / a.win_scale[q];
    return t;
}
// Whole-warp refresh for every lane with `need` set (all 32 lanes must call this converged).
__device__ __forceinline__ void refresh_thresholds_warp(const ScanArgs& a, bool need, int q, int lane) {
    uint32_t trig = __ballot_sync(0xFFFFFFFFu, need);
    while (trig) {
        const int src = __ffs((int)trig) - 1;
        trig &= trig - 1u;
        const int qq = __shfl_sync(0xFFFFFFFFu, q, src);
        const uint4* h = reinterpret_cast<const uint4*>(a.hist + (size_t)qq * kHistBins + lane * 8);
        const uint4 h0 = __ldcg(h), h1 = __ldcg(h + 1);
        const uint32_t c[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
        uint32_t sum = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += c[j];
        uint32_t suf = sum;  // inclusive suffix sum over lanes (lane 31 holds the top buckets)
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_down_sync(0xFFFFFFFFu, suf, off);
            if (lane + off < 32) suf += t;
        }
        const uint32_t k = (uint32_t)a.k_thr;
        if (suf >= k && suf - sum < k) {
            uint32_t above = suf - sum;
            int b = lane * 8;
#pragma unroll
            for (int j = 7; j >= 0; --j) {
                if (above + c[j] >= k) { b = lane * 8 + j; break; }
                above += c[j];
            }
            if (b > 0) raise_threshold(a.thr + qq, threshold_from_bucket(a, qq, (uint32_t)b));
        }
    }
}
// Append one candidate (tail of the streaming kernel's epilogue).  Returns true when the list just
// grew past a multiple of kRefreshEvery entries: the caller then refreshes the query's threshold
// (refresh_thresholds_warp, once the warp has reconverged).
__device__ __forceinline__ bool append_candidate(const ScanArgs& a, int q, float score,
                                                 uint32_t row) {
    uint32_t slot = atomicAdd(a.cnt + q, 1u);
    if (slot >= a.cap) return false;
    a.cand[(size_t)q * a.cap + slot] = make_raw_key(score, row);
    if (!a.hist) return false;
    atomicAdd(a.hist + (size_t)q * kHistBins + hist_bucket(__float_as_uint(score), a.hbase[q]), 1u);
    return ((slot + 1u) & (kRefreshEvery - 1u)) == 0u;
}

// Per-search reset of the candidate lists and flag words (select.cuh flag_slot), one element per
// thread i.  `sticky` (may be null): the summary words an earlier, never checked (asynchronous)
// search left behind are folded into it before they are cleared, so fip_search_status still sees
// them.  Run by init_lists_kernel, or fused into prepare_queries_kernel (one launch fewer).
struct ListInit {
    float* thr;
    uint32_t* cnt;
    uint32_t* flags;
    uint32_t* sticky;
    int nq;
};
__device__ __forceinline__ void init_lists_thread(const ListInit& a, int i) {
    if (i < a.nq) { a.thr[i] = kNegInf; a.cnt[i] = 0u; }
    if (i < 2 && a.sticky && a.flags[i]) atomicOr(a.sticky, 1u << i);
    if (i < 2 + 2 * a.nq) a.flags[i] = 0u;
}

inline int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

}  // namespace fip
