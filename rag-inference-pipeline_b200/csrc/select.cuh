// K3: exact top-k selection over per-query candidate lists, and the cross-shard k-way merge.
//
// Replaces faiss's result handlers (heap_replace_top / reservoir shrink / heap_reorder, reached
// from /root/reference/src/pipeline/components/faiss_store.py:152) with one primitive: sort the
// 64-bit candidate keys of a query in shared memory (bitonic, descending) and keep the first k.
// Because a key packs (score, row) so that "larger key" == "higher score, then lower row", the
// result order is the fixed tie-break (score desc, id asc).  The same primitive is used
//   * between scan phases  (compact the list to its k best, publish the new threshold),
//   * at the end of a search (emit distances / ids), and
//   * after the all-gather (merge g sorted shard lists).
#pragma once

#include <type_traits>

#include "common.cuh"
#include "stream_scan.cuh"

namespace fip {

// Bitonic sort, descending, of n (power of two) keys in shared memory by the whole CTA.
__device__ __forceinline__ void bitonic_sort_desc(uint64_t* s, int n) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int i = threadIdx.x; i < (n >> 1); i += blockDim.x) {
                int pos = 2 * i - (i & (stride - 1));
                uint64_t a = s[pos], b = s[pos + stride];
                bool desc = (pos & size) == 0;
                if ((a < b) == desc) { s[pos] = b; s[pos + stride] = a; }
            }
            __syncthreads();
        }
    }
}

struct SelectArgs {
    uint64_t* cand;       // [nq, cap]
    uint32_t* cnt;        // [nq]
    float* thr;           // [nq] out: k-th best score so far (or -inf when fewer than k)
    uint32_t cap;
    int k;
    int final_pass;       // 1: also write D / I
    int dense_count;      // >= 0: every list holds exactly this many entries (after a dense phase)
    float* out_d;         // [nq, k]
    int64_t* out_i;       // [nq, k]
    int64_t id_base;      // global id of local row 0
    const float* out_scale;  // [nq] or null: multiply reported scores (undoes the query scaling)
    uint32_t* overflow;   // flag words (flag_slot): [0] = any overflow, [2 + 2q] = query q overflowed
    // Exactness window (tensor-core variant only; null = exact scores, keep exactly k).
    // The scan's scores are APPROXIMATE: |approx - exact| <= eps[q] for every stored row (see
    // prepare_queries_kernel).  Let A be the k-th best approximate score so far.  The k-th best
    // exact score T is >= A - eps, so a row can only belong to the exact top-k if its approximate
    // score exceeds A - 2 eps.  Intermediate passes therefore keep EVERY candidate above
    // A - 2 eps (not a fixed count) and publish that as the filter threshold; after exact
    // rescoring the final pass picks the k best of them.  The result is exact by construction.
    // Only when more than keep_max candidates crowd that window (a cluster of near-ties tighter
    // than the filter's resolution) is the list truncated and the query flagged "uncertified";
    // the host redoes it with a tighter variant.
    const float* win_eps;     // [nq] eps in unscaled score units
    const float* win_scale;   // [nq] approx-score scale (out_scale of the scan)
    int keep_max;             // truncation limit for the window (<= cap)
    uint32_t* uncert;         // like `overflow`; callers pass overflow + 1, which interleaves the two sets
    // Live thresholds (common.cuh, ScanArgs::hist): an intermediate pass re-seeds the query's score
    // histogram with the candidates it keeps and publishes the bucket origin.  Null = off.
    uint32_t* hist;           // [nq, 256]
    uint32_t* hbase;          // [nq]
    // Fused exact rescoring (final pass of the tensor-core variant; null = keys are final): every
    // candidate's score is recomputed from the stored row and the original fp32 query before the k
    // best are picked -- the rescoring kernel and one launch + one pass over the lists saved.
    const void* rs_corpus;    // [n_rows, dim] storage dtype of the kernel instantiation
    const float* rs_q32;      // [nq, dim]
    int rs_dim;
};

// Exact fp32 scores of G candidate rows for one query, by one warp: lane l accumulates the 16-byte
// vectors l, l + 32, ... of every row in order, then the 32 partial sums are combined with xor
// shuffles.  ONE definition for the stand-alone rescoring kernel and the fused rescoring of the
// final select pass, so both give the same bits.  Rows are loaded G at a time: the rows are
// scattered (one per candidate), the work is latency-bound, memory-level parallelism is what counts.
template <typename T, int G>
__device__ __forceinline__ void rescore_group(const T* __restrict__ corpus, const float* __restrict__ qv, int dim,
                                              const uint32_t (&rows)[G], float (&acc)[G], int lane) {
    constexpr int EPV = 16 / sizeof(T);
    const int nvec = dim / EPV;
    const T* x[G];
#pragma unroll
    for (int g = 0; g < G; ++g) { x[g] = corpus + (size_t)rows[g] * dim; acc[g] = 0.f; }
    for (int v = lane; v < nvec; v += 32) {
        uint4 raw[G];
#pragma unroll
        for (int g = 0; g < G; ++g) raw[g] = *reinterpret_cast<const uint4*>(x[g] + (size_t)v * EPV);
        float4 qq[EPV / 4];
#pragma unroll
        for (int h = 0; h < EPV / 4; ++h) qq[h] = *reinterpret_cast<const float4*>(qv + (size_t)v * EPV + 4 * h);
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float xf[EPV];
            VecOps<T>::unpack(raw[g], xf);
#pragma unroll
            for (int h = 0; h < EPV / 4; ++h) {
                acc[g] = fmaf(xf[4 * h + 0], qq[h].x, acc[g]);
                acc[g] = fmaf(xf[4 * h + 1], qq[h].y, acc[g]);
                acc[g] = fmaf(xf[4 * h + 2], qq[h].z, acc[g]);
                acc[g] = fmaf(xf[4 * h + 3], qq[h].w, acc[g]);
            }
        }
    }
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc[g] += __shfl_xor_sync(0xFFFFFFFFu, acc[g], off);
}
constexpr int kRescoreGroup = 4;

constexpr int kSelectThreads = 256;

// Flag layout shared by the flat and IVF paths: F[0] = some list overflowed, F[1] = some query
// could not be certified, then per query q: F[2 + 2q] overflow, F[3 + 2q] uncertified.  The two
// summary words are adjacent so that ONE 8-byte read (K4, or the D2H copy that fetches the
// results) tells whether anything has to be redone.
__host__ __device__ inline size_t flag_slot(int q) { return 2 + 2 * (size_t)q; }
__host__ __device__ inline size_t flag_words(int nq) { return 2 + 2 * (size_t)nq; }

// k-th largest of the c distinct keys in s_keys (1 <= k <= c) by most-significant-digit radix
// select, 8 bits per round, starting at the highest bit in which the keys differ at all.  Keys are
// distinct (the low word encodes the row), so the answer is unique.  Whole CTA; returns the key.
template <int kThreads>
__device__ __forceinline__ uint64_t kth_largest_key(const uint64_t* s_keys, int c, int k,
                                                    uint32_t* s_hist /*[256]*/, uint64_t* s_red /*[2 * warps + 4]*/) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kThreads / 32;
    constexpr int kSelectThreads = kThreads;   // (shadows the default CTA size inside this function)
    // CTA-wide min / max -> the bits above `hi` are common to every key
    uint64_t mn = ~0ull, mx = 0ull;
    for (int i = tid; i < c; i += kSelectThreads) { uint64_t x = s_keys[i]; mn = min(mn, x); mx = max(mx, x); }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, off));
        mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, off));
    }
    if (lane == 0) { s_red[warp] = mn; s_red[kWarps + warp] = mx; }
    __syncthreads();
    mn = s_red[0]; mx = s_red[kWarps];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) { mn = min(mn, s_red[w]); mx = max(mx, s_red[kWarps + w]); }
    __syncthreads();
    if (k == 1) return mx;
    if (k == c) return mn;
    int undecided = 64 - __clzll((long long)(mn ^ mx));  // low bits not yet fixed (>= 1: keys differ)
    uint64_t prefix = undecided >= 64 ? 0ull : (mx >> undecided);
    int kk = k;                                          // rank among the keys matching `prefix`
    while (true) {
        const int shift = undecided > 8 ? undecided - 8 : 0;
        const int nbits = undecided - shift;             // 1..8 bits decided this round
        if (tid < 256) s_hist[tid] = 0u;                 // 256 bins
        __syncthreads();
        for (int i = tid; i < c; i += kSelectThreads) {
            const uint64_t x = s_keys[i];
            const bool match = undecided >= 64 || (x >> undecided) == prefix;
            if (match) atomicAdd(&s_hist[(uint32_t)(x >> shift) & ((1u << nbits) - 1u)], 1u);
        }
        __syncthreads();
        if (warp == 0) {  // bin holding the kk-th largest: walk bins from the top
            uint32_t h[8], sum = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) { h[j] = s_hist[lane * 8 + j]; sum += h[j]; }
            uint32_t suf = sum;  // inclusive suffix sum over lanes (lane 31 = highest bins)
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                uint32_t t = __shfl_down_sync(0xFFFFFFFFu, suf, off);
                if (lane + off < 32) suf += t;
            }
            if (suf >= (uint32_t)kk && suf - sum < (uint32_t)kk) {
                uint32_t above = suf - sum;
                int b = lane * 8;
#pragma unroll
                for (int j = 7; j >= 0; --j) {
                    if (above + h[j] >= (uint32_t)kk) { b = lane * 8 + j; sum = h[j]; break; }
                    above += h[j];
                }
                s_red[0] = (uint64_t)b; s_red[1] = above; s_red[2] = sum;
            }
        }
        __syncthreads();
        const uint32_t b = (uint32_t)s_red[0], above = (uint32_t)s_red[1], in_bin = (uint32_t)s_red[2];
        __syncthreads();
        prefix = (prefix << nbits) | b;
        kk -= (int)above;
        undecided = shift;
        if (undecided == 0) return prefix;
        if (in_bin == 1u) {  // a single key carries this prefix: it is the answer
            for (int i = tid; i < c; i += kSelectThreads)
                if ((s_keys[i] >> undecided) == prefix) s_red[3] = s_keys[i];
            __syncthreads();
            const uint64_t r = s_red[3];
            __syncthreads();
            return r;
        }
    }
}

// One CTA (kThreads threads: 256, or 1024 for the long lists of large k) per query.  Dynamic smem:
// (cap + next_pow2(k)) * 8 bytes.  Intermediate passes only need the k best as a SET (any order)
// and the new threshold: radix select + compaction.  The final pass sorts those k (bitonic over
// next_pow2(k) keys) and emits.
struct NoRescore {};
template <int kThreads, typename T>
__device__ __forceinline__ void select_topk_body(const SelectArgs& a) {
    constexpr int kSelectThreads = kThreads;
    extern __shared__ __align__(16) uint64_t s_keys[];
    __shared__ uint32_t s_hist[256];
    __shared__ uint64_t s_red[2 * (kThreads / 32) + 4];
    __shared__ uint32_t s_count;
    const int q = blockIdx.x, tid = threadIdx.x;
    uint32_t c = a.dense_count >= 0 ? (uint32_t)a.dense_count : a.cnt[q];
    if (c > a.cap) {  // the scan dropped candidates: this query must be redone (host fallback)
        if (tid == 0) { a.overflow[0] = 1u; a.overflow[flag_slot(q)] = 1u; }
        c = a.cap;
    }
    uint64_t* list = a.cand + (size_t)q * a.cap;
    for (int i = tid; i < (int)c; i += kSelectThreads) s_keys[i] = raw_to_key(list[i]);
    if (tid == 0) s_count = 0u;
    __syncthreads();
    if constexpr (!std::is_same<T, NoRescore>::value) {
        if (a.win_eps && (int)c > a.k) {
            // FINISH mode (the pass after the last filter launch): first the exactness window on
            // the APPROXIMATE scores -- exactly what an intermediate pass would keep -- compacted
            // into the spare region behind the list, so that only those candidates are rescored.
            const uint64_t kth_a = kth_largest_key<kThreads>(s_keys, (int)c, a.k, s_hist, s_red);
            uint64_t lim = kth_a - 1ull;
            const float win = 2.0f * a.win_eps[q] / a.win_scale[q];
            const float ta = key_score(kth_a);
            if (win > 0.f && ta > kNegInf)
                lim = ((uint64_t)orderable_f32(nextafterf(ta - win, kNegInf)) << 32) | 0xFFFFFFFFull;
            uint64_t* s_tmp = s_keys + a.cap;
            for (int i = tid; i < (int)c; i += kSelectThreads) {
                const uint64_t x = s_keys[i];
                if (x > lim) {
                    const uint32_t slot = atomicAdd(&s_count, 1u);
                    if ((int)slot < a.keep_max) s_tmp[slot] = x;
                }
            }
            __syncthreads();
            uint32_t kept = s_count;
            if ((int)kept > a.keep_max) {   // crowded window: flagged, the host redoes the query
                if (tid == 0) { a.uncert[0] = 1u; a.uncert[flag_slot(q)] = 1u; }
                kept = (uint32_t)a.keep_max;
            }
            __syncthreads();
            for (int i = tid; i < (int)kept; i += kSelectThreads) s_keys[i] = s_tmp[i];
            if (tid == 0) s_count = 0u;
            c = kept;
            __syncthreads();
        }
        // exact scores for everything (left) in the list (the filter's scores were approximate)
        const int lane = tid & 31, warp = tid >> 5;
        const float* qv = a.rs_q32 + (size_t)q * a.rs_dim;
        for (int g0 = warp * kRescoreGroup; g0 < (int)c; g0 += (kSelectThreads / 32) * kRescoreGroup) {
            uint32_t rows[kRescoreGroup];
            float acc[kRescoreGroup];
#pragma unroll
            for (int g = 0; g < kRescoreGroup; ++g) rows[g] = key_row(s_keys[min(g0 + g, (int)c - 1)]);
            rescore_group<T, kRescoreGroup>(reinterpret_cast<const T*>(a.rs_corpus), qv, a.rs_dim, rows, acc, lane);
            if (lane < kRescoreGroup && g0 + lane < (int)c) {
                float mine = acc[0];
#pragma unroll
                for (int g = 1; g < kRescoreGroup; ++g) if (lane == g) mine = acc[g];
                s_keys[g0 + lane] = make_key(mine, key_row(s_keys[g0 + lane]));
            }
        }
        __syncthreads();
    }
    const int keep = min((int)c, a.k);
    uint64_t kth = 0ull;  // every key >= kth is kept (all of them when c <= k)
    if ((int)c > a.k) kth = kth_largest_key<kThreads>(s_keys, (int)c, a.k, s_hist, s_red);
    if (!a.final_pass) {
        // threshold and the limit key: candidates with key > lim stay in the list
        float t = kNegInf;
        uint64_t lim = 0ull;      // key 0 is below every real key: keep everything
        uint32_t kept = c;
        if ((int)c >= a.k) {
            if ((int)c == a.k) {  // exactly k: the k-th best is the smallest key
                uint64_t mn = ~0ull;
                for (int i = tid; i < (int)c; i += kSelectThreads) mn = min(mn, s_keys[i]);
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) mn = min(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, off));
                if ((tid & 31) == 0) s_red[tid >> 5] = mn;
                __syncthreads();
                mn = s_red[0];
                for (int w = 1; w < kSelectThreads / 32; ++w) mn = min(mn, s_red[w]);
                __syncthreads();
                kth = mn;
            }
            t = key_score(kth);
            lim = kth - 1ull;
            kept = (uint32_t)a.k;
            if (a.win_eps) {
                const float win = 2.0f * a.win_eps[q] / a.win_scale[q];
                if (win > 0.f && t > kNegInf) {
                    t = nextafterf(t - win, kNegInf);
                    lim = ((uint64_t)orderable_f32(t) << 32) | 0xFFFFFFFFull;  // keep score > t
                    kept = 0xFFFFFFFFu;                                         // counted below
                }
            }
        }
        uint64_t keep_from = lim + 1ull;   // every key >= keep_from stays in the list
        if ((int)c > a.k || kept == 0xFFFFFFFFu) {
            for (int i = tid; i < (int)c; i += kSelectThreads) {
                const uint64_t x = s_keys[i];
                if (x > lim) list[atomicAdd(&s_count, 1u)] = key_to_raw(x);
            }
            __syncthreads();
            kept = s_count;
            if ((int)kept > a.keep_max && a.win_eps) {
                // the window is crowded: truncate to the keep_max best and flag the query
                __syncthreads();
                if (tid == 0) { s_count = 0u; a.uncert[0] = 1u; a.uncert[flag_slot(q)] = 1u; }
                const uint64_t kth2 = kth_largest_key<kThreads>(s_keys, (int)c, a.keep_max, s_hist, s_red);
                for (int i = tid; i < (int)c; i += kSelectThreads) {
                    const uint64_t x = s_keys[i];
                    if (x >= kth2) list[atomicAdd(&s_count, 1u)] = key_to_raw(x);
                }
                t = key_score(kth2);
                kept = (uint32_t)a.keep_max;
                keep_from = kth2;
            }
        }
        if (a.hist) {
            __syncthreads();
            if (tid < kHistBins) s_hist[tid] = 0u;
            __syncthreads();
            const uint32_t hb = orderable_f32(t);
            for (int i = tid; i < (int)c; i += kSelectThreads) {
                const uint64_t x = s_keys[i];
                if (x >= keep_from) {
                    const uint32_t d = ((uint32_t)(x >> 32) - hb) >> kHistShift;
                    atomicAdd(&s_hist[d < (uint32_t)(kHistBins - 1) ? d : (uint32_t)(kHistBins - 1)], 1u);
                }
            }
            __syncthreads();
            if (tid < kHistBins) a.hist[(size_t)q * kHistBins + tid] = s_hist[tid];
            if (tid == 0) a.hbase[q] = hb;
        }
        if (tid == 0) {
            a.cnt[q] = kept;
            a.thr[q] = t;
        }
        return;
    }
    // final pass: gather the survivors, sort them, emit
    int n = 1;
    while (n < keep) n <<= 1;
    uint64_t* s_out = s_keys + a.cap;
    for (int i = tid; i < n; i += kSelectThreads) s_out[i] = 0ull;
    __syncthreads();
    for (int i = tid; i < (int)c; i += kSelectThreads) {
        const uint64_t x = s_keys[i];
        if (x >= kth) s_out[atomicAdd(&s_count, 1u)] = x;
    }
    __syncthreads();
    bitonic_sort_desc(s_out, n);
    const float scale = a.out_scale ? a.out_scale[q] : 1.0f;
    for (int i = tid; i < a.k; i += kSelectThreads) {
        uint64_t key = i < keep ? s_out[i] : 0ull;
        bool valid = key != 0ull;
        a.out_d[(size_t)q * a.k + i] = valid ? key_score(key) * scale : kLowestF32;
        a.out_i[(size_t)q * a.k + i] = valid ? a.id_base + (int64_t)key_row(key) : -1;
    }
}
__global__ void __launch_bounds__(kSelectThreads) select_topk_kernel(SelectArgs a) { select_topk_body<kSelectThreads, NoRescore>(a); }
constexpr int kSelectThreadsBig = 1024;
__global__ void __launch_bounds__(kSelectThreadsBig) select_topk_kernel_big(SelectArgs a) { select_topk_body<kSelectThreadsBig, NoRescore>(a); }
// final pass with fused exact rescoring, per storage type
template <typename T>
__global__ void __launch_bounds__(kSelectThreads) select_rescore_kernel(SelectArgs a) { select_topk_body<kSelectThreads, T>(a); }
template <typename T>
__global__ void __launch_bounds__(kSelectThreadsBig) select_rescore_kernel_big(SelectArgs a) { select_topk_body<kSelectThreadsBig, T>(a); }
// Lists of 8 K entries and more (k >= 256) get the 1024-thread form: one CTA per SM either way (the
// list fills the shared memory), so the per-key passes run 4x wider.
inline bool select_wants_big(int cap) { return cap >= 8192; }

// Exact rescoring of the surviving candidates (tensor-core variant).  The tcgen05 scan sees the
// fp32 queries rounded to 16 bits, so its scores carry ~1e-4 relative error; that is fine for
// FILTERING (the lists keep a margin of extra candidates) but not for the reported distances or
// the final order.  One warp per (query, candidate): re-read the stored row (exact 16-bit values)
// and the original fp32 query, accumulate in fp32, and rewrite the key with the exact score.
// Work: nq * k' rows of dim elements -- microseconds next to the scan.
template <typename T>
__global__ void rescore_kernel(uint64_t* __restrict__ cand, const uint32_t* __restrict__ cnt,
                               uint32_t cap, int nq, int kprime, const T* __restrict__ corpus,
                               const float* __restrict__ q32, int dim) {
    constexpr int G = kRescoreGroup;
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int groups = (kprime + G - 1) / G;            // per query: slots [G g, G g + G)
    for (int64_t w = warp; w < (int64_t)nq * groups; w += nwarps) {
        const int q = (int)(w / groups), s0 = (int)(w - (int64_t)q * groups) * G;
        const int have = (int)min(cnt[q], cap);
        if (s0 >= have) continue;  // warp-uniform
        uint64_t key[G];
        uint32_t rows[G];
        float acc[G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
            key[g] = cand[(size_t)q * cap + min(s0 + g, have - 1)];
            rows[g] = key_row(key[g]);
        }
        rescore_group<T, G>(corpus, q32 + (size_t)q * dim, dim, rows, acc, lane);
#pragma unroll
        for (int g = 0; g < G; ++g)
            if (lane == 0 && s0 + g < have) cand[(size_t)q * cap + s0 + g] = make_raw_key(acc[g], rows[g]);
    }
}

// Cross-shard merge.  in_d / in_i: [g, nq, k] (rank-major, as an all-gather leaves them); every
// shard list is best-first with id -1 padding, and shard r holds lower ids than shard r+1
// (contiguous row sharding), so (shard, position) ascending == id ascending among equal scores.
struct MergeArgs {
    const float* in_d;
    const int64_t* in_i;
    int g, nq, k_in, k_out;
    int64_t stride_d, stride_i;  // elements between shard r and shard r+1 (nq * k_in when packed)
    float* out_d;      // [nq, k_out]
    int64_t* out_i;    // [nq, k_out]
};

__global__ void merge_topk_kernel(MergeArgs a) {
    extern __shared__ __align__(16) uint64_t s_keys[];
    const int q = blockIdx.x;
    const int total = a.g * a.k_in;
    int n = 1;
    while (n < total) n <<= 1;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        uint64_t key = 0ull;
        if (i < total) {
            int r = i / a.k_in, p = i - r * a.k_in;
            size_t off = (size_t)q * a.k_in + p;
            if (a.in_i[(size_t)r * a.stride_i + off] >= 0)
                key = make_key(a.in_d[(size_t)r * a.stride_d + off], (uint32_t)i);
        }
        s_keys[i] = key;
    }
    __syncthreads();
    bitonic_sort_desc(s_keys, n);
    for (int i = threadIdx.x; i < a.k_out; i += blockDim.x) {
        uint64_t key = i < n ? s_keys[i] : 0ull;
        float d = kLowestF32;
        int64_t id = -1;
        if (key != 0ull) {
            int src = (int)key_row(key);
            int r = src / a.k_in, p = src - r * a.k_in;
            size_t off = (size_t)q * a.k_in + p;
            d = a.in_d[(size_t)r * a.stride_d + off];
            id = a.in_i[(size_t)r * a.stride_i + off];
        }
        a.out_d[(size_t)q * a.k_out + i] = d;
        a.out_i[(size_t)q * a.k_out + i] = id;
    }
}

}  // namespace fip
