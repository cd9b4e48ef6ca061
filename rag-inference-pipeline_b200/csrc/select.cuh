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
    float* out_d;         // [nq, k]
    int64_t* out_i;       // [nq, k]
    int64_t id_base;      // global id of local row 0
    const float* out_scale;  // [nq] or null: multiply reported scores (undoes the query scaling)
    uint32_t* overflow;   // [1 + nq]: [0] = any overflow, [1+q] = query q overflowed
};

// One CTA per query.  Dynamic smem: next_pow2(cap) * 8 bytes.
__global__ void select_topk_kernel(SelectArgs a) {
    extern __shared__ __align__(16) uint64_t s_keys[];
    const int q = blockIdx.x;
    uint32_t c = a.cnt[q];
    if (c > a.cap) {  // the scan dropped candidates: this query must be redone (host fallback)
        if (threadIdx.x == 0) { a.overflow[0] = 1u; a.overflow[1 + q] = 1u; }
        c = a.cap;
    }
    int n = 1;
    while (n < (int)c) n <<= 1;
    const uint64_t* src = a.cand + (size_t)q * a.cap;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_keys[i] = i < (int)c ? src[i] : 0ull;
    __syncthreads();
    bitonic_sort_desc(s_keys, n);
    const int keep = min((int)c, a.k);
    if (!a.final_pass) {
        uint64_t* dst = a.cand + (size_t)q * a.cap;
        for (int i = threadIdx.x; i < keep; i += blockDim.x) dst[i] = s_keys[i];
        if (threadIdx.x == 0) {
            a.cnt[q] = (uint32_t)keep;
            a.thr[q] = keep >= a.k ? key_score(s_keys[a.k - 1]) : kNegInf;
        }
    } else {
        const float scale = a.out_scale ? a.out_scale[q] : 1.0f;
        for (int i = threadIdx.x; i < a.k; i += blockDim.x) {
            uint64_t key = i < keep ? s_keys[i] : 0ull;
            bool valid = key != 0ull;
            a.out_d[(size_t)q * a.k + i] = valid ? key_score(key) * scale : kLowestF32;
            a.out_i[(size_t)q * a.k + i] = valid ? a.id_base + (int64_t)key_row(key) : -1;
        }
    }
}

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
    constexpr int EPV = 16 / sizeof(T);
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nvec = dim / EPV;
    for (int64_t w = warp; w < (int64_t)nq * kprime; w += nwarps) {
        const int q = (int)(w / kprime), slot = (int)(w - (int64_t)q * kprime);
        if (slot >= (int)min(cnt[q], cap)) continue;  // warp-uniform
        const uint64_t key = cand[(size_t)q * cap + slot];
        const uint32_t row = key_row(key);
        const T* x = corpus + (size_t)row * dim;
        const float* qv = q32 + (size_t)q * dim;
        float acc = 0.f;
        for (int v = lane; v < nvec; v += 32) {
            uint4 raw = *reinterpret_cast<const uint4*>(x + (size_t)v * EPV);
            float xf[EPV];
            VecOps<T>::unpack(raw, xf);
#pragma unroll
            for (int h = 0; h < EPV / 4; ++h) {
                float4 qq = *reinterpret_cast<const float4*>(qv + (size_t)v * EPV + 4 * h);
                acc = fmaf(xf[4 * h + 0], qq.x, acc);
                acc = fmaf(xf[4 * h + 1], qq.y, acc);
                acc = fmaf(xf[4 * h + 2], qq.z, acc);
                acc = fmaf(xf[4 * h + 3], qq.w, acc);
            }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
        if (lane == 0) cand[(size_t)q * cap + slot] = make_key(acc, row);
    }
}

// Cross-shard merge.  in_d / in_i: [g, nq, k] (rank-major, as an all-gather leaves them); every
// shard list is best-first with id -1 padding, and shard r holds lower ids than shard r+1
// (contiguous row sharding), so (shard, position) ascending == id ascending among equal scores.
struct MergeArgs {
    const float* in_d;
    const int64_t* in_i;
    int g, nq, k_in, k_out;
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
            size_t off = ((size_t)r * a.nq + q) * a.k_in + p;
            if (a.in_i[off] >= 0) key = make_key(a.in_d[off], (uint32_t)i);
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
            size_t off = ((size_t)r * a.nq + q) * a.k_in + p;
            d = a.in_d[off];
            id = a.in_i[off];
        }
        a.out_d[(size_t)q * a.k_out + i] = d;
        a.out_i[(size_t)q * a.k_out + i] = id;
    }
}

}  // namespace fip
