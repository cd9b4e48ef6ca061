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
