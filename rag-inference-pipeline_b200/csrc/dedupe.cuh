// In-batch exact-duplicate elimination of queries, on the device.
//
// The reference's caller keeps an LRU keyed by sha256(embedding.tobytes())
// (/root/reference/src/pipeline/services/retrieval/api.py:392-425): two queries are "the same"
// exactly when their fp32 bytes are equal.  A coalesced batch of many callers is where such
// duplicates meet (popular questions, retries), and in the tensor-bound regime every query costs.
// Before the search the batch is reduced to its distinct rows, afterwards every caller's row is
// expanded from its representative's result:
//   hash     one warp per query: 64-bit multiplicative hash of the row's 32-bit words;
//   rep      one warp per query q: first j < q with equal hash AND bit-equal row (the compare is
//            exact, a hash collision can only cost time), else q itself;
//   compact  one block: distinct rows keep their order; sel[] (distinct -> original row) and
//            map[] (original row -> distinct index), count to the host with the plan;
//   gather   distinct rows -> contiguous query matrix;        expand   result rows -> every caller.
#pragma once

#include "common.cuh"

namespace fip {

__global__ void query_hash_kernel(const float* __restrict__ q, int nq, int dim, uint64_t* __restrict__ hash) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= nq) return;
    const uint32_t* w = reinterpret_cast<const uint32_t*>(q + (size_t)row * dim);
    uint64_t h = 0x9E3779B97F4A7C15ull * (uint64_t)(lane + 1);
    for (int c = lane; c < dim; c += 32) h = (h ^ w[c]) * 0xFF51AFD7ED558CCDull + 0xC4CEB9FE1A85EC53ull;
    h ^= h >> 29;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) h = (h * 0x2545F4914F6CDD1Dull) ^ __shfl_xor_sync(0xFFFFFFFFu, h, off);
    if (lane == 0) hash[row] = h;
}

__global__ void query_rep_kernel(const float* __restrict__ q, const uint64_t* __restrict__ hash, int nq,
                                 int dim, int* __restrict__ rep) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= nq) return;
    const uint64_t h = hash[row];
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(q + (size_t)row * dim);
    int found = row;
    for (int j0 = 0; j0 < row && found == row; j0 += 32) {
        const int j = j0 + lane;
        uint32_t cand = __ballot_sync(0xFFFFFFFFu, j < row && hash[j] == h);
        while (cand) {   // ascending j: the FIRST equal row becomes the representative
            const int jj = j0 + __ffs((int)cand) - 1;
            cand &= cand - 1u;
            const uint32_t* other = reinterpret_cast<const uint32_t*>(q + (size_t)jj * dim);
            bool same = true;
            for (int c = lane; c < dim; c += 32) same = same && mine[c] == other[c];
            if (__all_sync(0xFFFFFFFFu, same)) { found = jj; break; }
        }
    }
    if (lane == 0) rep[row] = found;
}

// One block of 1024 threads.  out_count[0] = number of distinct rows.
__global__ void query_compact_kernel(const int* __restrict__ rep, int nq, int* __restrict__ pos,
                                     int* __restrict__ sel, int* __restrict__ map, int* __restrict__ out_count) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int q0 = 0; q0 < nq; q0 += 1024) {
        const int q = q0 + tid;
        const int flag = (q < nq && rep[q] == q) ? 1 : 0;
        int inc = flag;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int t = __shfl_up_sync(0xFFFFFFFFu, inc, off);
            if (lane >= off) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int v = s_warp[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, v, off);
                if (lane >= off) v += t;
            }
            s_warp[lane] = v;
        }
        __syncthreads();
        const int before = s_base + (warp ? s_warp[warp - 1] : 0) + inc - flag;
        if (flag) { pos[q] = before; sel[before] = q; }
        __syncthreads();
        if (tid == 0) s_base += s_warp[31];
        __syncthreads();
    }
    for (int q = tid; q < nq; q += 1024) map[q] = pos[rep[q]];   // rep[q] <= q is distinct: its pos is set
    if (tid == 0) out_count[0] = s_base;
}

// out rows [q] = in rows [map[q]]  (distances and ids of the representative's result)
__global__ void expand_results_kernel(const float* __restrict__ sd, const int64_t* __restrict__ si,
                                      const int* __restrict__ map, float* __restrict__ dd,
                                      int64_t* __restrict__ di, int nq, int k) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < (int64_t)nq * k) {
        const int q = (int)(i / k), c = (int)(i - (int64_t)q * k);
        dd[i] = sd[(size_t)map[q] * k + c];
        di[i] = si[(size_t)map[q] * k + c];
    }
}

}  // namespace fip
