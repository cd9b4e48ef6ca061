// Device-side synthetic corpus generator.  Bit-identical to synthetic.py (numpy) and to the C
// restatement in oracle/: Philox4x32-10, counter (row_lo, row_hi, col/2, stream), key = seed,
// element = (sum of the four 16-bit halves of two output words - 131070) * scale, then RNE
// rounding to the storage dtype.  Stands in for the reference's random-normal corpus
// (/root/reference/scripts/create_test_docs.py:87-96) without 150 GB of host I/O.
#pragma once

#include "common.cuh"

namespace fip {

__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

__device__ __forceinline__ float irwin_hall4(uint32_t a, uint32_t b) {
    int s = (int)((a & 0xFFFFu) + (a >> 16) + (b & 0xFFFFu) + (b >> 16)) - 131070;
    return __int2float_rn(s) * (float)(1.0 / 37837.22723876);
}

template <typename T> __device__ __forceinline__ T to_storage(float v);
template <> __device__ __forceinline__ float to_storage<float>(float v) { return v; }
template <> __device__ __forceinline__ __half to_storage<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 to_storage<__nv_bfloat16>(float v) {
    return __float2bfloat16_rn(v);
}

// Fills local rows [0, nrows) of `out` with global rows [global_row0, global_row0 + nrows).
template <typename T>
__global__ void synthetic_fill_kernel(T* __restrict__ out, int64_t nrows, int dim,
                                      int64_t global_row0, uint64_t seed, uint32_t stream) {
    const int npairs = (dim + 1) >> 1;
    const int64_t total = nrows * (int64_t)npairs;
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / npairs;
        int p = (int)(i - r * npairs);
        uint64_t row = (uint64_t)(global_row0 + r);
        uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)p, stream};
        philox4x32_10(c, k0, k1);
        T* o = out + r * (int64_t)dim + 2 * p;
        o[0] = to_storage<T>(irwin_hall4(c[0], c[1]));
        if (2 * p + 1 < dim) o[1] = to_storage<T>(irwin_hall4(c[2], c[3]));
    }
}

// fp32 host/device rows -> storage dtype (RNE), used by fip_add_*.
template <typename T>
__global__ void convert_rows_kernel(const float* __restrict__ in, T* __restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = to_storage<T>(in[i]);
}

template <typename T> __device__ __forceinline__ float from_storage(T v);
template <> __device__ __forceinline__ float from_storage<float>(float v) { return v; }
template <> __device__ __forceinline__ float from_storage<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float from_storage<__nv_bfloat16>(__nv_bfloat16 v) {
    return __bfloat162float(v);
}

// max over rows of ||x_r||^2 (fp32, from the stored values): the exactness certificate of the
// tensor-core variant bounds |<dq, x_r>| by ||dq|| * max ||x_r||.  One warp per row; squared norms
// are non-negative, so their bit patterns order like unsigned integers (atomicMax on the bits).
template <typename T>
__global__ void row_norm_max_kernel(const T* __restrict__ rows, int64_t n, int dim,
                                    float* __restrict__ max_norm2) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    float best = 0.f;
    for (int64_t r = warp; r < n; r += nwarps) {
        const T* x = rows + r * (int64_t)dim;
        float acc = 0.f;
        for (int c = lane; c < dim; c += 32) {
            const float v = from_storage<T>(x[c]);
            acc = fmaf(v, v, acc);
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
        best = fmaxf(best, acc);
    }
    if (lane == 0 && best > 0.f) atomicMax(reinterpret_cast<unsigned int*>(max_norm2), __float_as_uint(best));
}

template <typename T>
__global__ void widen_rows_kernel(const T* __restrict__ in, float* __restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = from_storage<T>(in[i]);
}

}  // namespace fip
