// K2: bandwidth-bound streaming scan (CUDA cores, exact fp32 FMA accumulation).
//
// Replaces faiss's exhaustive_inner_product_seq / the small-batch end of
// exhaustive_inner_product_blas behind IndexFlatIP.search
// (/root/reference/src/pipeline/components/faiss_store.py:152).  Used when the batch is so small
// that the scan is HBM-bound on CUDA cores, and for fp32 storage (exact fp32 products).
//
// Layout: the 32 lanes of a warp own consecutive 16-byte vectors of a corpus row, so every global
// load instruction of a warp covers 512 contiguous bytes (ld.global.nc, L1 no-allocate: the corpus
// is read exactly once).  A warp keeps R rows x QB queries of partial sums in registers, reads the
// query vectors from shared memory (fp32, staged once per CTA), and finishes each row group with a
// shuffle reduce-scatter so that lane l ends up owning (row, query) pair l: that lane compares the
// score with the query's running threshold and appends the rare survivors to the candidate list.
#pragma once

#include "common.cuh"

namespace fip {

__device__ __forceinline__ uint4 ldg_stream_16B(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

template <typename T> struct VecOps;
template <> struct VecOps<float> {
    static constexpr int EPV = 4;
    __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[4]) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
        f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    }
};
template <> struct VecOps<__half> {
    static constexpr int EPV = 8;
    __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[8]) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
            float2 t = __half22float2(h);
            f[2 * i] = t.x; f[2 * i + 1] = t.y;
        }
    }
};
template <> struct VecOps<__nv_bfloat16> {
    static constexpr int EPV = 8;
    __device__ static __forceinline__ void unpack(const uint4& v, float (&f)[8]) {
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
        }
    }
};

// Reduce V per-lane partial sums over the 32 lanes.  V >= 32: reduce-scatter, lane l ends with the
// totals of flat indices l*(V/32) + t in v[t].  V < 32: lanes l with l % (32/V) == 0 hold the total
// of flat index l / (32/V) in v[0].
template <int V>
__device__ __forceinline__ void warp_reduce_scatter(float (&v)[V], int lane) {
    constexpr unsigned kFull = 0xFFFFFFFFu;
#pragma unroll
    for (int step = 0; step < 5; ++step) {
        const int off = 16 >> step;
        const int n = V >> step;  // live values before this step (may reach 0 -> treat as 1)
        if (n >= 2) {
            const bool upper = (lane & off) != 0;
#pragma unroll
            for (int i = 0; i < (n >> 1); ++i) {
                float send = upper ? v[i] : v[i + (n >> 1)];
                float keep = upper ? v[i + (n >> 1)] : v[i];
                v[i] = keep + __shfl_xor_sync(kFull, send, off);
            }
        } else {
            v[0] += __shfl_xor_sync(kFull, v[0], off);
        }
    }
}

// queries: fp32 [nq, dim] on device.  This launch handles queries [q0, q0 + QB).
template <typename T, int QB, int R>
__global__ void __launch_bounds__(256)
stream_scan_kernel(ScanArgs a, const float* __restrict__ queries, int q0) {
    constexpr int EPV = VecOps<T>::EPV;
    constexpr int V = QB * R;
    constexpr int PER = V >= 32 ? V / 32 : 1;
    constexpr int REP = V >= 32 ? 1 : 32 / V;
    extern __shared__ __align__(16) float s_q[];  // [QB][dim]
    __shared__ float s_thr[QB];

    const int dim = a.dim;
    for (int i = threadIdx.x; i < QB * dim; i += blockDim.x) {
        int j = i / dim, c = i - j * dim;
        s_q[i] = (q0 + j < a.nq) ? queries[(size_t)(q0 + j) * dim + c] : 0.f;
    }
    if (threadIdx.x < QB)
        s_thr[threadIdx.x] = (q0 + (int)threadIdx.x < a.nq) ? a.thr[q0 + threadIdx.x] : 0.f;
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int warps_per_cta = blockDim.x >> 5;
    const int64_t gwarp = (int64_t)blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * warps_per_cta;
    const int nvec = dim / EPV;
    const T* __restrict__ corpus = reinterpret_cast<const T*>(a.corpus);
    const int64_t ngroups = ((int64_t)(a.row_end - a.row_begin) + R - 1) / R;

    for (int64_t g = gwarp; g < ngroups; g += nwarps) {
        const uint32_t r0 = a.row_begin + (uint32_t)(g * R);
        float acc[V];
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = 0.f;

        for (int v = lane; v < nvec; v += 32) {
            uint4 xv[R];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                xv[r] = make_uint4(0u, 0u, 0u, 0u);
                if (r0 + r < a.row_end)
                    xv[r] = ldg_stream_16B(corpus + (size_t)(r0 + r) * dim + (size_t)v * EPV);
            }
            float xf[R][EPV];
#pragma unroll
            for (int r = 0; r < R; ++r) VecOps<T>::unpack(xv[r], xf[r]);
#pragma unroll
            for (int j = 0; j < QB; ++j) {
                const float4* qp = reinterpret_cast<const float4*>(s_q + (size_t)j * dim + v * EPV);
#pragma unroll
                for (int h = 0; h < EPV / 4; ++h) {
                    float4 qv = qp[h];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float s = acc[r * QB + j];
                        s = fmaf(xf[r][4 * h + 0], qv.x, s);
                        s = fmaf(xf[r][4 * h + 1], qv.y, s);
                        s = fmaf(xf[r][4 * h + 2], qv.z, s);
                        s = fmaf(xf[r][4 * h + 3], qv.w, s);
                        acc[r * QB + j] = s;
                    }
                }
            }
        }
        warp_reduce_scatter<V>(acc, lane);

        if (lane % REP == 0) {
#pragma unroll
            for (int t = 0; t < PER; ++t) {
                const int flat = (V >= 32) ? lane * PER + t : lane / REP;
                const int r = flat / QB, j = flat - r * QB;
                const uint32_t row = r0 + r;
                const int q = q0 + j;
                if (row < a.row_end && q < a.nq) {
                    const float s = acc[t];
                    if (a.dense) {
                        a.cand[(size_t)q * a.cap + (row - a.row_begin)] = make_raw_key(s, row);
                    } else if (s > s_thr[j]) {
                        append_candidate(a, q, s, row);
                    }
                }
            }
        }
    }
}

}  // namespace fip
