// K2: bandwidth-bound streaming scan (CUDA cores, exact fp32 FMA accumulation).
//
// Replaces faiss's exhaustive_inner_product_seq / the small-batch end of
// exhaustive_inner_product_blas behind IndexFlatIP.search
// (/root/reference/src/pipeline/components/faiss_store.py:152).  Used when the batch is so small
// that the scan is HBM-bound on CUDA cores, and for fp32 storage (exact fp32 products).
//
// Layout: corpus rows stream HBM -> shared memory through per-warp rings of 1-D bulk copies
// (cp.async.bulk, the TMA engine: no registers held by loads in flight), the 32 lanes of a warp
// then read consecutive 16-byte vectors of a row (128-bit, conflict-free).  A warp keeps R rows x
// QB queries of partial sums in registers, reads the query vectors from shared memory (fp32,
// staged once per CTA), and finishes each row block with a shuffle reduce-scatter so that lane l
// ends up owning (row, query) pair l: that lane compares the score with the query's running
// threshold and appends the rare survivors to the candidate list.
#pragma once

#include "common.cuh"

namespace fip {

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

// ---------------------------------------------------------------------------------------------
// Bulk-copy plumbing (raw PTX; K1 has its own copies in tc_ptx.cuh for the tensor-map forms).
__device__ __forceinline__ uint32_t ss_smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void ss_mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void ss_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ss_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n\t"
        "@P1 bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t"
        "}" ::"r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
}
// 1-D bulk copy global -> shared (TMA engine, SASS UBLKCP): `bytes` contiguous bytes, 16-byte
// aligned on both sides, completion counted in bytes on `bar`.
__device__ __forceinline__ void ss_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
        "l"(src), "r"(bytes), "r"(bar)
        : "memory");
}
__device__ __forceinline__ uint4 ss_lds_16B(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "r"(addr));
    return r;
}

constexpr int kStreamMaxWarps = 16;

// Host-side launch plan of the streaming kernel: warps per CTA and ring depth per warp.
struct StreamPlan {
    int warps = 0, slots = 0;
    size_t smem = 0;
};
// A warp's ring holds `slots` blocks of R rows; all rings + the fp32 queries must fit in 227 KB.
// More warps hide FMA / LDS latency, more slots hide HBM latency: prefer >= 3 slots, then warps.
inline StreamPlan plan_stream(int qb, int rows_per_block, size_t row_bytes, int dim) {
    const size_t budget = 227 * 1024 - 1024;
    const size_t qbytes = ((size_t)qb * dim * sizeof(float) + 127) / 128 * 128;
    const size_t blk = (size_t)rows_per_block * row_bytes;
    StreamPlan best;
    for (int w : {16, 8, 4, 2, 1}) {
        if (qbytes + (size_t)w * 2 * (blk + 8) > budget) continue;
        int s = (int)((budget - qbytes) / ((size_t)w * (blk + 8)));
        s = s > 8 ? 8 : s;
        if (s >= 3 || w == 1) { best.warps = w; best.slots = s; break; }
        if (best.warps == 0) { best.warps = w; best.slots = s; }  // 2 slots: keep looking for 3+
    }
    if (best.warps) best.smem = qbytes + (size_t)best.warps * best.slots * (blk + 8);
    return best;
}

// queries: fp32 [nq, dim] on device.  This launch handles queries [q0, q0 + QB).
//
// One persistent CTA per SM.  Every WARP owns a private ring of `slots` shared-memory blocks of
// R consecutive corpus rows (rows are contiguous in HBM, so a block is ONE 1-D bulk copy) and its
// own mbarriers: lane 0 keeps `slots` copies in flight, the warp consumes them in order.  No
// cross-warp synchronisation after the prologue, and the bytes in flight per SM
// (warps x slots x R x row_bytes, ~190 KB) do not depend on register allocation or occupancy.
// Shared memory reads are 16 bytes per lane, consecutive lanes consecutive vectors: conflict-free.
template <typename T, int QB, int R>
__global__ void __launch_bounds__(kStreamMaxWarps * 32, 1)
stream_scan_kernel(ScanArgs a, const float* __restrict__ queries, int q0, int slots) {
    constexpr int EPV = VecOps<T>::EPV;
    constexpr int V = QB * R;
    constexpr int PER = V >= 32 ? V / 32 : 1;
    constexpr int REP = V >= 32 ? 1 : 32 / V;
    extern __shared__ __align__(128) uint8_t s_raw[];

    const int dim = a.dim;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warps_per_cta = blockDim.x >> 5;
    const uint32_t row_bytes = (uint32_t)dim * sizeof(T);
    const uint32_t blk = R * row_bytes;
    float* s_q = reinterpret_cast<float*>(s_raw);  // [QB][dim]
    const uint32_t qbytes = ((uint32_t)(QB * dim * sizeof(float)) + 127u) / 128u * 128u;
    const uint32_t ring0 = ss_smem_u32(s_raw) + qbytes + (uint32_t)warp * (uint32_t)slots * blk;
    const uint32_t bar0 = ss_smem_u32(s_raw) + qbytes + (uint32_t)warps_per_cta * (uint32_t)slots * blk +
                          (uint32_t)warp * (uint32_t)slots * 8u;

    for (int i = threadIdx.x; i < QB * dim; i += blockDim.x) {
        int j = i / dim, c = i - j * dim;
        s_q[i] = (q0 + j < a.nq) ? queries[(size_t)(q0 + j) * dim + c] : 0.f;
    }
    if (lane == 0)
        for (int s = 0; s < slots; ++s) ss_mbar_init(bar0 + 8u * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    const int64_t gwarp = (int64_t)blockIdx.x * warps_per_cta + warp;
    const int64_t nwarps = (int64_t)gridDim.x * warps_per_cta;
    const int nvec = dim / EPV;
    const uint32_t n_rows = a.row_end - a.row_begin;
    const int64_t nblocks = ((int64_t)n_rows + R - 1) / R;
    const int64_t n_my = gwarp < nblocks ? (nblocks - gwarp + nwarps - 1) / nwarps : 0;
    const uint8_t* __restrict__ corpus = reinterpret_cast<const uint8_t*>(a.corpus);
    const uint32_t s_q_addr = ss_smem_u32(s_q);

    auto issue = [&](int64_t i) {  // lane 0 only
        const int slot = (int)(i % slots);
        const int64_t b = gwarp + i * nwarps;
        const uint32_t r0 = (uint32_t)(b * R);
        const uint32_t rows_in = min((uint32_t)R, n_rows - r0);
        const uint32_t bytes = rows_in * row_bytes;
        ss_mbar_expect_tx(bar0 + 8u * slot, bytes);
        ss_bulk_g2s(ring0 + (uint32_t)slot * blk, corpus + (size_t)(a.row_begin + r0) * row_bytes, bytes,
                    bar0 + 8u * slot);
    };
    if (lane == 0)
        for (int64_t i = 0; i < min((int64_t)slots, n_my); ++i) issue(i);

    for (int64_t i = 0; i < n_my; ++i) {
        const int slot = (int)(i % slots);
        const uint32_t parity = (uint32_t)((i / slots) & 1);
        const uint32_t r0 = a.row_begin + (uint32_t)((gwarp + i * nwarps) * R);
        // this lane's (row, query) pairs are fixed: fetch their thresholds now (L2: they are
        // raised live by whoever appends candidates), the latency hides behind the block's FMAs
        float thr_reg[PER];
#pragma unroll
        for (int t = 0; t < PER; ++t) {
            const int flat = (V >= 32) ? lane * PER + t : lane / REP;
            const int q = q0 + flat % QB;
            thr_reg[t] = (q < a.nq && !a.dense) ? __ldcg(a.thr + q) : 0.f;
        }
        ss_mbar_wait(bar0 + 8u * slot, parity);
        const uint32_t src = ring0 + (uint32_t)slot * blk;

        float acc[V];
#pragma unroll
        for (int t = 0; t < V; ++t) acc[t] = 0.f;
#pragma unroll 2
        for (int v = lane; v < nvec; v += 32) {
            float xf[R][EPV];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint4 raw = ss_lds_16B(src + (uint32_t)r * row_bytes + (uint32_t)v * 16u);
                VecOps<T>::unpack(raw, xf[r]);
            }
#pragma unroll
            for (int j = 0; j < QB; ++j) {
#pragma unroll
                for (int h = 0; h < EPV / 4; ++h) {
                    const uint4 qr = ss_lds_16B(s_q_addr + (uint32_t)((j * dim + v * EPV + 4 * h) * 4));
                    const float qx = __uint_as_float(qr.x), qy = __uint_as_float(qr.y);
                    const float qz = __uint_as_float(qr.z), qw = __uint_as_float(qr.w);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float s = acc[r * QB + j];
                        s = fmaf(xf[r][4 * h + 0], qx, s);
                        s = fmaf(xf[r][4 * h + 1], qy, s);
                        s = fmaf(xf[r][4 * h + 2], qz, s);
                        s = fmaf(xf[r][4 * h + 3], qw, s);
                        acc[r * QB + j] = s;
                    }
                }
            }
        }
        // the slot is consumed: refill it before the (short) reduction and emit
        __syncwarp();
        if (lane == 0 && i + slots < n_my) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(i + slots);
        }
        warp_reduce_scatter<V>(acc, lane);

        bool need = false;   // one of this lane's appends crossed a refresh boundary
        int need_q = 0;
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
                    } else if (s > thr_reg[t]) {
                        if (append_candidate(a, q, s, row)) { need = true; need_q = q; }
                    }
                }
            }
        }
        __syncwarp();
        if (a.hist) refresh_thresholds_warp(a, need, need_q, lane);
    }
}

}  // namespace fip
