// Inline-PTX wrappers for the sm_100a features K1 uses: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (alloc / mma / commit / ld), thread-block-cluster helpers for the CTA-pair (cta_group::2)
// variant.  No CUTLASS/CuTe: these are the raw instructions.
#pragma once

#include <cuda.h>

#include "common.cuh"

namespace fip {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

// ------------------------------------------------------------------ cluster
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address -> shared::cluster address of the same offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}
// The two "accumulator drained" arrivals below are RELAXED on purpose.  What they order is the
// epilogue's TMEM reads before the next MMA into the same accumulator, and that is already done:
// tcgen05.wait::ld has completed (the values are in registers) and tcgen05.fence::before_thread_sync
// precedes the arrive.  A release arrive would additionally drain the thread's outstanding GLOBAL
// stores (the survivors just appended to the candidate lists) -- MEMBAR.ALL.GPU + ERRBAR in SASS,
// 22 % of all stall samples in the k = 1024 profile -- although nobody reads those lists before
// the kernel ends.
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// arrive on a barrier given by a shared::cluster address (own or peer CTA)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
    return ok != 0;
}
// try_wait with cluster-scope acquire: used where the arrivals come from the peer CTA.
__device__ __forceinline__ bool mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%1], %2, %3;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity), "r"(0x989680u)
        : "memory");
    return ok != 0;
}
// Blocking wait with a watchdog: a pipeline bug traps (CUDA error on the host) instead of hanging
// the GPU.  No legitimate wait in these kernels lasts anywhere near 2^32 cycles (> 2 s).
template <bool kClusterScope = false>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    auto probe = [&]() {
        return kClusterScope ? mbar_try_wait_cluster(bar, parity) : mbar_try_wait(bar, parity);
    };
    if (probe()) return;
    const long long t0 = clock64();
    while (!probe()) {
        if (clock64() - t0 > (1ll << 32)) {
            printf("flatip_b200: mbarrier wait timed out (block %d thread %d bar 0x%x parity %u)\n",
                   (int)blockIdx.x, (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

// ------------------------------------------------------------------ TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)map) : "memory");
}
// 2-D tile load into this CTA's shared memory, completion bytes to this CTA's `bar`.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst), "l"((uint64_t)map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// CTA-pair form: data lands in this CTA's shared memory, completion bytes go to `cluster_bar`, a
// shared::cluster address that may name the LEADER CTA's barrier (both CTAs of a pair feed one
// full-barrier, because one thread of the leader issues the MMAs for both).
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map,
                                                 uint32_t cluster_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst), "l"((uint64_t)map), "r"(cluster_bar), "r"(c0),
        "r"(c1)
        : "memory");
}

// ------------------------------------------------------------------ tcgen05
__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
template <int kCtas>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    if (kCtas == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                     "r"(ncols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    } else {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                     "r"(ncols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
}
template <int kCtas>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    if (kCtas == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                     : "memory");
    else
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                     : "memory");
}
// tcgen05.commit: arrive on `bar` once every MMA issued so far by this thread has completed.
// Pair form: the arrive is multicast to the barrier at the same offset in both CTAs of the pair.
template <int kCtas>
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    if (kCtas == 1) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
                     : "memory");
    } else {
        asm volatile(
            "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
            "[%0], %1;" ::"r"(bar), "h"((uint16_t)3)
            : "memory");
    }
}
// D[tmem] (+)= A[smem] * B[smem]^T.  kKind 0: kind::f16 (fp16 / bf16 operands, K = 16 per
// instruction), 1: kind::tf32 (fp32 operands read as tf32, K = 8).  fp32 accumulate.
template <int kCtas, int kKind>
__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                     uint32_t idesc, uint32_t accumulate) {
    if (kCtas == 1 && kKind == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
    else if (kCtas == 1 && kKind == 1)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
    else if (kCtas == 2 && kKind == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                     "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
                     "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
                     : "memory");
}
// 32 lanes x 32 columns of fp32 accumulators -> 32 registers per thread (lane = TMEM lane).
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

// Shared-memory matrix descriptor for a K-major operand tile laid out by TMA with SWIZZLE_128B:
// rows of 128 bytes, 8-row groups 1024 bytes apart (SBO); LBO is unused for swizzled K-major.
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);  // start address       bits [0,14)
    d |= (uint64_t)1 << 16;                        // leading byte offset bits [16,30) (ignored)
    d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset  bits [32,46)
    d |= (uint64_t)1 << 46;                        // descriptor version 1 (sm_100)
    d |= (uint64_t)2 << 61;                        // layout type SWIZZLE_128B
    return d;
}
// Instruction descriptor: fp32 accumulate, both operands K-major; fmt 0 = fp16, 1 = bf16, 2 = tf32.
__host__ __device__ constexpr uint32_t make_idesc(int fmt_a, int fmt_b, int m, int n) {
    return (1u << 4) | ((uint32_t)fmt_a << 7) | ((uint32_t)fmt_b << 10) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// Predicated 8-byte global store (no branch): the dense phase's writes.
__device__ __forceinline__ void st_global_v2_pred(uint64_t* ptr, uint32_t lo, uint32_t hi, bool p) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %3, 0;\n\t@p st.global.v2.u32 [%0], {%1, %2};\n\t}" ::"l"(ptr),
        "r"(lo), "r"(hi), "r"((uint32_t)p)
        : "memory");
}
// Epilogue pass 1, one score: if (v > thr) mask |= bit;   (two SASS instructions, no branch)
__device__ __forceinline__ void mask_or_if_gt(uint32_t& mask, float v, float thr, uint32_t bit) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.gt.f32 p, %1, %2;\n\t"
        "@p or.b32 %0, %0, %3;\n\t"
        "}"
        : "+r"(mask)
        : "f"(v), "f"(thr), "r"(bit));
}
// Epilogue pass 2, one score: if (take) { *list = {lo, hi}; ++list; }   (predicated, no branch)
__device__ __forceinline__ void list_append_if(uint64_t*& list, uint32_t lo, uint32_t hi, uint32_t take) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %3, 0;\n\t"
        "@p st.global.v2.u32 [%0], {%1, %2};\n\t"
        "@p add.u64 %0, %0, 8;\n\t"
        "}"
        : "+l"(list)
        : "r"(lo), "r"(hi), "r"(take)
        : "memory");
}

}  // namespace tc
}  // namespace fip
