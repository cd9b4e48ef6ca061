// K4: fused cross-shard exchange + merge over NVLink peer memory.
//
// After the local search every rank holds its shard's best k per query.  The NCCL path
// (sharded.py) all-gathers those lists and then launches the K3 merge.  This kernel does both in
// ONE launch per rank, tile by tile (tile = one query):
//   publish  CTA q stores its query's k (score, id) pairs straight into the exchange buffer of
//            EVERY rank (peer-mapped memory: plain st.global over NVLink 5 / NVSwitch), then
//            fence.sys and a release store of the epoch into that rank's flag[src rank][q];
//   wait     the same CTA acquires flag[r][q] == epoch for every source rank r in its OWN buffer
//            (local memory: the spin costs no NVLink traffic);
//   merge    g sorted lists -> global best k under (score desc, id asc), bitonic in shared memory.
// No collective, no global barrier: query q of rank A only waits for query q of the other ranks,
// so the transfer of later queries overlaps the merge of earlier ones.
//
// Liveness does not depend on how the hardware dispatches CTAs: the grid is bounded by the number
// of CTAs that can be resident at once and every CTA first PUBLISHES all of its queries (grid-stride
// loop, never blocks) and only then waits / merges them.  A wait therefore only ever depends on
// publish phases, which always complete -- also when a grid is larger than what is resident.
//
// Status: every rank also ships the two summary flag words of its local search (select.cuh
// flag_slot: list overflowed / query not certified) with query 0; the CTA that merges query 0 ORs
// the g statuses into *out_status, so every rank learns whether ANY shard needs a redo without a
// host round trip between the local search and the exchange.
//
// Buffers are double-buffered by epoch parity: a rank can be at most one search ahead of a peer
// (it cannot finish search e+1 before the peer has published e+1, which the peer only does after
// its own kernel e has completed), so parity (e+2)&1 == e&1 is free again by then.
//
// The reference has no counterpart (one process, faiss-cpu:
// /root/reference/src/pipeline/components/faiss_store.py:152); upstream faiss's IndexShards does
// the merge on the host.
#pragma once

#include "common.cuh"
#include "select.cuh"

namespace fip {

constexpr int kMaxRanks = 16;

struct ExchangeArgs {
    uint8_t* peer[kMaxRanks];   // base of every rank's exchange buffer as mapped HERE (peer[rank] is local)
    int g, rank;
    int nq, k;
    uint32_t epoch;             // > 0, increments per call on every rank in lock-step
    int64_t slot_bytes;         // capacity of one (parity, source rank) slot
    int64_t flags_off;          // byte offset of the flag area
    int nq_max;
    int off_i;                  // offset of the ids inside a slot for THIS call (pack layout of nq, k)
    const float* d_loc;         // [nq, k] this rank's local result
    const int64_t* i_loc;
    float* out_d;               // [nq, k] global result
    int64_t* out_i;
    int64_t status_off;         // byte offset of the status words: [parity][source rank] u32
    const uint32_t* flags_loc;  // this rank's two summary flag words (null = clean)
    uint32_t* out_status;       // OR-accumulated: bit 0 some shard overflowed, bit 1 uncertified (null = unused)
    int phase;                  // 0 = publish, wait, merge (the normal single launch); 1 = publish only;
                                // 2 = wait + merge only.  1 then 2, with every rank's publish launch
                                // finished in between, is for ranks that SHARE a device (tests, smoke):
                                // nothing then depends on two kernels being resident together.
};

__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// grid <= resident CTAs (any size is live, see above), block = a power of two >= 64; dynamic smem =
// next_pow2(g * k) * 8 bytes.
__global__ void exchange_merge_kernel(ExchangeArgs a) {
    extern __shared__ __align__(16) uint64_t s_keys[];
    const int tid = threadIdx.x;
    const uint32_t parity = a.epoch & 1u;
    const int64_t parity_off = (int64_t)parity * a.g * a.slot_bytes;

    // ---- publish: my rows q -> slot [parity][my rank] of every rank's buffer
    for (int q = blockIdx.x; q < a.nq && a.phase != 2; q += gridDim.x) {
        const float* src_d = a.d_loc + (size_t)q * a.k;
        const int64_t* src_i = a.i_loc + (size_t)q * a.k;
        for (int p = 0; p < a.g; ++p) {
            uint8_t* slot = a.peer[p] + parity_off + (int64_t)a.rank * a.slot_bytes;
            float* dst_d = reinterpret_cast<float*>(slot) + (size_t)q * a.k;
            int64_t* dst_i = reinterpret_cast<int64_t*>(slot + a.off_i) + (size_t)q * a.k;
            for (int i = tid; i < a.k; i += blockDim.x) {
                dst_d[i] = src_d[i];
                dst_i[i] = src_i[i];
            }
        }
        if (q == 0 && tid < a.g) {
            const uint32_t st = a.flags_loc ? ((a.flags_loc[0] ? 1u : 0u) | (a.flags_loc[1] ? 2u : 0u)) : 0u;
            reinterpret_cast<uint32_t*>(a.peer[tid] + a.status_off)[(size_t)parity * a.g + a.rank] = st;
        }
        __threadfence_system();
        __syncthreads();
        if (tid < a.g) {
            uint32_t* flags = reinterpret_cast<uint32_t*>(a.peer[tid] + a.flags_off);
            st_release_sys_u32(flags + ((size_t)parity * a.g + a.rank) * a.nq_max + q, a.epoch);
        }
    }

    if (a.phase == 1) return;
    uint8_t* mine = a.peer[a.rank];
    const int total = a.g * a.k;
    int n = 1;
    while (n < total) n <<= 1;
    const uint8_t* base = mine + parity_off;
    for (int q = blockIdx.x; q < a.nq; q += gridDim.x) {
        // ---- wait: every source rank's row q has landed in MY buffer
        if (tid < a.g) {
            const uint32_t* flag = reinterpret_cast<const uint32_t*>(mine + a.flags_off) +
                                   ((size_t)parity * a.g + tid) * a.nq_max + q;
            const long long t0 = clock64();
            while (ld_acquire_sys_u32(flag) != a.epoch) {
                if (clock64() - t0 > (60ll << 30)) {  // ~30 s: a peer died or never launched
                    printf("flatip_b200: exchange wait timed out (rank %d waits for rank %d, query %d, epoch %u)\n",
                           a.rank, tid, q, a.epoch);
                    __trap();
                }
                __nanosleep(64);
            }
            if (q == 0 && a.out_status) {
                const uint32_t st = __ldcg(reinterpret_cast<const uint32_t*>(mine + a.status_off) +
                                           (size_t)parity * a.g + tid);
                if (st) atomicOr(a.out_status, st);
            }
        }
        __syncthreads();

        // ---- merge (same keys as merge_topk_kernel: position in shard-major order breaks ties,
        //      which equals id order because shard r holds lower ids than shard r + 1)
        for (int i = tid; i < n; i += blockDim.x) {
            uint64_t key = 0ull;
            if (i < total) {
                const int r = i / a.k, p = i - r * a.k;
                const uint8_t* slot = base + (int64_t)r * a.slot_bytes;
                const size_t off = (size_t)q * a.k + p;
                // __ldcg: peers wrote these lines into this GPU's L2; never trust a stale L1 copy
                if (__ldcg(reinterpret_cast<const long long*>(slot + a.off_i) + off) >= 0)
                    key = make_key(__ldcg(reinterpret_cast<const float*>(slot) + off), (uint32_t)i);
            }
            s_keys[i] = key;
        }
        __syncthreads();
        bitonic_sort_desc(s_keys, n);
        for (int i = tid; i < a.k; i += blockDim.x) {
            const uint64_t key = i < n ? s_keys[i] : 0ull;
            float d = kLowestF32;
            int64_t id = -1;
            if (key != 0ull) {
                const int src = (int)key_row(key);
                const int r = src / a.k, p = src - r * a.k;
                const uint8_t* slot = base + (int64_t)r * a.slot_bytes;
                const size_t off = (size_t)q * a.k + p;
                d = __ldcg(reinterpret_cast<const float*>(slot) + off);
                id = (int64_t)__ldcg(reinterpret_cast<const long long*>(slot + a.off_i) + off);
            }
            a.out_d[(size_t)q * a.k + i] = d;
            a.out_i[(size_t)q * a.k + i] = id;
        }
        __syncthreads();   // s_keys is reused by the next query
    }
}

}  // namespace fip
