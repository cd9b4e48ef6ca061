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
    const float* thr;          // [nq] running k-th best score per query (strict > passes)
    uint32_t* cnt;             // [nq] candidates currently in each list
    uint64_t* cand;            // [nq, cap] candidate keys
    uint32_t cap;
};

// Append one candidate (common tail of every scan kernel's epilogue).
__device__ __forceinline__ void append_candidate(const ScanArgs& a, int q, float score,
                                                 uint32_t row) {
    uint32_t slot = atomicAdd(a.cnt + q, 1u);
    if (slot < a.cap) a.cand[(size_t)q * a.cap + slot] = make_raw_key(score, row);
}

inline int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

}  // namespace fip
