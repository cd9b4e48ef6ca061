// C ABI of the IVF-Flat (L2) operator (include/ivfflat_b200.h) and its host-side orchestration.
// Included at the end of flatip_api.cu (one translation unit: it reuses the select kernel, the
// candidate-list workspace conventions, make_tmap, DevBuf and the error plumbing).
//
// Reference anchors: the index the reference builds and loads is an IndexIVFFlat / METRIC_L2 with
// nlist = 4096 and nprobe = 64 (/root/reference/scripts/create_test_docs.py:83-104,
// /root/reference/src/pipeline/components/faiss_store.py:66-93); it is searched through the same
// `index.search(x, k)` call as the flat index (faiss_store.py:152).
//
// A search is: query norms -> coarse scores (all centroids) -> select nprobe -> [host: invert the
// (query, list) pairs into list-major tiles] -> dense scan of each query's first list(s) -> select
// (threshold + histogram) -> ONE filter scan of everything else with live thresholds -> select ->
// finalize (distances, ids).  Overflowing candidate lists (adversarial data) redo the batch in
// "safe mode" (phases too short to overflow), exactly like the flat path.
#pragma once

#include <chrono>
#include <cmath>
#include <cstring>
#include <numeric>

#include "../../include/ivfflat_b200.h"
#include "ivf.cuh"
#include "tc_ivf.cuh"
#include "ivf_plan.cuh"

struct fivf_index {
    int dim = 0, nlist = 0, device = 0, num_sms = 0;
    cudaStream_t stream = nullptr;
    std::mutex mu;
    bool trained = false;
    DevBuf<float> cent, cnorm;              // [nlist, dim], [nlist]
    // sorted-by-list storage (valid when !dirty)
    float* x = nullptr;                      // [n_sorted, dim]
    int64_t* ids = nullptr;                  // [n_sorted]
    float* xnorm = nullptr;                  // [n_sorted]
    int64_t n_sorted = 0;
    std::vector<int64_t> off;                // [nlist + 1] (host copy; lists are planned on the host)
    // rows added since the last rebuild, in add order
    float* pend_x = nullptr;
    int64_t pend_n = 0, pend_cap = 0;
    std::vector<int64_t> pend_ids;
    std::vector<int32_t> pend_assign;
    int64_t next_id = 0;
    CUtensorMap tmap_x{};
    bool tmap_valid = false;
    // options
    int force_safe = 0, live_thr = 1, qb_max = 16, profile = 0;
    int device_plan = 1;   // normal-mode searches are planned by the ivf_plan.cuh kernels (0: host planner, 1: batches of >= 1024 (query, probe) pairs, 2: always)
    int tc_min = 1;    // a list probed by >= tc_min queries of a batch goes to the tensor-core kernel (0 = never)
    CUtensorMap tmap_x128{};   // SWIZZLE_128B map of the sorted storage (tensor-core kernel)
    // fp16 FILTER COPY of the sorted storage (tensor-core kernel, kind::f16): built on first use,
    // dropped by every rebuild; the fp32 rows stay the truth (exact rescoring reads them)
    __half* x16 = nullptr;
    CUtensorMap tmap_x16{};
    bool x16_failed = false;       // the copy did not fit: stay on tf32
    float x2max_host = 0.f;        // max ||x||^2 (fp16 range check)
    int tensor_format = 0;         // 0: fp16 filter copy when representable, 1: tf32 on the fp32 rows
    // Coarse quantizer on the tensor cores with fp32-accurate scores: every fp32 value v is split
    // into three bf16 terms v = v1 + v2 + v3 (24 mantissa bits, exact), and <q, c> ~= q1c1 + q1c2 +
    // q2c1 + q1c3 + q2c2 + q3c1 (the dropped terms are below 2^-24 relative) becomes ONE bf16 GEMM
    // by laying the six products side by side along K: q' = [q1 q1 q2 q1 q2 q3], c' = [c1 c2 c1 c3 c2
    // c1] (K' = 6 dim).  `coarse` holds the c' rows as a bf16 flat index; K1's dense mode
    // (dense_scores, flatip_api.cu) writes the whole [nq, nlist] score matrix as candidate keys,
    // one small kernel turns <q,c> into 2<q,c> - ||c||^2, and the usual select picks the nprobe
    // best.  fp32 accumulation in the tensor core: the scores differ from the FFMA kernel's by
    // summation order only (~1e-7 relative), like any two fp32 GEMMs.
    fip_index* coarse = nullptr;
    bool coarse_valid = false;
    int coarse_tensor = 1;         // 0: FFMA coarse_scores_kernel (always used for assignment at add time), 1: tensor cores for batches >= 512, 2: always
    DevBuf<float> q_aug;
    int64_t st_coarse_tensor = 0;
    cudaEvent_t ev[8] = {};   // profile: start/stop pairs around the scan launches of up to 4 phases
    // workspace
    DevBuf<uint64_t> ckeys, cand;
    DevBuf<float> qnorm, probe_d, thr, sc, q32, out_d, stage, x2max, eps, ones;
    DevBuf<int64_t> probe_i, pos, out_i, perm, doff;
    DevBuf<int32_t> assign;
    DevBuf<ivf::TileQuery> tile_q;
    DevBuf<uint32_t> cnt, flags, hist, hbase;
    DevBuf<ivf::Tile> tiles;
    DevBuf<ivf::TcItem> tc_items;
    DevBuf<int32_t> qmap;
    DevBuf<uint32_t> qskip;
    DevBuf<float> qprep;
    DevBuf<int64_t> loff;                    // [nlist + 1] list boundaries on the device (device planner)
    bool loff_valid = false;
    DevBuf<uint32_t> plan_z, plan_take, plan_bases, plan_tile0base;   // device planner workspace
    PinnedBuf<ivf::PlanCounts> h_counts;
    PinnedBuf<float> h_qnorm;
    PinnedBuf<int64_t> h_probe, h_i;
    PinnedBuf<uint8_t> h_plan;
    PinnedBuf<uint32_t> h_flags;
    PinnedBuf<float> h_q, h_d;
    void* plan_ws = nullptr;   // IvfPlan: host plan workspace, reused by every search (build_plan)
    // stats of the last search
    int64_t st_launches = 0, st_phases = 0, st_tiles = 0, st_rows = 0, st_pairs = 0, st_overflow = 0,
            st_cap = 0, st_plan_us = 0, st_tiles16 = 0, st_occ16 = -1, st_distinct_rows = 0, st_scan_ns = 0, st_tc_items = 0, st_tc_pairs = 0, st_crowded = 0, st_crowded_final = 0, st_device_plan = 0, st_f16 = 0;
    int64_t ntotal() const { return n_sorted + pend_n; }
};

namespace {

int fivf_use(const fivf_index* h) {
    FIP_CUDA(cudaSetDevice(h->device));
    return FIP_OK;
}

constexpr float kF16NormLimit = 4.0e9f;   // ||v||^2 below this => every component below 65504 (fp16 max)
constexpr int kAssignChunk = 16384;   // rows per coarse-assignment chunk (keys: chunk * nlist * 8 B)

// Nearest centroid of n device rows -> host `out` (chunked: coarse scores + row argmax).
int fivf_assign_device(fivf_index* h, const float* x_dev, int64_t n, int32_t* out_host) {
    const int nlist = h->nlist, dim = h->dim;
    FIP_CUDA(h->ckeys.ensure((size_t)std::min<int64_t>(n, kAssignChunk) * nlist));
    FIP_CUDA(h->assign.ensure((size_t)std::min<int64_t>(n, kAssignChunk)));
    for (int64_t r0 = 0; r0 < n; r0 += kAssignChunk) {
        const int m = (int)std::min<int64_t>(kAssignChunk, n - r0);
        dim3 grid((nlist + ivf::CO_T - 1) / ivf::CO_T, (m + ivf::CO_T - 1) / ivf::CO_T);
        ivf::coarse_scores_kernel<<<grid, 256, 0, h->stream>>>(x_dev + (size_t)r0 * dim, m, h->cent.p, nlist,
                                                              h->cnorm.p, dim, h->ckeys.p, (uint32_t)nlist);
        FIP_CUDA(cudaGetLastError());
        ivf::row_argmax_kernel<<<(m + 7) / 8, 256, 0, h->stream>>>(h->ckeys.p, m, nlist, (uint32_t)nlist, h->assign.p);
        FIP_CUDA(cudaGetLastError());
        FIP_CUDA(cudaMemcpyAsync(out_host + r0, h->assign.p, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, h->stream));
        FIP_CUDA(cudaStreamSynchronize(h->stream));
    }
    return FIP_OK;
}

// v = v1 + v2 + v3 in bf16 (exact for fp32 v); row layouts along K' = 6 dim:
//   kQuery: [v1 v1 v2 v1 v2 v3]   (written as fp32 values that ARE bf16 numbers: the tensor-core
//                                   path's own rounding of the queries is then exact)
//   else  : [v1 v2 v1 v3 v2 v1]   (written as bf16 into the coarse index's storage)
template <bool kQuery, typename Out>
__global__ void split3_rows_kernel(const float* __restrict__ src, int n, int dim, Out* __restrict__ dst) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)n * dim) return;
    const int r = (int)(i / dim), c = (int)(i - (int64_t)r * dim);
    const float v = src[i];
    const float v1 = __bfloat162float(__float2bfloat16_rn(v));
    const float v2 = __bfloat162float(__float2bfloat16_rn(v - v1));
    const float v3 = __bfloat162float(__float2bfloat16_rn(v - v1 - v2));
    const float seg[6] = {v1, kQuery ? v1 : v2, kQuery ? v2 : v1, kQuery ? v1 : v3, v2, kQuery ? v3 : v1};
    Out* o = dst + (size_t)r * 6 * dim + c;
#pragma unroll
    for (int t = 0; t < 6; ++t) o[(size_t)t * dim] = to_storage<Out>(seg[t]);
}
// keys[q][c]: score <q,c>  ->  2 <q,c> - ||c||^2 (the coarse score; same raw-key layout)
__global__ void coarse_bias_kernel(uint64_t* __restrict__ keys, const float* __restrict__ cnorm, int nq, int nlist) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= (int64_t)nq * nlist) return;
    const int c = (int)(i % nlist);
    const uint64_t k = keys[i];
    const float s = fmaf(2.0f, __uint_as_float((uint32_t)(k >> 32)), -cnorm[c]);
    keys[i] = ((uint64_t)__float_as_uint(s) << 32) | (k & 0xFFFFFFFFull);
}

// (Re)builds the bf16 x 3 centroid table.  Enqueue only.
int fivf_build_coarse(fivf_index* h) {
    const int wide = 6 * h->dim;
    if (!h->coarse) {
        fip_config c{};
        c.dim = wide; c.dtype = FIP_BF16; c.device = h->device; c.capacity_rows = h->nlist;
        int rc = fip_create(&c, &h->coarse);
        if (rc) return rc;
        h->coarse->co.enabled = 0;
    }
    fip_index* ci = h->coarse;
    int rc = reserve_rows(ci, h->nlist);
    if (rc) return rc;
    const int64_t tot = (int64_t)h->nlist * h->dim;
    split3_rows_kernel<false, __nv_bfloat16><<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(
        h->cent.p, h->nlist, h->dim, (__nv_bfloat16*)ci->corpus);
    FIP_CUDA(cudaGetLastError());
    ci->n_rows = h->nlist;
    ci->tmap_x_valid = false; ci->corpus_gen++;
    h->coarse_valid = true;
    return FIP_OK;
}

int fivf_update_cnorm(fivf_index* h) {
    h->coarse_valid = false;
    FIP_CUDA(h->cnorm.ensure(h->nlist));
    ivf::row_norm2_kernel<<<(h->nlist + 7) / 8, 256, 0, h->stream>>>(h->cent.p, h->nlist, h->dim, h->cnorm.p);
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

// Stable counting sort of rows by list: off[nlist + 1], perm[pos] = source row.
void counting_sort(const int32_t* assign, int64_t n, int nlist, std::vector<int64_t>* off,
                   std::vector<int64_t>* perm) {
    off->assign((size_t)nlist + 1, 0);
    for (int64_t i = 0; i < n; ++i) (*off)[(size_t)assign[i] + 1]++;
    for (int l = 0; l < nlist; ++l) (*off)[(size_t)l + 1] += (*off)[l];
    perm->resize((size_t)n);
    std::vector<int64_t> cur(off->begin(), off->end() - 1);
    for (int64_t i = 0; i < n; ++i) (*perm)[(size_t)cur[assign[i]]++] = i;
}

// Merge the pending rows into the sorted-by-list storage (old rows of a list first, then the new
// ones in add order, so positions inside a list follow add order = id order for sequential ids).
int fivf_rebuild(fivf_index* h) {
    if (h->pend_n == 0) return FIP_OK;
    const int nlist = h->nlist, dim = h->dim;
    const int64_t n_old = h->n_sorted, n_new = n_old + h->pend_n;
    if (n_new >= (int64_t)0xFFFFFFFFll) return fail(FIP_ERR_INVALID, "an IVF index holds at most 2^32 - 2 rows per GPU");
    std::vector<int64_t> cnt_new((size_t)nlist, 0);
    for (int64_t i = 0; i < h->pend_n; ++i) cnt_new[h->pend_assign[(size_t)i]]++;
    std::vector<int64_t> off_new((size_t)nlist + 1, 0), perm((size_t)n_new);
    if (h->off.empty()) h->off.assign((size_t)nlist + 1, 0);
    for (int l = 0; l < nlist; ++l)
        off_new[(size_t)l + 1] = off_new[l] + (h->off[(size_t)l + 1] - h->off[l]) + cnt_new[l];
    std::vector<int64_t> cur((size_t)nlist);
    for (int l = 0; l < nlist; ++l) {   // virtual source = [old sorted rows ; pending rows]
        int64_t p = off_new[l];
        for (int64_t s = h->off[l]; s < h->off[(size_t)l + 1]; ++s) perm[(size_t)p++] = s;
        cur[l] = p;
    }
    for (int64_t i = 0; i < h->pend_n; ++i) perm[(size_t)cur[h->pend_assign[(size_t)i]]++] = n_old + i;
    // virtual source ids
    DevBuf<int64_t> src_ids;
    FIP_CUDA(src_ids.ensure((size_t)n_new));
    if (n_old) FIP_CUDA(cudaMemcpyAsync(src_ids.p, h->ids, sizeof(int64_t) * n_old, cudaMemcpyDeviceToDevice, h->stream));
    FIP_CUDA(cudaMemcpyAsync(src_ids.p + n_old, h->pend_ids.data(), sizeof(int64_t) * h->pend_n, cudaMemcpyHostToDevice, h->stream));
    FIP_CUDA(h->perm.ensure((size_t)n_new));
    FIP_CUDA(cudaMemcpyAsync(h->perm.p, perm.data(), sizeof(int64_t) * n_new, cudaMemcpyHostToDevice, h->stream));
    float *nx = nullptr, *nn = nullptr;
    int64_t* ni = nullptr;
    FIP_CUDA(cudaMalloc((void**)&nx, (size_t)n_new * dim * sizeof(float)));
    cudaError_t e = cudaMalloc((void**)&ni, (size_t)n_new * sizeof(int64_t));
    if (e == cudaSuccess) e = cudaMalloc((void**)&nn, (size_t)n_new * sizeof(float));
    if (e != cudaSuccess) { cudaFree(nx); cudaFree(ni); src_ids.release(); FIP_CUDA(e); }
    // gather in two launches (rows whose source is old storage / pending storage share one perm;
    // the kernel takes one source, so the pending rows are addressed through an offset pointer)
    ivf::gather_sorted2_kernel<<<(unsigned)((n_new + 7) / 8), 256, 0, h->stream>>>(
        h->x, h->pend_x, n_old, src_ids.p, h->perm.p, n_new, dim, nx, ni, nn);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    src_ids.release();
    if (e != cudaSuccess) { cudaFree(nx); cudaFree(ni); cudaFree(nn); FIP_CUDA(e); }
    cudaFree(h->x); cudaFree(h->ids); cudaFree(h->xnorm); cudaFree(h->pend_x);
    h->x = nx; h->ids = ni; h->xnorm = nn; h->n_sorted = n_new;
    h->pend_x = nullptr; h->pend_n = 0; h->pend_cap = 0;
    h->pend_ids.clear(); h->pend_ids.shrink_to_fit();
    h->pend_assign.clear(); h->pend_assign.shrink_to_fit();
    h->off.swap(off_new);
    h->tmap_valid = false;
    h->loff_valid = false;
    FIP_CUDA(h->x2max.ensure(1));
    FIP_CUDA(cudaMemsetAsync(h->x2max.p, 0, sizeof(float), h->stream));
    ivf::max_norm_kernel<<<h->num_sms * 4, 256, 0, h->stream>>>(h->xnorm, n_new, h->x2max.p);
    FIP_CUDA(cudaGetLastError());
    FIP_CUDA(cudaMemcpyAsync(&h->x2max_host, h->x2max.p, sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    if (h->x16) { cudaFree(h->x16); h->x16 = nullptr; }
    h->x16_failed = false;
    return FIP_OK;
}

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------
// Host plan: invert the probe table into list-major tiles, phase by phase.
struct Seg { int32_t list; uint32_t r0, r1; int32_t q; uint32_t skip; };
struct IvfPlan;
void fivf_free_plan(void* p);

// Lives in the index and is reused by every search: the vectors keep their capacity, so planning
// costs no allocations (fresh multi-megabyte vectors meant mmap + page faults on every call).
struct IvfPlan {
    struct Phase { size_t t16_begin = 0, t16_n = 0, t4_begin = 0, t4_n = 0; };
    std::vector<Phase> phases;
    std::vector<ivf::Tile> tiles;
    std::vector<ivf::TileQuery> tile_q;
    int64_t rows = 0, pairs = 0, distinct_rows = 0;   // distinct_rows: rows of the lists probed by >= 1 query
    // tensor-core path (phase 1 of the normal mode only): 128-query groups of heavily probed lists
    std::vector<ivf::TcItem> tc_items;
    std::vector<int32_t> qmap;     // prepared-query row -> query number (-1 = padding)
    std::vector<uint32_t> qskip;   // prepared-query row -> TileQuery::skip
    int64_t tc_pairs = 0;
    // scratch of build_plan
    std::vector<std::vector<Seg>> whole, part;   // per phase
    std::vector<int32_t> lcount, cur;
    std::vector<ivf::TileQuery> lq;
    std::vector<ivf::Tile> t16, t4;
    std::vector<uint8_t> seen;
    void reset() {
        phases.clear(); tiles.clear(); tile_q.clear(); tc_items.clear(); qmap.clear(); qskip.clear();
        rows = pairs = distinct_rows = tc_pairs = 0;
        for (auto& v : whole) v.clear();
        for (auto& v : part) v.clear();
        lq.clear(); t16.clear(); t4.clear();
    }
};


void fivf_free_plan(void* p) { delete static_cast<IvfPlan*>(p); }

// probes: [nq, nprobe] list numbers (-1 = none).
//   normal mode: phase 0 (dense) scans only the first `target0` rows of each query's probe
//     sequence -- enough to seed its threshold and score histogram; phase 1 scans EVERY probed list
//     as a whole, grouped across the queries that probe it, with live thresholds; a query skips the
//     rows of a list it already saw in phase 0 (TileQuery::skip), so nothing is scored twice and a
//     list is still read once per query group;
//   safe mode: the probe sequence is cut into phases of `cap` (dense) / `cap - k` rows so that no
//     candidate list can overflow; partial lists become single-query tiles.
void build_plan(const fivf_index* h, const int64_t* probes, int nq, int nprobe, int k, int cap,
                bool safe, int qb_big, int tc_min, IvfPlan* plan) {
    const int nlist = h->nlist;
    const std::vector<int64_t>& off = h->off;
    plan->reset();
    std::vector<std::vector<Seg>>& whole = plan->whole;
    std::vector<std::vector<Seg>>& part = plan->part;
    auto phase_vecs = [&](size_t p) {
        if (whole.size() <= p) { whole.resize(p + 1); part.resize(p + 1); }
    };
    const int64_t target0 = std::min<int64_t>(cap, std::max<int64_t>(4 * (int64_t)k, 256));
    if (!safe) {
        phase_vecs(1);
        whole[1].reserve((size_t)nq * nprobe);
        part[0].reserve((size_t)nq * 2);
        const size_t pairs_max = (size_t)nq * nprobe;
        plan->tile_q.reserve(pairs_max + (size_t)nq * 2 + 64);
        if (tc_min > 0) {
            const size_t rows_max = pairs_max + 32 * std::min<size_t>((size_t)nlist, pairs_max);
            plan->qmap.reserve(rows_max);
            plan->qskip.reserve(rows_max);
            plan->tc_items.reserve(pairs_max / 4 + 1024);
        }
        plan->tiles.reserve(pairs_max / 4 + (size_t)nq * 2 + 1024);
    }
    {
        std::vector<uint8_t>& seen = plan->seen;
        seen.assign((size_t)nlist, 0);
        for (size_t i = 0; i < (size_t)nq * nprobe; ++i) {
            const int64_t l = probes[i];
            if (l >= 0 && l < nlist && !seen[(size_t)l]) { seen[(size_t)l] = 1; plan->distinct_rows += off[(size_t)l + 1] - off[l]; }
        }
    }
    for (int q = 0; q < nq; ++q) {
        if (!safe) {
            int64_t cum = 0;
            for (int j = 0; j < nprobe; ++j) {
                const int64_t l = probes[(size_t)q * nprobe + j];
                if (l < 0 || l >= nlist) continue;
                const int64_t len = off[(size_t)l + 1] - off[l];
                if (len == 0) continue;
                int64_t take = 0;
                if (cum < target0) {
                    take = std::min(len, target0 - cum);
                    phase_vecs(0);
                    part[0].push_back(Seg{(int32_t)l, 0u, (uint32_t)take, q, 0u});
                    cum += take;
                }
                if (take < len) {
                    phase_vecs(1);
                    whole[1].push_back(Seg{(int32_t)l, 0u, (uint32_t)len, q, (uint32_t)take});
                }
            }
            continue;
        }
        size_t phase = 0;
        int64_t used = 0, budget = cap;
        for (int j = 0; j < nprobe; ++j) {
            const int64_t l = probes[(size_t)q * nprobe + j];
            if (l < 0 || l >= nlist) continue;
            const int64_t len = off[(size_t)l + 1] - off[l];
            int64_t r = 0;
            while (r < len) {
                const int64_t room = budget - used;
                if (room <= 0) { phase++; used = 0; budget = std::max<int64_t>(256, cap - k); continue; }
                const int64_t take = std::min(len - r, room);
                phase_vecs(phase);
                Seg s{(int32_t)l, (uint32_t)r, (uint32_t)(r + take), q, 0u};
                (r == 0 && take == len ? whole : part)[phase].push_back(s);
                r += take; used += take;
                if (phase == 0 && used >= target0) { phase++; used = 0; budget = std::max<int64_t>(256, cap - k); }
            }
        }
    }
    size_t n_phases = whole.size();
    while (n_phases > 0 && whole[n_phases - 1].empty() && part[n_phases - 1].empty() && n_phases > 2) n_phases--;
    plan->phases.assign(n_phases, IvfPlan::Phase{});
    std::vector<int32_t>& lcount = plan->lcount;
    lcount.assign((size_t)nlist + 1, 0);
    std::vector<ivf::TileQuery>& lq = plan->lq;
    std::vector<ivf::Tile>& t16 = plan->t16;
    std::vector<ivf::Tile>& t4 = plan->t4;
    t16.reserve(4096); t4.reserve(4096);
    auto emit = [&](std::vector<ivf::Tile>& dst, int64_t list0, int64_t r0, int64_t r1, uint32_t qoff, uint32_t gq) {
        for (int64_t r = r0; r < r1; r += ivf::SC_ROWS) {
            const uint32_t nr = (uint32_t)std::min<int64_t>(ivf::SC_ROWS, r1 - r);
            dst.push_back(ivf::Tile{(uint32_t)(list0 + r), nr, qoff, gq, (uint32_t)r});
            plan->rows += nr;
            plan->pairs += (int64_t)nr * gq;
        }
    };
    for (size_t p = 0; p < n_phases; ++p) {
        t16.clear(); t4.clear();
        // whole-list segments: bucket the queries by list (counting sort, O(pairs + nlist))
        std::fill(lcount.begin(), lcount.end(), 0);
        for (const Seg& s : whole[p]) lcount[(size_t)s.list + 1]++;
        for (int l = 0; l < nlist; ++l) lcount[(size_t)l + 1] += lcount[l];
        lq.resize(whole[p].size());
        {
            std::vector<int32_t>& cur = plan->cur;
            cur.assign(lcount.begin(), lcount.end() - 1);
            for (const Seg& s : whole[p]) lq[(size_t)cur[s.list]++] = ivf::TileQuery{s.q, s.skip};
        }
        for (int l = 0; l < nlist; ++l) {
            const int m = lcount[(size_t)l + 1] - lcount[l];
            if (m == 0) continue;
            const int64_t len = off[(size_t)l + 1] - off[l];
            if (p == 1 && !safe && tc_min > 0 && m >= tc_min) {
                // heavily probed list: its queries go to the tensor-core kernel in groups of up to 128
                // (a query that pre-scanned the head of the list in phase 0 carries its skip count)
                ivf::TileQuery* b = lq.data() + lcount[l];
                const int nfree = m;
                int taken = 0, ngroups = 0;
                uint32_t g_qrow0[64], g_qrows[64];   // a list has at most nq / 128 + 1 groups per round
                while (taken < nfree) {   // the last group may be small: its tile comes from L2 anyway
                    const int gq = std::min(128, nfree - taken);
                    const int padded = (gq + 31) / 32 * 32;
                    const size_t base = plan->qmap.size();
                    plan->qmap.resize(base + (size_t)padded, -1);
                    plan->qskip.resize(base + (size_t)padded, 0u);
                    for (int i = 0; i < gq; ++i) {
                        plan->qmap[base + (size_t)i] = b[taken + i].q;
                        plan->qskip[base + (size_t)i] = b[taken + i].skip;
                    }
                    plan->rows += len;
                    plan->pairs += len * gq;
                    plan->tc_pairs += len * gq;
                    taken += gq;
                    g_qrow0[ngroups] = (uint32_t)base;
                    g_qrows[ngroups] = (uint32_t)padded;
                    if (++ngroups == 64 || taken == nfree) {
                        // row-tile major: the groups of one 256-row tile are adjacent items, i.e. run
                        // on different SMs at the same time: the tile comes from HBM once, then from L2
                        for (int64_t r = 0; r < len; r += tc::BN)
                            for (int g = 0; g < ngroups; ++g)
                                plan->tc_items.push_back(ivf::TcItem{(uint32_t)(off[l] + r), (uint32_t)(off[l] + len),
                                                                     g_qrow0[g], (uint32_t)off[l], g_qrows[g]});
                        ngroups = 0;
                    }
                }
                if (taken) {   // the rest of the bucket stays on the FFMA2 path
                    for (int i = taken; i < m; ++i) b[i - taken] = b[i];
                }
                const int rest = m - taken;
                for (int g0 = 0; g0 < rest;) {
                    int gq = std::min(qb_big, rest - g0);
                    const bool big = gq > 4;
                    if (!big) gq = std::min(4, rest - g0);
                    const uint32_t qoff = (uint32_t)plan->tile_q.size();
                    for (int i = 0; i < gq; ++i) plan->tile_q.push_back(b[g0 + i]);
                    emit(big ? t16 : t4, off[l], 0, len, qoff, (uint32_t)gq);
                    g0 += gq;
                }
                continue;
            }
            for (int g0 = 0; g0 < m;) {
                int gq = std::min(qb_big, m - g0);
                const bool big = gq > 4;
                if (!big) gq = std::min(4, m - g0);
                const uint32_t qoff = (uint32_t)plan->tile_q.size();
                for (int i = 0; i < gq; ++i) plan->tile_q.push_back(lq[(size_t)lcount[l] + g0 + i]);
                emit(big ? t16 : t4, off[l], 0, len, qoff, (uint32_t)gq);
                g0 += gq;
            }
        }
        for (const Seg& s : part[p]) {
            const uint32_t qoff = (uint32_t)plan->tile_q.size();
            plan->tile_q.push_back(ivf::TileQuery{s.q, 0u});
            emit(t4, off[s.list], s.r0, s.r1, qoff, 1u);
        }
        IvfPlan::Phase& ph = plan->phases[p];
        ph.t16_begin = plan->tiles.size(); ph.t16_n = t16.size();
        plan->tiles.insert(plan->tiles.end(), t16.begin(), t16.end());
        ph.t4_begin = plan->tiles.size(); ph.t4_n = t4.size();
        plan->tiles.insert(plan->tiles.end(), t4.begin(), t4.end());
    }
}

template <int QB>
int launch_ivf_scan(fivf_index* h, const ivf::ScanIvfArgs& a, int* launches) {
    if (a.n_tiles == 0) return FIP_OK;
    auto kern = ivf::ivf_scan_kernel<QB>;
    const size_t smem = ivf::scan_smem_bytes(QB, h->dim);
    const int stages = ivf::scan_stages(QB, h->dim);
    if (stages < 2) return fail(FIP_ERR_INVALID, "dim too large for the IVF scan kernel's shared-memory ring");
    FIP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (QB == 16 && h->st_occ16 < 0) {
        int occ = 0;
        FIP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, ivf::SC_THREADS, smem));
        h->st_occ16 = occ;
    }
    const int grid = std::min(a.n_tiles, 2 * h->num_sms);   // two CTAs per SM
    kern<<<grid, ivf::SC_THREADS, smem, h->stream>>>(h->tmap_x, a, stages);
    FIP_CUDA(cudaGetLastError());
    (*launches)++;
    return FIP_OK;
}

int ivf_select(fivf_index* h, int nq, int k, int cap, bool final_pass, bool live, int keep_max, int* launches) {
    SelectArgs s{};
    if (!final_pass) {   // exactness window (see ivf_eps_kernel)
        s.win_eps = h->eps.p; s.win_scale = h->ones.p; s.keep_max = keep_max; s.uncert = h->flags.p + 1;
    }
    s.cand = h->cand.p; s.cnt = h->cnt.p; s.thr = h->thr.p; s.cap = (uint32_t)cap; s.k = k;
    s.final_pass = final_pass ? 1 : 0; s.out_d = h->sc.p; s.out_i = h->pos.p; s.id_base = 0;
    s.overflow = h->flags.p; s.dense_count = -1;
    if (live && !final_pass) { s.hist = h->hist.p; s.hbase = h->hbase.p; }
    const size_t smem = ((size_t)cap + (size_t)next_pow2(std::min(k, cap))) * sizeof(uint64_t);
    FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    select_topk_kernel<<<nq, kSelectThreads, smem, h->stream>>>(s);
    FIP_CUDA(cudaGetLastError());
    (*launches)++;
    return FIP_OK;
}

// One search of device-resident queries; results to device out_d / out_i.  Synchronises (the probe
// table is planned on the host).  *overflowed = some candidate list overflowed (redo in safe mode).
int fivf_run(fivf_index* h, const float* q_dev, int nq, int k, int nprobe, float* out_d, int64_t* out_i,
             bool safe, bool allow_tc, bool* overflowed, bool* crowded, bool wide_window = false) {
    const int dim = h->dim, nlist = h->nlist;
    int launches = 0;
    nprobe = std::max(1, std::min(nprobe, nlist));
    FIP_CUDA(h->qnorm.ensure(nq));
    FIP_CUDA(h->ckeys.ensure((size_t)nq * nlist));
    FIP_CUDA(h->probe_d.ensure((size_t)nq * nprobe));
    FIP_CUDA(h->probe_i.ensure((size_t)nq * nprobe));
    FIP_CUDA(h->flags.ensure(flag_words(nq)));
    FIP_CUDA(h->thr.ensure(nq));
    FIP_CUDA(h->cnt.ensure(nq));
    FIP_CUDA(h->sc.ensure((size_t)nq * k));
    FIP_CUDA(h->pos.ensure((size_t)nq * k));
    FIP_CUDA(h->h_probe.ensure((size_t)nq * nprobe));
    FIP_CUDA(h->eps.ensure(nq));
    FIP_CUDA(h->ones.ensure(nq));
    FIP_CUDA(h->x2max.ensure(1));
    if (h->n_sorted == 0) FIP_CUDA(cudaMemsetAsync(h->x2max.p, 0, sizeof(float), h->stream));
    ivf::row_norm2_kernel<<<(nq + 7) / 8, 256, 0, h->stream>>>(q_dev, nq, dim, h->qnorm.p);
    FIP_CUDA(cudaGetLastError());
    init_lists_kernel<<<(2 * nq + 2 + 255) / 256, 256, 0, h->stream>>>(h->thr.p, h->cnt.p, h->flags.p, nullptr, nq);
    FIP_CUDA(cudaGetLastError());
    // coarse quantizer: every centroid scored, the nprobe nearest selected (nearest first)
    h->st_coarse_tensor = 0;
    // (measured at the reference's index shape, 4096 x 768: 2.26 vs 2.36 ms per search at B = 1024,
    //  1.19 vs 1.17 ms at B = 32 -- the K' = 6 dim pass is latency-bound on few tiles -- hence auto = large batches)
    if (nlist >= 256 && (h->coarse_tensor == 2 || (h->coarse_tensor == 1 && nq >= 512))) {
        // ... on the tensor cores (bf16 x 3 split, see fivf_index::coarse): dense K1 pass, bias, select
        if (!h->coarse_valid) { int rc = fivf_build_coarse(h); if (rc) return rc; }
        const int wide = 6 * dim;
        FIP_CUDA(h->q_aug.ensure((size_t)nq * wide));
        const int64_t tot = (int64_t)nq * dim;
        split3_rows_kernel<true, float><<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(q_dev, nq, dim, h->q_aug.p);
        FIP_CUDA(cudaGetLastError());
        int rc = dense_scores(h->coarse, h->q_aug.p, nq, h->ckeys.p, (uint32_t)nlist, h->stream, &launches);
        if (rc) return rc;
        const int64_t nk = (int64_t)nq * nlist;
        coarse_bias_kernel<<<(unsigned)((nk + 255) / 256), 256, 0, h->stream>>>(h->ckeys.p, h->cnorm.p, nq, nlist);
        FIP_CUDA(cudaGetLastError());
        SelectArgs s{};
        s.cand = h->ckeys.p; s.cnt = h->cnt.p; s.thr = h->thr.p; s.cap = (uint32_t)nlist; s.k = nprobe;
        s.final_pass = 1; s.out_d = h->probe_d.p; s.out_i = h->probe_i.p; s.id_base = 0;
        s.overflow = h->flags.p; s.dense_count = nlist;
        const size_t smem = ((size_t)nlist + (size_t)next_pow2(nprobe)) * sizeof(uint64_t);
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel<<<nq, kSelectThreads, smem, h->stream>>>(s);
        FIP_CUDA(cudaGetLastError());
        launches += 3;
        h->st_coarse_tensor = 1;
    } else {
        dim3 grid((nlist + ivf::CO_T - 1) / ivf::CO_T, (nq + ivf::CO_T - 1) / ivf::CO_T);
        ivf::coarse_scores_kernel<<<grid, 256, 0, h->stream>>>(q_dev, nq, h->cent.p, nlist, h->cnorm.p, dim,
                                                              h->ckeys.p, (uint32_t)nlist);
        FIP_CUDA(cudaGetLastError());
        SelectArgs s{};
        s.cand = h->ckeys.p; s.cnt = h->cnt.p; s.thr = h->thr.p; s.cap = (uint32_t)nlist; s.k = nprobe;
        s.final_pass = 1; s.out_d = h->probe_d.p; s.out_i = h->probe_i.p; s.id_base = 0;
        s.overflow = h->flags.p; s.dense_count = nlist;
        const size_t smem = ((size_t)nlist + (size_t)next_pow2(nprobe)) * sizeof(uint64_t);
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel<<<nq, kSelectThreads, smem, h->stream>>>(s);
        FIP_CUDA(cudaGetLastError());
        launches += 4;
    }
    const int cap = safe ? kMaxCap : std::min(std::max(4096, 32 * next_pow2(k)), kMaxCap);
    const bool live = h->live_thr && !safe;
    // truncation limit of the exactness window; wide_window (last resort for clusters of duplicates
    // or near-ties larger than 2k + 64 rows): half the list capacity
    const int keep_max = wide_window ? cap / 2 : std::min(cap / 2, 2 * k + 64);
    const bool big_ok = h->qb_max >= 16 && ivf::scan_stages(16, dim) >= 2;
    const bool tc_ok = allow_tc && h->tc_min > 0 && dim % 32 == 0;
    const int k_plan = std::max(k, keep_max);
    if (!h->plan_ws) h->plan_ws = new IvfPlan();
    IvfPlan& plan = *static_cast<IvfPlan*>(h->plan_ws);
    bool q_fits_f16 = false;   // known only on the device-plan path (PlanCounts::q_over)
    std::vector<IvfPlan::Phase> dev_phases;
    const std::vector<IvfPlan::Phase>* phases = &plan.phases;
    size_t n_tc_items = 0, n_qrows = 0, n_tiles = 0;
    int64_t pl_rows = 0, pl_pairs = 0, pl_distinct = 0, pl_tc_pairs = 0;
    // (a handful of pairs is planned faster on the host than five more launches take)
    bool dev_plan = h->device_plan && !safe && (int64_t)nq * nprobe < (int64_t)1 << 30 &&
                    (h->device_plan > 1 || (int64_t)nq * nprobe >= 1024);
    if (h->profile) FIP_CUDA(cudaStreamSynchronize(h->stream));   // "plan_us" then excludes the coarse quantizer
    const auto t0 = std::chrono::steady_clock::now();
    if (dev_plan) {
        // ---- plan on the device (ivf_plan.cuh); the host only reads the work-list sizes back
        const size_t pairs_max = (size_t)nq * nprobe;
        const size_t zwords = sizeof(ivf::PlanCounts) / 4 + 3 * (size_t)nlist;
        FIP_CUDA(h->plan_z.ensure(zwords));
        FIP_CUDA(h->plan_take.ensure(pairs_max));
        FIP_CUDA(h->plan_bases.ensure((size_t)ivf::PLAN_BASES * ((size_t)nlist + 1)));
        FIP_CUDA(h->plan_tile0base.ensure(nq));
        FIP_CUDA(h->h_counts.ensure(1));
        FIP_CUDA(h->tile_q.ensure(pairs_max + (size_t)nq + 1));
        if (!h->loff_valid) {
            if (h->off.empty()) h->off.assign((size_t)nlist + 1, 0);
            FIP_CUDA(h->loff.ensure((size_t)nlist + 1));
            FIP_CUDA(cudaMemcpyAsync(h->loff.p, h->off.data(), sizeof(int64_t) * ((size_t)nlist + 1),
                                     cudaMemcpyHostToDevice, h->stream));
            FIP_CUDA(cudaStreamSynchronize(h->stream));   // h->off is pageable
            h->loff_valid = true;
        }
        ivf::PlanCounts* d_counts = reinterpret_cast<ivf::PlanCounts*>(h->plan_z.p);
        uint32_t* lcount = h->plan_z.p + sizeof(ivf::PlanCounts) / 4;
        uint32_t* cur = lcount + nlist;
        uint32_t* probed = cur + nlist;
        const uint32_t target0 = (uint32_t)std::min<int64_t>(cap, std::max<int64_t>(4 * (int64_t)k_plan, 256));
        const int tc_min = tc_ok ? h->tc_min : 0, qb_big = big_ok ? 16 : 4;
        FIP_CUDA(cudaMemsetAsync(h->plan_z.p, 0, zwords * 4, h->stream));
        ivf::plan_query_kernel<<<(nq + 3) / 4, 128, 0, h->stream>>>(   // one warp per query
            h->probe_i.p, nq, nprobe, nlist, h->loff.p, target0, h->plan_take.p, lcount, probed,
            h->plan_tile0base.p, h->qnorm.p, kF16NormLimit, d_counts);
        FIP_CUDA(cudaGetLastError());
        ivf::plan_lists_kernel<<<1, ivf::PLAN_LIST_THREADS, 0, h->stream>>>(
            lcount, probed, h->loff.p, nlist, tc_min, qb_big, h->plan_bases.p, d_counts);
        FIP_CUDA(cudaGetLastError());
        FIP_CUDA(cudaMemcpyAsync(h->h_counts.p, d_counts, sizeof(ivf::PlanCounts), cudaMemcpyDeviceToHost, h->stream));
        FIP_CUDA(cudaStreamSynchronize(h->stream));
        const ivf::PlanCounts c = *h->h_counts.p;
        if (c.tiles_total + c.n_tiles0 >= (1ull << 30)) {
            dev_plan = false;   // work lists beyond 32-bit indexing: let the host planner (64-bit) decide
        } else {
            n_tiles = (size_t)c.n_t16 + c.n_t4 + c.n_tiles0;
            n_tc_items = c.n_items;
            n_qrows = c.qrows;
            FIP_CUDA(h->tiles.ensure(n_tiles + 1));
            FIP_CUDA(h->tc_items.ensure(n_tc_items + 1));
            FIP_CUDA(h->qmap.ensure(n_qrows + 1));
            FIP_CUDA(h->qskip.ensure(n_qrows + 1));
            ivf::Tile* t16 = h->tiles.p;
            ivf::Tile* t4 = t16 + c.n_t16;
            ivf::Tile* t0 = t4 + c.n_t4;
            if (c.n_pairs1) {
                ivf::plan_scatter_kernel<<<(unsigned)((pairs_max + 255) / 256), 256, 0, h->stream>>>(
                    h->probe_i.p, h->plan_take.p, (int64_t)pairs_max, nprobe, h->loff.p, h->plan_bases.p, cur, h->tile_q.p);
                FIP_CUDA(cudaGetLastError());
                ivf::plan_emit_kernel<<<nlist, 128, 0, h->stream>>>(h->plan_bases.p, h->loff.p, nlist, tc_min, qb_big,
                                                                   h->tile_q.p, t16, t4, h->tc_items.p, h->qmap.p, h->qskip.p);
                FIP_CUDA(cudaGetLastError());
                launches += 2;
            }
            ivf::plan_tiles0_kernel<<<(nq + 127) / 128, 128, 0, h->stream>>>(
                h->probe_i.p, h->plan_take.p, nq, nprobe, h->loff.p, h->plan_tile0base.p, (uint32_t)pairs_max, t0, h->tile_q.p);
            FIP_CUDA(cudaGetLastError());
            launches += 3;
            dev_phases.assign(2, IvfPlan::Phase{});
            dev_phases[0].t16_begin = 0; dev_phases[0].t16_n = 0;
            dev_phases[0].t4_begin = (size_t)c.n_t16 + c.n_t4; dev_phases[0].t4_n = c.n_tiles0;
            dev_phases[1].t16_begin = 0; dev_phases[1].t16_n = c.n_t16;
            dev_phases[1].t4_begin = c.n_t16; dev_phases[1].t4_n = c.n_t4;
            phases = &dev_phases;
            pl_rows = (int64_t)c.rows; pl_pairs = (int64_t)c.pairs; pl_distinct = (int64_t)c.distinct_rows;
            pl_tc_pairs = (int64_t)c.tc_pairs;
            q_fits_f16 = c.q_over == 0;
        }
    }
    if (!dev_plan) {
        FIP_CUDA(cudaMemcpyAsync(h->h_probe.p, h->probe_i.p, sizeof(int64_t) * (size_t)nq * nprobe,
                                 cudaMemcpyDeviceToHost, h->stream));
        FIP_CUDA(h->h_qnorm.ensure(nq));
        FIP_CUDA(cudaMemcpyAsync(h->h_qnorm.p, h->qnorm.p, sizeof(float) * (size_t)nq, cudaMemcpyDeviceToHost, h->stream));
        FIP_CUDA(cudaStreamSynchronize(h->stream));
        q_fits_f16 = true;
        for (int i = 0; i < nq; ++i) q_fits_f16 = q_fits_f16 && h->h_qnorm.p[i] < kF16NormLimit;
        build_plan(h, h->h_probe.p, nq, nprobe, k_plan, cap, safe, big_ok ? 16 : 4, tc_ok ? h->tc_min : 0, &plan);
        const size_t tile_bytes = plan.tiles.size() * sizeof(ivf::Tile), tq_bytes = plan.tile_q.size() * sizeof(ivf::TileQuery);
        FIP_CUDA(h->h_plan.ensure(tile_bytes + tq_bytes + 16));
        FIP_CUDA(h->tiles.ensure(plan.tiles.size() + 1));
        FIP_CUDA(h->tile_q.ensure(plan.tile_q.size() + 1));
        if (tile_bytes) {
            memcpy(h->h_plan.p, plan.tiles.data(), tile_bytes);
            FIP_CUDA(cudaMemcpyAsync(h->tiles.p, h->h_plan.p, tile_bytes, cudaMemcpyHostToDevice, h->stream));
        }
        if (tq_bytes) {
            memcpy(h->h_plan.p + tile_bytes, plan.tile_q.data(), tq_bytes);
            FIP_CUDA(cudaMemcpyAsync(h->tile_q.p, h->h_plan.p + tile_bytes, tq_bytes, cudaMemcpyHostToDevice, h->stream));
        }
        n_tc_items = plan.tc_items.size();
        n_qrows = plan.qmap.size();
        n_tiles = plan.tiles.size();
        if (n_tc_items) {
            FIP_CUDA(h->qmap.ensure(n_qrows));
            FIP_CUDA(h->tc_items.ensure(n_tc_items));
            FIP_CUDA(h->qskip.ensure(n_qrows));
            FIP_CUDA(cudaMemcpyAsync(h->qmap.p, plan.qmap.data(), n_qrows * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
            FIP_CUDA(cudaMemcpyAsync(h->qskip.p, plan.qskip.data(), n_qrows * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
            FIP_CUDA(cudaMemcpyAsync(h->tc_items.p, plan.tc_items.data(), n_tc_items * sizeof(ivf::TcItem),
                                     cudaMemcpyHostToDevice, h->stream));
        }
        pl_rows = plan.rows; pl_pairs = plan.pairs; pl_distinct = plan.distinct_rows; pl_tc_pairs = plan.tc_pairs;
    }
    h->st_plan_us = std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count();
    const bool use_tc = n_tc_items > 0;
    // tensor-core operand format: the fp16 filter copy when every component of the lists and of the
    // batch's queries is inside fp16's range (norm test), else tf32 straight from the fp32 rows
    bool use_f16 = use_tc && h->tensor_format == 0 && dim % 64 == 0 && q_fits_f16 &&
                   h->x2max_host < kF16NormLimit && !h->x16_failed;
    if (use_f16 && !h->x16) {
        const size_t elems = (size_t)h->n_sorted * dim;
        if (cudaMalloc((void**)&h->x16, elems * sizeof(__half)) != cudaSuccess) {
            cudaGetLastError();
            h->x16 = nullptr; h->x16_failed = true; use_f16 = false;
        } else {
            ivf::rows_to_f16_kernel<<<h->num_sms * 8, 256, 0, h->stream>>>(h->x, h->x16, (int64_t)(elems / 8));
            FIP_CUDA(cudaGetLastError());
            int rc16 = make_tmap(&h->tmap_x16, h->x16, kF16, dim, (uint64_t)h->n_sorted);
            if (rc16) return rc16;
        }
    }
    ivf::ivf_eps_kernel<<<(nq + 255) / 256, 256, 0, h->stream>>>(h->qnorm.p, h->x2max.p, nq, dim,
                                                               use_tc ? (use_f16 ? 2 : 1) : 0, h->eps.p, h->ones.p);
    FIP_CUDA(cudaGetLastError());
    CUtensorMap tmap_qg{};
    if (use_tc) {   // prepared queries of the tensor-core groups + the item list
        const size_t rows = n_qrows;
        FIP_CUDA(h->qprep.ensure(rows * (size_t)dim));
        const int64_t vecs = (int64_t)rows * (dim / 4);
        if (use_f16)
            ivf::gather_queries_kernel<__half><<<(unsigned)((vecs + 255) / 256), 256, 0, h->stream>>>(
                q_dev, h->qmap.p, (int64_t)rows, dim, reinterpret_cast<__half*>(h->qprep.p));
        else
            ivf::gather_queries_kernel<float><<<(unsigned)((vecs + 255) / 256), 256, 0, h->stream>>>(
                q_dev, h->qmap.p, (int64_t)rows, dim, h->qprep.p);
        FIP_CUDA(cudaGetLastError());
        launches++;
        {   // [rows, dim], box = 128 bytes of K x 32 rows, SWIZZLE_128B
            EncodeTiledFn enc = get_encode_fn();
            if (!enc) return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
            const int esize = use_f16 ? 2 : 4;
            cuuint64_t dims[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
            cuuint64_t strides[1] = {(cuuint64_t)dim * esize};
            cuuint32_t box[2] = {(cuuint32_t)(tc::BK_BYTES / esize), 32u};
            cuuint32_t estr[2] = {1, 1};
            CUresult r = enc(&tmap_qg, use_f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                             h->qprep.p, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS)
                return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
        }
    }
    FIP_CUDA(h->cand.ensure((size_t)nq * cap));
    if (live) {
        FIP_CUDA(h->hist.ensure((size_t)nq * kHistBins));
        FIP_CUDA(h->hbase.ensure(nq));
    }
    if (!h->tmap_valid && h->n_sorted > 0) {
        // [n_sorted, dim] fp32, box = 16 floats (64 bytes) x 128 rows, SWIZZLE_64B
        EncodeTiledFn enc = get_encode_fn();
        if (!enc) return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
        cuuint64_t dims[2] = {(cuuint64_t)dim, (cuuint64_t)h->n_sorted};
        cuuint64_t strides[1] = {(cuuint64_t)dim * 4};
        cuuint32_t box[2] = {(cuuint32_t)ivf::SC_CHUNK, (cuuint32_t)ivf::SC_BOX_ROWS};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = enc(&h->tmap_x, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, h->x, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
            return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
        int rc128 = make_tmap(&h->tmap_x128, h->x, kF32, dim, (uint64_t)h->n_sorted);
        if (rc128) return rc128;
        h->tmap_valid = true;
    }
    // the coarse select used cnt / thr: reset the lists
    init_lists_kernel<<<(2 * nq + 2 + 255) / 256, 256, 0, h->stream>>>(h->thr.p, h->cnt.p, h->flags.p, nullptr, nq);
    FIP_CUDA(cudaGetLastError());
    launches++;

    ivf::ScanIvfArgs a{};
    a.tile_q = h->tile_q.p; a.q32 = q_dev; a.xnorm = h->xnorm; a.dim = dim;
    a.lists.nq = nq; a.lists.thr = h->thr.p; a.lists.cnt = h->cnt.p; a.lists.cand = h->cand.p;
    a.lists.cap = (uint32_t)cap; a.lists.k_thr = k;
    if (live) { a.lists.hist = h->hist.p; a.lists.hbase = h->hbase.p; a.lists.win_eps = h->eps.p; a.lists.win_scale = h->ones.p; }
    const size_t n_phases = phases->size();
    for (size_t p = 0; p < n_phases; ++p) {
        const IvfPlan::Phase& ph = (*phases)[p];
        a.dense = p == 0 ? 1 : 0;
        const bool timed = h->profile && p < 4;
        if (timed) {
            if (!h->ev[0]) for (auto& e : h->ev) FIP_CUDA(cudaEventCreate(&e));
            FIP_CUDA(cudaEventRecord(h->ev[2 * p], h->stream));
        }
        if (p == 1 && use_tc) {   // heavily probed lists: tcgen05 tiles
            ivf::TcIvfArgs ta{};
            ta.items = h->tc_items.p; ta.n_items = (int)n_tc_items; ta.qmap = h->qmap.p; ta.qskip = h->qskip.p;
            ta.xnorm = h->xnorm; ta.dim = dim; ta.lists = a.lists;
            constexpr int smem = tc::Cfg<2, 1, 1>::SMEM_BYTES;
            static_assert(tc::Cfg<0, 1, 1>::SMEM_BYTES == smem, "same ring for both formats");
            const int grid = std::min(ta.n_items, h->num_sms);
            if (use_f16) {
                FIP_CUDA(cudaFuncSetAttribute(ivf::tc_ivf_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                ivf::tc_ivf_kernel<0><<<grid, ivf::TCI_THREADS, smem, h->stream>>>(tmap_qg, h->tmap_x16, ta);
            } else {
                FIP_CUDA(cudaFuncSetAttribute(ivf::tc_ivf_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                ivf::tc_ivf_kernel<2><<<grid, ivf::TCI_THREADS, smem, h->stream>>>(tmap_qg, h->tmap_x128, ta);
            }
            FIP_CUDA(cudaGetLastError());
            launches++;
        }
        a.tiles = h->tiles.p + ph.t16_begin; a.n_tiles = (int)ph.t16_n;
        int rc = launch_ivf_scan<16>(h, a, &launches);
        if (rc) return rc;
        a.tiles = h->tiles.p + ph.t4_begin; a.n_tiles = (int)ph.t4_n;
        rc = launch_ivf_scan<4>(h, a, &launches);
        if (rc) return rc;
        if (timed) FIP_CUDA(cudaEventRecord(h->ev[2 * p + 1], h->stream));
        rc = ivf_select(h, nq, k, cap, false, live, keep_max, &launches);
        if (rc) return rc;
    }
    if (n_phases > 0) {   // exact distances for the survivors, then the final ranking
        const int64_t warps = (int64_t)nq * keep_max;
        const unsigned grid = (unsigned)std::min<int64_t>((warps + 7) / 8, (int64_t)h->num_sms * 8);
        ivf::ivf_rescore_kernel<<<grid, 256, 0, h->stream>>>(h->cand.p, h->cnt.p, (uint32_t)cap, nq, keep_max, h->x, q_dev, dim);
        FIP_CUDA(cudaGetLastError());
        launches++;
    }
    {   // (nothing scanned -- empty index or empty probed lists -- gives all padding)
        int rc = ivf_select(h, nq, k, cap, true, false, 0, &launches);
        if (rc) return rc;
    }
    const int64_t tot = (int64_t)nq * k;
    ivf::ivf_finalize_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, h->stream>>>(
        h->sc.p, h->pos.p, h->ids, nq, k, out_d, out_i);
    FIP_CUDA(cudaGetLastError());
    launches++;
    FIP_CUDA(h->h_flags.ensure(4));
    FIP_CUDA(cudaMemcpyAsync(h->h_flags.p, h->flags.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    *overflowed = h->h_flags.p[0] != 0;
    // some query's exactness window held more than keep_max candidates and was truncated: the
    // caller redoes the batch on the (much narrower) fp32 window, then with a wide window
    *crowded = h->h_flags.p[1] != 0;
    h->st_launches = launches; h->st_phases = (int64_t)n_phases; h->st_tiles = (int64_t)n_tiles;
    h->st_rows = pl_rows; h->st_pairs = pl_pairs; h->st_cap = cap;
    h->st_distinct_rows = pl_distinct;
    h->st_tc_items = (int64_t)n_tc_items; h->st_tc_pairs = pl_tc_pairs;
    h->st_device_plan = dev_plan ? 1 : 0;
    h->st_f16 = use_f16 ? 1 : 0;
    h->st_scan_ns = 0;
    if (h->profile)
        for (size_t p = 0; p < n_phases && p < 4; ++p) {
            float ms = 0.f;
            FIP_CUDA(cudaEventElapsedTime(&ms, h->ev[2 * p], h->ev[2 * p + 1]));
            h->st_scan_ns += (int64_t)(ms * 1e6);
        }
    h->st_tiles16 = 0;
    for (const auto& ph : *phases) h->st_tiles16 += (int64_t)ph.t16_n;
    return FIP_OK;
}

int fivf_search_checked(fivf_index* h, const float* q_dev, int nq, int k, int nprobe, float* out_d,
                        int64_t* out_i) {
    int rc = fivf_rebuild(h);
    if (rc) return rc;
    bool over = false, crowded = false;
    rc = fivf_run(h, q_dev, nq, k, nprobe, out_d, out_i, h->force_safe != 0, true, &over, &crowded);
    if (rc) return rc;
    h->st_overflow = over ? 1 : 0;
    h->st_crowded = crowded ? 1 : 0;
    if (crowded && !over) {   // near-tie clusters tighter than tf32 resolves: FFMA2 (fp32) filter only
        rc = fivf_run(h, q_dev, nq, k, nprobe, out_d, out_i, h->force_safe != 0, false, &over, &crowded);
        if (rc) return rc;
    }
    if (over) {
        rc = fivf_run(h, q_dev, nq, k, nprobe, out_d, out_i, true, false, &over, &crowded);
        if (rc) return rc;
        if (over) return fail(FIP_ERR_STATE, "internal error: IVF safe mode overflowed a candidate list");
    }
    if (crowded) {
        // more than 2k + 64 candidates inside the fp32 rounding window (large clusters of duplicate
        // or nearly duplicate rows): safe mode with the window limit at half the list capacity
        rc = fivf_run(h, q_dev, nq, k, nprobe, out_d, out_i, true, false, &over, &crowded, true);
        if (rc) return rc;
        if (over) return fail(FIP_ERR_STATE, "internal error: IVF safe mode overflowed a candidate list");
    }
    // still crowded: over 8192 rows tie within fp32 rounding of one another for some query; the
    // result is then exact only up to that rounding (reported, never silent)
    h->st_crowded_final = crowded ? 1 : 0;
    return FIP_OK;
}

int fivf_check_search(const fivf_index* h, const void* q, int64_t nq, int32_t k, const void* d, const void* i) {
    if (!h) return fail(FIP_ERR_INVALID, "null index");
    if (nq < 0 || (nq > 0 && !q)) return fail(FIP_ERR_INVALID, "bad query buffer");
    if (k < 1 || k > kMaxK) return fail(FIP_ERR_INVALID, "k must be in [1, 2048]");
    if (nq > 0 && (!d || !i)) return fail(FIP_ERR_INVALID, "null output buffer");
    if (nq > (int64_t)1 << 20) return fail(FIP_ERR_INVALID, "at most 2^20 queries per call");
    if (!h->trained) return fail(FIP_ERR_STATE, "IVF index is not trained (no centroids)");
    return FIP_OK;
}

}  // namespace

extern "C" {

int fivf_create(int32_t dim, int32_t nlist, int32_t device, fivf_index** out) {
    if (!out) return fail(FIP_ERR_INVALID, "null output handle");
    *out = nullptr;
    if (dim < 32 || dim % 32 != 0 || dim > 4096)
        return fail(FIP_ERR_INVALID, "IVF-Flat needs 32 <= dim <= 4096 and dim % 32 == 0 (128-byte TMA chunks)");
    if (nlist < 1 || nlist > kMaxCap) return fail(FIP_ERR_INVALID, "nlist must be in [1, 16384]");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(FIP_ERR_NODEVICE, "no CUDA device visible (this library has no CPU fallback)");
    }
    if (device < 0 || device >= count) return fail(FIP_ERR_INVALID, "device index out of range");
    cudaDeviceProp prop{};
    FIP_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(FIP_ERR_NODEVICE, std::string("device is ") + prop.name + ", this library is built for sm_100a only");
    fivf_index* h = new fivf_index();
    h->dim = dim; h->nlist = nlist; h->device = device; h->num_sms = prop.multiProcessorCount;
    h->off.assign((size_t)nlist + 1, 0);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = h->cent.ensure((size_t)nlist * dim);
    if (e == cudaSuccess) e = h->cnorm.ensure(nlist);
    if (e != cudaSuccess) { delete h; FIP_CUDA(e); }
    *out = h;
    return FIP_OK;
}

void fivf_destroy(fivf_index* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->x); cudaFree(h->ids); cudaFree(h->xnorm); cudaFree(h->pend_x); cudaFree(h->x16);
    if (h->coarse) fip_destroy(h->coarse);
    h->q_aug.release();
    h->cent.release(); h->cnorm.release(); h->ckeys.release(); h->cand.release(); h->qnorm.release();
    h->probe_d.release(); h->thr.release(); h->sc.release(); h->q32.release(); h->out_d.release();
    h->stage.release(); h->x2max.release(); h->eps.release(); h->ones.release(); h->probe_i.release(); h->pos.release(); h->out_i.release(); h->perm.release();
    h->loff.release(); h->plan_z.release(); h->plan_take.release(); h->plan_bases.release(); h->plan_tile0base.release();
    h->doff.release(); h->assign.release(); h->tile_q.release(); h->cnt.release(); h->flags.release();
    h->hist.release(); h->hbase.release(); h->tiles.release(); h->tc_items.release(); h->qmap.release(); h->qskip.release(); h->qprep.release(); h->h_probe.release(); h->h_i.release();
    h->h_plan.release(); h->h_flags.release(); h->h_q.release(); h->h_d.release();
    fivf_free_plan(h->plan_ws);
    for (auto& e : h->ev) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int64_t fivf_ntotal(const fivf_index* h) { return h ? h->ntotal() : -1; }
int32_t fivf_nlist(const fivf_index* h) { return h ? h->nlist : -1; }
int32_t fivf_dim(const fivf_index* h) { return h ? h->dim : -1; }
int32_t fivf_is_trained(const fivf_index* h) { return h && h->trained ? 1 : 0; }

int fivf_set_centroids(fivf_index* h, const float* centroids) {
    if (!h || !centroids) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    if (h->ntotal() > 0) return fail(FIP_ERR_STATE, "centroids cannot change once vectors were added");
    int rc = fivf_use(h);
    if (rc) return rc;
    FIP_CUDA(cudaMemcpyAsync(h->cent.p, centroids, sizeof(float) * (size_t)h->nlist * h->dim, cudaMemcpyHostToDevice, h->stream));
    rc = fivf_update_cnorm(h);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    h->trained = true;
    return FIP_OK;
}

int fivf_get_centroids(fivf_index* h, float* out) {
    if (!h || !out) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->trained) return fail(FIP_ERR_STATE, "IVF index is not trained (no centroids)");
    int rc = fivf_use(h);
    if (rc) return rc;
    FIP_CUDA(cudaMemcpyAsync(out, h->cent.p, sizeof(float) * (size_t)h->nlist * h->dim, cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    return FIP_OK;
}

// k-means for the coarse quantizer (faiss Clustering::train as IndexIVF::train drives it: at most
// 256 training points per centroid, random distinct rows as initial centroids, `niter` Lloyd
// iterations, empty clusters re-seeded by splitting a large one with a +-1/1024 perturbation).
// The random choices differ from faiss's (its RNG stream is not reproduced): trained centroids are
// a property of this library; an index trained by faiss is imported with fivf_set_centroids /
// fivf_load_lists instead.
int fivf_train(fivf_index* h, const float* x, int64_t n, int32_t niter, uint64_t seed) {
    if (!h || !x) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    const int nlist = h->nlist, dim = h->dim;
    if (h->ntotal() > 0) return fail(FIP_ERR_STATE, "train() must precede add()");
    if (n < nlist) return fail(FIP_ERR_INVALID, "Number of training points (" + std::to_string(n) +
                                                ") should be at least as large as number of clusters (" + std::to_string(nlist) + ")");
    if (niter < 0) niter = 10;
    int rc = fivf_use(h);
    if (rc) return rc;
    // deterministic shuffle (splitmix64-driven Fisher-Yates) for the subsample and the seeds
    std::vector<int64_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    uint64_t st = seed;
    auto next = [&]() { uint64_t z = (st += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
                        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); };
    for (int64_t i = n - 1; i > 0; --i) std::swap(order[(size_t)i], order[(size_t)(next() % (uint64_t)(i + 1))]);
    const int64_t nt = std::min<int64_t>(n, (int64_t)256 * nlist);
    DevBuf<float> tx;
    FIP_CUDA(tx.ensure((size_t)nt * dim));
    {
        std::vector<float> stage((size_t)std::min<int64_t>(nt, 65536) * dim);
        for (int64_t r0 = 0; r0 < nt; r0 += 65536) {
            const int64_t m = std::min<int64_t>(65536, nt - r0);
            for (int64_t i = 0; i < m; ++i)
                memcpy(stage.data() + (size_t)i * dim, x + (size_t)order[(size_t)(r0 + i)] * dim, sizeof(float) * dim);
            cudaError_t e = cudaMemcpy(tx.p + (size_t)r0 * dim, stage.data(), sizeof(float) * (size_t)m * dim, cudaMemcpyHostToDevice);
            if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
        }
    }
    // initial centroids: the first nlist rows of the shuffled training set
    cudaError_t e = cudaMemcpyAsync(h->cent.p, tx.p, sizeof(float) * (size_t)nlist * dim, cudaMemcpyDeviceToDevice, h->stream);
    if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
    std::vector<int32_t> assign((size_t)nt);
    std::vector<int64_t> off, perm;
    std::vector<float> hc;
    auto cleanup = [&](int code) { tx.release(); return code; };
    for (int it = 0; it < niter; ++it) {
        rc = fivf_update_cnorm(h);
        if (rc) return cleanup(rc);
        rc = fivf_assign_device(h, tx.p, nt, assign.data());
        if (rc) return cleanup(rc);
        counting_sort(assign.data(), nt, nlist, &off, &perm);
        e = h->perm.ensure((size_t)nt);
        if (e == cudaSuccess) e = h->doff.ensure((size_t)nlist + 1);
        if (e == cudaSuccess) e = cudaMemcpyAsync(h->perm.p, perm.data(), sizeof(int64_t) * nt, cudaMemcpyHostToDevice, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(h->doff.p, off.data(), sizeof(int64_t) * ((size_t)nlist + 1), cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
        ivf::centroid_mean_kernel<<<dim3(nlist, (dim + 255) / 256), 256, 0, h->stream>>>(tx.p, h->perm.p, h->doff.p, dim, h->cent.p);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
        // empty clusters: split the (currently) largest cluster, symmetric perturbation
        std::vector<int64_t> size((size_t)nlist);
        int n_empty = 0;
        for (int l = 0; l < nlist; ++l) { size[l] = off[(size_t)l + 1] - off[l]; n_empty += size[l] == 0; }
        if (n_empty) {
            hc.resize((size_t)nlist * dim);
            e = cudaMemcpy(hc.data(), h->cent.p, sizeof(float) * hc.size(), cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
            for (int l = 0; l < nlist; ++l) {
                if (size[l] != 0) continue;
                const int big = (int)(std::max_element(size.begin(), size.end()) - size.begin());
                const float eps = 1.0f / 1024.0f;
                for (int c = 0; c < dim; ++c) {
                    const float v = hc[(size_t)big * dim + c];
                    const float s = (c & 1) ? eps : -eps;
                    hc[(size_t)l * dim + c] = v * (1.0f + s);
                    hc[(size_t)big * dim + c] = v * (1.0f - s);
                }
                size[l] = size[big] / 2;
                size[big] -= size[l];
            }
            e = cudaMemcpy(h->cent.p, hc.data(), sizeof(float) * hc.size(), cudaMemcpyHostToDevice);
            if (e != cudaSuccess) { tx.release(); FIP_CUDA(e); }
        }
    }
    tx.release();
    rc = fivf_update_cnorm(h);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    h->trained = true;
    return FIP_OK;
}

// Nearest centroid (list number) of n host rows.
int fivf_assign(fivf_index* h, const float* x, int64_t n, int32_t* out) {
    if (!h || (n > 0 && (!x || !out))) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->trained) return fail(FIP_ERR_STATE, "IVF index is not trained (no centroids)");
    int rc = fivf_use(h);
    if (rc) return rc;
    for (int64_t r0 = 0; r0 < n; r0 += kAssignChunk) {
        const int64_t m = std::min<int64_t>(kAssignChunk, n - r0);
        FIP_CUDA(h->stage.ensure((size_t)m * h->dim));
        FIP_CUDA(cudaMemcpyAsync(h->stage.p, x + (size_t)r0 * h->dim, sizeof(float) * (size_t)m * h->dim, cudaMemcpyHostToDevice, h->stream));
        rc = fivf_assign_device(h, h->stage.p, m, out + r0);
        if (rc) return rc;
    }
    return FIP_OK;
}

static int fivf_reserve_pending(fivf_index* h, int64_t want) {
    if (want <= h->pend_cap) return FIP_OK;
    const int64_t new_cap = std::max<int64_t>(want, h->pend_cap + h->pend_cap / 2);
    float* p = nullptr;
    FIP_CUDA(cudaMalloc((void**)&p, (size_t)new_cap * h->dim * sizeof(float)));
    if (h->pend_x && h->pend_n) {
        cudaError_t e = cudaMemcpyAsync(p, h->pend_x, (size_t)h->pend_n * h->dim * sizeof(float), cudaMemcpyDeviceToDevice, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { cudaFree(p); FIP_CUDA(e); }
    }
    cudaFree(h->pend_x);
    h->pend_x = p;
    h->pend_cap = new_cap;
    return FIP_OK;
}

// IndexIVF::add / add_with_ids: assign each row to its nearest centroid and append it to that list.
// ids == null: sequential ids starting at ntotal (faiss add()).  `assign` (optional, [n]) overrides
// the coarse assignment (an imported index keeps the lists it was built with).
static int fivf_add_impl(fivf_index* h, const float* x, int64_t n, const int64_t* ids, const int32_t* assign) {
    if (!h || (n > 0 && !x)) return fail(FIP_ERR_INVALID, "null argument");
    if (n < 0) return fail(FIP_ERR_INVALID, "negative row count");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->trained) return fail(FIP_ERR_STATE, "IVF index is not trained (no centroids)");
    int rc = fivf_use(h);
    if (rc) return rc;
    rc = fivf_reserve_pending(h, h->pend_n + n);
    if (rc) return rc;
    const size_t a0 = h->pend_assign.size();
    h->pend_assign.resize(a0 + (size_t)n);
    for (int64_t r0 = 0; r0 < n; r0 += 65536) {
        const int64_t m = std::min<int64_t>(65536, n - r0);
        float* dst = h->pend_x + (size_t)(h->pend_n + r0) * h->dim;
        cudaError_t e = cudaMemcpyAsync(dst, x + (size_t)r0 * h->dim, sizeof(float) * (size_t)m * h->dim, cudaMemcpyHostToDevice, h->stream);
        if (e != cudaSuccess) { h->pend_assign.resize(a0); FIP_CUDA(e); }
        if (assign) {
            for (int64_t i = 0; i < m; ++i) {
                const int32_t l = assign[r0 + i];
                if (l < 0 || l >= h->nlist) { h->pend_assign.resize(a0); cudaStreamSynchronize(h->stream); return fail(FIP_ERR_INVALID, "list number out of range"); }
                h->pend_assign[a0 + (size_t)(r0 + i)] = l;
            }
        } else {
            rc = fivf_assign_device(h, dst, m, h->pend_assign.data() + a0 + r0);
            if (rc) { h->pend_assign.resize(a0); return rc; }
        }
    }
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    for (int64_t i = 0; i < n; ++i) h->pend_ids.push_back(ids ? ids[i] : h->next_id + i);
    h->next_id += n;
    h->pend_n += n;
    return FIP_OK;
}
int fivf_add(fivf_index* h, const float* x, int64_t n, const int64_t* ids) { return fivf_add_impl(h, x, n, ids, nullptr); }
int fivf_add_preassigned(fivf_index* h, const float* x, int64_t n, const int64_t* ids, const int32_t* assign) {
    if (n > 0 && !assign) return fail(FIP_ERR_INVALID, "null assignment");
    return fivf_add_impl(h, x, n, ids, assign);
}

int fivf_list_sizes(fivf_index* h, int64_t* out) {
    if (!h || !out) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = fivf_use(h);
    if (rc) return rc;
    rc = fivf_rebuild(h);
    if (rc) return rc;
    for (int l = 0; l < h->nlist; ++l) out[l] = h->off[(size_t)l + 1] - h->off[l];
    return FIP_OK;
}

// The inverted lists, concatenated in list order: codes [ntotal, dim] fp32, ids [ntotal].
int fivf_export_lists(fivf_index* h, float* codes, int64_t* ids) {
    if (!h) return fail(FIP_ERR_INVALID, "null index");
    std::lock_guard<std::mutex> lk(h->mu);
    int rc = fivf_use(h);
    if (rc) return rc;
    rc = fivf_rebuild(h);
    if (rc) return rc;
    if (h->n_sorted == 0) return FIP_OK;
    if (codes) FIP_CUDA(cudaMemcpyAsync(codes, h->x, sizeof(float) * (size_t)h->n_sorted * h->dim, cudaMemcpyDeviceToHost, h->stream));
    if (ids) FIP_CUDA(cudaMemcpyAsync(ids, h->ids, sizeof(int64_t) * (size_t)h->n_sorted, cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    return FIP_OK;
}

int fivf_search_device(fivf_index* h, const float* queries_dev, int64_t nq, int32_t k, int32_t nprobe,
                       float* distances_dev, int64_t* ids_dev) {
    int rc = fivf_check_search(h, queries_dev, nq, k, distances_dev, ids_dev);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    rc = fivf_use(h);
    if (rc) return rc;
    return fivf_search_checked(h, queries_dev, (int)nq, k, nprobe, distances_dev, ids_dev);
}

int fivf_search(fivf_index* h, const float* queries, int64_t nq, int32_t k, int32_t nprobe,
                float* distances, int64_t* ids) {
    int rc = fivf_check_search(h, queries, nq, k, distances, ids);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    std::lock_guard<std::mutex> lk(h->mu);
    rc = fivf_use(h);
    if (rc) return rc;
    const size_t nqd = (size_t)nq * h->dim, nqk = (size_t)nq * k;
    FIP_CUDA(h->h_q.ensure(nqd));
    FIP_CUDA(h->h_d.ensure(nqk));
    FIP_CUDA(h->h_i.ensure(nqk));
    FIP_CUDA(h->q32.ensure(nqd));
    FIP_CUDA(h->out_d.ensure(nqk));
    FIP_CUDA(h->out_i.ensure(nqk));
    memcpy(h->h_q.p, queries, sizeof(float) * nqd);
    FIP_CUDA(cudaMemcpyAsync(h->q32.p, h->h_q.p, sizeof(float) * nqd, cudaMemcpyHostToDevice, h->stream));
    rc = fivf_search_checked(h, h->q32.p, (int)nq, k, nprobe, h->out_d.p, h->out_i.p);
    if (rc) return rc;
    FIP_CUDA(cudaMemcpyAsync(h->h_d.p, h->out_d.p, sizeof(float) * nqk, cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaMemcpyAsync(h->h_i.p, h->out_i.p, sizeof(int64_t) * nqk, cudaMemcpyDeviceToHost, h->stream));
    FIP_CUDA(cudaStreamSynchronize(h->stream));
    memcpy(distances, h->h_d.p, sizeof(float) * nqk);
    memcpy(ids, h->h_i.p, sizeof(int64_t) * nqk);
    return FIP_OK;
}

int fivf_set_option(fivf_index* h, const char* key, int64_t value) {
    if (!h || !key) return fail(FIP_ERR_INVALID, "null argument");
    const std::string k(key);
    if (k == "force_safe") h->force_safe = (int)value;
    else if (k == "live_thresholds") h->live_thr = (int)value;
    else if (k == "max_queries_per_tile") h->qb_max = (int)value;
    else if (k == "profile_events") h->profile = (int)value;
    else if (k == "tensor_min_queries") h->tc_min = (int)value;
    else if (k == "tensor_format") h->tensor_format = value ? 1 : 0;
    else if (k == "coarse_tensor") h->coarse_tensor = value < 0 ? 0 : (value > 2 ? 2 : (int)value);
    else if (k == "device_plan") h->device_plan = value < 0 ? 0 : (value > 2 ? 2 : (int)value);
    else return fail(FIP_ERR_INVALID, "unknown option " + k);
    return FIP_OK;
}

int64_t fivf_get_stat(const fivf_index* h, const char* key) {
    if (!h || !key) return -1;
    const std::string k(key);
    if (k == "launches") return h->st_launches;
    if (k == "phases") return h->st_phases;
    if (k == "tiles") return h->st_tiles;
    if (k == "tiles16") return h->st_tiles16;
    if (k == "rows_scanned") return h->st_rows;
    if (k == "pairs_scored") return h->st_pairs;
    if (k == "overflow") return h->st_overflow;
    if (k == "candidate_capacity") return h->st_cap;
    if (k == "plan_us") return h->st_plan_us;
    if (k == "distinct_rows") return h->st_distinct_rows;
    if (k == "scan_ns") return h->st_scan_ns;
    if (k == "tensor_items") return h->st_tc_items;
    if (k == "tensor_pairs") return h->st_tc_pairs;
    if (k == "crowded_window") return h->st_crowded;
    if (k == "crowded_window_unresolved") return h->st_crowded_final;
    if (k == "device_plan") return h->st_device_plan;
    if (k == "coarse_tensor") return h->st_coarse_tensor;
    if (k == "tensor_f16") return h->st_f16;
    if (k == "scan16_ctas_per_sm") return h->st_occ16;
    return -1;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// Test hook (no GPU needed): runs the host planner on a synthetic probe table and checks its
// invariant -- every (query, probed list) pair is scored over exactly the rows of that list, once
// (dense prefix + skip-aware main phase, FFMA2 tiles + tensor items) -- and times it.
// sizes [nlist]; probes [nq, nprobe] (list numbers, -1 = none, no duplicates within a query).
// out[0..7] = phases, tiles, tensor_items, rows, pairs, plan_us (best of iters), bad_pairs, qmap_rows.
extern "C" int fivf_debug_plan(int32_t nlist, const int64_t* sizes, const int64_t* probes, int32_t nq,
                               int32_t nprobe, int32_t k, int32_t cap, int32_t safe, int32_t tc_min,
                               int32_t iters, int64_t* out) {
    if (nlist < 1 || !sizes || !probes || !out || nq < 1 || nprobe < 1 || k < 1 || cap < k) return FIP_ERR_INVALID;
    fivf_index h;
    h.nlist = nlist;
    h.off.assign((size_t)nlist + 1, 0);
    for (int l = 0; l < nlist; ++l) h.off[(size_t)l + 1] = h.off[l] + sizes[l];
    IvfPlan plan;
    int64_t best = INT64_MAX;
    for (int it = 0; it < std::max(1, iters); ++it) {
        const auto t0 = std::chrono::steady_clock::now();
        build_plan(&h, probes, nq, nprobe, k, cap, safe != 0, 16, tc_min, &plan);
        best = std::min<int64_t>(best, std::chrono::duration_cast<std::chrono::microseconds>(
                                           std::chrono::steady_clock::now() - t0).count());
    }
    // coverage[q][j] = rows of probes[q][j] scored for q
    std::vector<int64_t> cov((size_t)nq * nprobe, 0);
    auto list_of = [&](uint32_t pos) {
        return (int)(std::upper_bound(h.off.begin(), h.off.end(), (int64_t)pos) - h.off.begin()) - 1;
    };
    auto credit = [&](int q, int l, int64_t rows) -> bool {
        for (int j = 0; j < nprobe; ++j)
            if (probes[(size_t)q * nprobe + j] == l) { cov[(size_t)q * nprobe + j] += rows; return true; }
        return false;
    };
    int64_t bad = 0;
    for (const ivf::Tile& t : plan.tiles) {
        const int l = list_of(t.row0);
        if (t.nrows == 0 || t.nrows > (uint32_t)ivf::SC_ROWS || t.nq == 0 || t.nq > 16 ||
            (int64_t)t.row0 + t.nrows > h.off[(size_t)l + 1] || (int64_t)t.row0 - t.rel0 != h.off[l]) { bad++; continue; }
        for (uint32_t i = 0; i < t.nq; ++i) {
            const ivf::TileQuery tq = plan.tile_q[t.qoff + i];
            const int64_t lo = std::max<int64_t>(t.rel0, tq.skip), hi = (int64_t)t.rel0 + t.nrows;
            if (!credit(tq.q, l, std::max<int64_t>(0, hi - lo))) bad++;
        }
    }
    for (const ivf::TcItem& t : plan.tc_items) {
        const int l = list_of(t.row0);
        if ((int64_t)t.list0 != h.off[l] || (int64_t)t.row_end != h.off[(size_t)l + 1] || t.qrows == 0 || t.qrows > 128 ||
            t.qrows % 32 != 0 || t.qrow0 + t.qrows > plan.qmap.size()) { bad++; continue; }
        const int64_t r0 = (int64_t)t.row0 - t.list0, r1 = std::min<int64_t>(r0 + tc::BN, (int64_t)t.row_end - t.list0);
        for (uint32_t i = 0; i < t.qrows; ++i) {
            const int q = plan.qmap[t.qrow0 + i];
            if (q < 0) continue;
            const int64_t lo = std::max<int64_t>(r0, plan.qskip[t.qrow0 + i]);
            if (!credit(q, l, std::max<int64_t>(0, r1 - lo))) bad++;
        }
    }
    // dense-phase prefixes were credited through their own tiles; every pair must now be complete
    for (int q = 0; q < nq; ++q)
        for (int j = 0; j < nprobe; ++j) {
            const int64_t l = probes[(size_t)q * nprobe + j];
            const int64_t want = (l >= 0 && l < nlist) ? sizes[l] : 0;
            if (cov[(size_t)q * nprobe + j] != want) bad++;
        }
    out[0] = (int64_t)plan.phases.size(); out[1] = (int64_t)plan.tiles.size(); out[2] = (int64_t)plan.tc_items.size();
    out[3] = plan.rows; out[4] = plan.pairs; out[5] = best; out[6] = bad; out[7] = (int64_t)plan.qmap.size();
    return FIP_OK;
}
