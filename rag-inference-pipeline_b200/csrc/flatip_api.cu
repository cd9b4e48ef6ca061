// C ABI (include/flatip_b200.h) and host-side orchestration of the B200 flat-IP search.
//
// A search is a short sequence of launches on one stream:
//   prepare queries -> [ scan phase p over rows [r_p, r_p+1)  ->  select/compact ] x P -> outputs
// The scan kernels (K1 tensor-core / K2 streaming) only FILTER: a score survives when it beats the
// query's running threshold (the k-th best score over all earlier phases) and is appended to that
// query's candidate list.  The select kernel (K3) then keeps the k best and publishes the new
// threshold.  Phase lengths grow geometrically (the expected number of survivors of a phase of m
// rows after n rows is k*m/n), so a 12.5 M-row shard needs 4 phases at k = 100.  If a list ever
// overflows (adversarial row order), the affected queries are redone in "safe mode", whose phases
// are too short to overflow by construction -- results are exact for every input.
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <vector>

#include "../../include/flatip_b200.h"
#include "common.cuh"
#include "dedupe.cuh"
#include "exchange.cuh"
#include "select.cuh"
#include "stream_scan.cuh"
#include "synthetic.cuh"
#include "tc_scan.cuh"

using namespace fip;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define FIP_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            cudaGetLastError();                                                              \
            return fail(_e == cudaErrorMemoryAllocation ? FIP_ERR_NOMEM : FIP_ERR_CUDA,      \
                        std::string(#expr) + ": " + cudaGetErrorString(_e));                 \
        }                                                                                    \
    } while (0)

constexpr int kMaxK = 2048;
constexpr int kMaxCap = 16384;  // select kernel sorts a whole list in shared memory (128 KB)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t ensure(size_t want) {
        if (want <= n) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) n = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
};
template <typename T>
struct PinnedBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t ensure(size_t want) {
        if (want <= n) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        n = 0;
        cudaError_t e = cudaMallocHost((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) n = want;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        n = 0;
    }
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) !=
                cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}

// 2-D K-major tile map: inner dim = `dim` elements, outer dim = rows; box = 128 bytes x 128 rows.
int make_tmap(CUtensorMap* map, const void* base, int dtype, int dim, uint64_t rows) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    const int esize = dtype_size(dtype);
    cuuint64_t dims[2] = {(cuuint64_t)dim, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)dim * esize};
    cuuint32_t box[2] = {(cuuint32_t)(tc::BK_BYTES / esize), (cuuint32_t)tc::BOX_ROWS};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMapDataType dt = dtype == kBF16  ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                             : dtype == kF16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                             : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUresult r = enc(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(FIP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return FIP_OK;
}

// flags: select.cuh flag_slot().  `sticky` (may be null): the summary words an earlier, never
// checked (asynchronous) search left behind are folded into it before they are cleared, so
// fip_search_status still sees them.  grid >= flag_words(nq) threads.
__global__ void init_lists_kernel(float* thr, uint32_t* cnt, uint32_t* flags, uint32_t* sticky, int nq) {
    init_lists_thread(ListInit{thr, cnt, flags, sticky, nq}, (int)(blockIdx.x * blockDim.x + threadIdx.x));
}
__global__ void fold_sticky_kernel(uint32_t* flags, uint32_t* sticky) {
    if (threadIdx.x < 2 && flags[threadIdx.x]) { atomicOr(sticky, 1u << threadIdx.x); flags[threadIdx.x] = 0u; }
}
__global__ void fill_padding_kernel(float* d, int64_t* id, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) { d[i] = kLowestF32; id[i] = -1; }
}
// dst rows[sel[j]] = src rows[j]
__global__ void gather_rows_f32_kernel(const float* src, const int* sel, float* dst, int nsel, int width) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < (int64_t)nsel * width) {
        int j = (int)(i / width), c = (int)(i - (int64_t)j * width);
        dst[i] = src[(size_t)sel[j] * width + c];
    }
}
__global__ void scatter_results_kernel(const float* sd, const int64_t* si, const int* sel, float* dd,
                                       int64_t* di, int nsel, int k) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < (int64_t)nsel * k) {
        int j = (int)(i / k), c = (int)(i - (int64_t)j * k);
        dd[(size_t)sel[j] * k + c] = sd[i];
        di[(size_t)sel[j] * k + c] = si[i];
    }
}

}  // namespace

// Request coalescing (fip_search, fip_group_search): host callers that arrive while a search is
// running are queued; whoever finds no leader becomes one, takes every queued request with the same
// k (up to max_batch queries) and runs them as ONE batch.  Below ~210 queries a flat scan is
// HBM-bound -- its cost does not depend on the batch size -- so riding along is free.
struct HostRequest {
    const float* q; int64_t nq; int k; float* d; int64_t* i;
    int rc = 0; bool done = false; std::string err;
};
struct Coalescer {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<HostRequest*> pending;
    bool leader_active = false;
    int enabled = 1;
    int64_t max_batch = 4096;
    int64_t st_batches = 0, st_calls = 0, st_coalesced_calls = 0, st_last_batch_calls = 0;
    int64_t st_queries = 0, st_distinct = 0;   // in-batch dedupe: queries received / actually searched

    void count(size_t calls) {
        st_batches++;
        st_calls += (int64_t)calls;
        st_last_batch_calls = (int64_t)calls;
        if (calls > 1) st_coalesced_calls += (int64_t)calls;
    }
    // run_batch(const std::vector<HostRequest*>&) -> int; its error text is taken from g_last_error
    template <typename F>
    int submit(HostRequest& me, F&& run_batch) {
        if (!enabled) {
            std::vector<HostRequest*> one{&me};
            return run_batch(one);
        }
        std::unique_lock<std::mutex> lk(mu);
        pending.push_back(&me);
        while (!me.done) {
            if (leader_active) {
                cv.wait(lk);
                continue;
            }
            // become the leader for one batch: the oldest request sets k; every queued request with
            // that k rides along, in arrival order, while the batch stays within max_batch queries
            leader_active = true;
            std::vector<HostRequest*> batch;
            const int bk = pending.front()->k;
            int64_t total = 0;
            for (auto it = pending.begin(); it != pending.end();) {
                HostRequest* p = *it;
                if (p->k == bk && (batch.empty() || total + p->nq <= max_batch)) {
                    batch.push_back(p);
                    total += p->nq;
                    it = pending.erase(it);
                } else {
                    ++it;
                }
            }
            lk.unlock();
            const int brc = run_batch(batch);
            const std::string berr = brc ? g_last_error : std::string();
            lk.lock();
            for (auto* p : batch) { p->rc = brc; p->err = berr; p->done = true; }
            leader_active = false;
            cv.notify_all();
        }
        lk.unlock();
        if (me.rc) g_last_error = me.err;
        return me.rc;
    }
};

struct fip_index {
    fip_config cfg{};
    int num_sms = 0;
    cudaStream_t stream = nullptr;
    void* corpus = nullptr;
    int64_t n_rows = 0;
    int64_t cap_rows = 0;
    std::mutex mu;

    Coalescer co;   // fip_search callers that arrive during a search share the next batch

    // options
    int variant = FIP_VARIANT_AUTO;
    int stream_max_batch = 0;  // 0 = by arithmetic intensity (nq <= sizeof(storage))
    int force_safe = 0;
    int query_split = 0;  // 0 = auto (1 for fp16 / fp32 storage, 2 for bf16)
    int cta_pair = 0;     // 0 = auto, 1 = single-CTA MMAs, 2 = CTA pairs (cta_group::2)
    int max_clusters[2] = {-1, -1};  // co-resident 2-CTA clusters of the pair kernel (plain / deferred epilogue)
    int defer_epilogue = 0;  // 0 = auto (k >= 256), 1 = never, 2 = always
    int fused_rescore = 1;   // exact rescoring inside the final select pass (0: separate kernel)
    int64_t st_deferred = 0;
    int profile = 0;      // 1: bracket every scan launch with CUDA events (bench roofline)
    int debug_counts = 0; // 1: read the list lengths back after every filter launch (synchronises)
    int64_t st_list_entries = 0, st_list_max = 0;
    std::vector<cudaEvent_t> ev;  // start/stop pairs
    // The events form a ring; every search records which slice it used (EvSet).  Timings are
    // COLLECTED when a search is finished / its status is read: asynchronous searches accumulate, so
    // the statistics below cover every search since the previous collection (one search for the
    // blocking calls): sums per scan-launch index, plus how many searches contributed.
    struct EvSet { uint64_t start; int count; std::vector<int64_t> rows; int launches; };
    uint64_t ev_head = 0;
    std::vector<EvSet> ev_sets;
    double st_scan_ms = 0.0;      // device time of the scan kernels of the collected searches
    int64_t st_scan_launches = 0, st_scan_searches = 0;
    std::vector<float> st_scan_each, st_gap_each;  // per scan launch index (summed); gap = time to the next scan
    std::vector<int64_t> st_scan_rows;             // corpus rows of each timed scan

    // per-search workspace
    DevBuf<float> thr, q_scale, q_eps, q32, out_d;
    DevBuf<float> max_norm2;  // [1] max squared row norm of the shard (certificate radius)
    int certify = 1;          // 0: skip the exactness certificate of the tensor-core variant
    int live_thr = 1;         // 0: thresholds only move between scan launches (geometric phases)
    DevBuf<uint32_t> hist, hbase;   // live thresholds: [nq, 256] score histogram, [nq] bucket origin
    DevBuf<uint32_t> cnt, overflow;   // overflow = the flag words (select.cuh flag_slot)
    DevBuf<uint32_t> sticky;          // [1] summary flags of asynchronous searches nobody checked yet
    struct PendingSearch {            // the search enqueued last and not finished yet
        const float* q32 = nullptr; int nq = 0, k = 0; float* out_d = nullptr; int64_t* out_i = nullptr;
        cudaStream_t stream = nullptr; bool safe = false; bool active = false;
        int variant = 0, phases = 0, cap = 0, split = 0, launches = 0;
    } pend;
    DevBuf<uint64_t> cand;
    DevBuf<uint8_t> qprep;  // prepared query terms in the MMA operand type
    DevBuf<int64_t> out_i;
    DevBuf<int> sel;
    DevBuf<float> sub_q, sub_d;
    DevBuf<int64_t> sub_i;
    PinnedBuf<float> h_q, h_d;
    PinnedBuf<int64_t> h_i;
    PinnedBuf<uint32_t> h_flags;
    // in-batch duplicate elimination (dedupe.cuh): 0 = off, 1 = coalesced batches (default), 2 = every
    // host batch
    int dedupe = 1;
    DevBuf<uint64_t> dd_hash;
    DevBuf<int> dd_rep, dd_pos, dd_sel, dd_map, dd_count;
    PinnedBuf<int> h_dd_count;
    DevBuf<float> dd_q, dd_d;
    DevBuf<int64_t> dd_i;
    DevBuf<float> add_stage;
    PinnedBuf<uint8_t> h_add[2];   // fip_add_host: double-buffered pinned staging

    CUtensorMap tmap_x{};
    bool tmap_x_valid = false;
    uint64_t corpus_gen = 1;   // bumped by everything that changes the stored rows

    // fp32 storage only: fp16 FILTER COPY of the corpus for the tensor-core variant.  The scan only
    // filters (survivors are rescored exactly from the fp32 rows), so it may read 2 bytes per
    // element instead of 4: half the HBM time and twice the tensor rate of the tf32 form.  Built on
    // first use after a change of the corpus (+50 % device memory; skipped when it does not fit or
    // when some row leaves fp16's range).
    int filter_copy = 1;
    __half* corpus16 = nullptr;
    CUtensorMap tmap_x16{};
    uint64_t c16_gen = 0;      // corpus_gen the copy (or the decision not to have one) belongs to
    bool c16_usable = false;
    float c16_xmax = 0.f;      // max row norm (host copy)
    PinnedBuf<float> h_mn2;

    // stats
    int64_t st_launches = 0, st_variant = 0, st_phases = 0, st_overflow = 0, st_cap = 0, st_pair = 0;
    int64_t st_uncert = 0, st_split = 0, st_stream_redo = 0, st_filter16 = 0;
    int64_t st_total_launches = 0, st_total_searches = 0;

    size_t row_bytes() const { return (size_t)cfg.dim * dtype_size(cfg.dtype); }
};

namespace {

int use_device(const fip_index* idx) {
    FIP_CUDA(cudaSetDevice(idx->cfg.device));
    return FIP_OK;
}

int reserve_rows(fip_index* idx, int64_t want) {
    if (want <= idx->cap_rows) return FIP_OK;
    int64_t new_cap = std::max<int64_t>(want, idx->cap_rows + idx->cap_rows / 2);
    void* p = nullptr;
    FIP_CUDA(cudaMalloc(&p, (size_t)new_cap * idx->row_bytes()));
    if (idx->corpus) {
        cudaError_t e = cudaMemcpyAsync(p, idx->corpus, (size_t)idx->n_rows * idx->row_bytes(),
                                        cudaMemcpyDeviceToDevice, idx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(idx->stream);
        if (e != cudaSuccess) { cudaFree(p); FIP_CUDA(e); }
        cudaFree(idx->corpus);
    }
    idx->corpus = p;
    idx->cap_rows = new_cap;
    idx->tmap_x_valid = false; idx->corpus_gen++;
    return FIP_OK;
}

// How many times the scanned row count may grow in ONE filter launch so that a query's candidate
// list is expected to stay within its capacity.
//   fixed thresholds (live = false): a launch over m rows after n rows leaves ~k m / n survivors,
//     so growth is limited to (cap/2 - k) / k  (lists half full on average data);
//   live thresholds: the launch raises its own thresholds (common.cuh), survivors grow like
//     k log(ratio).  A refresh happens every kRefreshEvery appended entries, i.e. with
//     x = kRefreshEvery / k the threshold lags by a factor (1 + x) in rows: x / log(1 + x) more
//     survivors than ideal, plus ~20 % for the bucket resolution; the expected list is kept below
//     3/4 of the capacity.
// Overflow means "redo in safe mode", so these are performance bounds, not correctness ones.
double phase_ratio(int cap, int k_plan, int keep_max, bool live) {
    const double keff = std::max(k_plan, 16);
    const double fixed = 1.0 + std::max(2.0, ((double)cap / 2 - keff) / keff);
    if (!live) return fixed;
    const double x = (double)kRefreshEvery / keff;
    const double lag = 1.2 * x / std::log1p(x);
    const double budget = 0.75 * cap - std::max(k_plan, keep_max) - (double)kRefreshEvery;
    return std::max(fixed, std::exp(std::min(40.0, std::max(0.0, budget) / (lag * keff))));
}
int keep_max_for(int cap, int k_out, int k_plan, bool certify) {
    return certify ? std::min(cap / 2, k_out + 2 * (k_plan - k_out)) : 0;
}
int64_t phase_rows(int64_t pos, double ratio) {
    return std::max<int64_t>(256, (int64_t)std::min((double)pos * (ratio - 1.0), 4.0e9) / 256 * 256);
}
int count_phases(int64_t n, int cap, double ratio) {
    int64_t pos = std::min<int64_t>(n, cap);
    int phases = 1;
    while (pos < n) {
        pos += phase_rows(pos, ratio);
        phases++;
    }
    return phases;
}
// Candidate-list capacity per query.  The floor (4096, or 32 x k for large k) keeps the lists of a
// big batch small; every phase boundary costs a select pass and a scan launch (~50 us), so a small
// batch -- whose lists are cheap and whose select passes fit one wave of CTAs -- takes a larger
// capacity when that saves a phase.
int cap_for_k(int k_out, int k_plan, bool certify, bool live, bool safe, int nq, int64_t n) {
    if (safe) return kMaxCap;
    int cap = std::min(std::max(4096, 32 * next_pow2(k_plan)), kMaxCap);
    int best = count_phases(n, cap, phase_ratio(cap, k_plan, keep_max_for(cap, k_out, k_plan, certify), live));
    for (int c = cap * 2; c <= kMaxCap; c *= 2) {
        if ((int64_t)nq * c > (int64_t)148 * kMaxCap) break;
        const int ph = count_phases(n, c, phase_ratio(c, k_plan, keep_max_for(c, k_out, k_plan, certify), live));
        if (ph < best) { best = ph; cap = c; }
    }
    return cap;
}

struct SearchCtx {
    fip_index* idx;
    cudaStream_t stream;
    int launches = 0;
    const CUtensorMap* tmap_scan = nullptr;   // corpus map of the tensor-core scan (storage or fp16 filter copy)
    bool level0 = false;   // the caller's own search (not a redo of a subset): folds unchecked flags
};

template <typename T, int QB, int R>
int launch_stream_pass(SearchCtx& c, const ScanArgs& a, const float* q32, int q0) {
    fip_index* idx = c.idx;
    auto kern = stream_scan_kernel<T, QB, R>;
    const StreamPlan plan = plan_stream(QB, R, (size_t)a.dim * sizeof(T), a.dim);
    if (plan.warps == 0 || plan.slots < 2)
        return fail(FIP_ERR_INVALID, "dim too large for the streaming kernel (a row block does not fit "
                                     "in shared memory twice)");
    FIP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    const int64_t blocks = ((int64_t)(a.row_end - a.row_begin) + R - 1) / R;
    int grid = (int)std::min<int64_t>((int64_t)idx->num_sms, (blocks + plan.warps - 1) / plan.warps);
    grid = std::max(grid, 1);
    kern<<<grid, plan.warps * 32, plan.smem, c.stream>>>(a, q32, q0, plan.slots);
    FIP_CUDA(cudaGetLastError());
    c.launches++;
    return FIP_OK;
}

// Rows per block: 4 for 16-bit storage, 2 for fp32 (same bytes per block); 16-query passes halve
// it to stay inside the register budget.
template <typename T>
int launch_stream_phase(SearchCtx& c, const ScanArgs& a, const float* q32) {
    constexpr int R = sizeof(T) == 4 ? 2 : 4;
    for (int q0 = 0; q0 < a.nq;) {
        int rem = a.nq - q0;
        int qb = rem >= 16 ? 16 : next_pow2(rem);
        while (qb > 1 && plan_stream(qb, qb == 16 ? R / 2 : R, (size_t)a.dim * sizeof(T), a.dim).slots < 2) qb >>= 1;
        int rc;
        switch (qb) {
            case 1: rc = launch_stream_pass<T, 1, R>(c, a, q32, q0); break;
            case 2: rc = launch_stream_pass<T, 2, R>(c, a, q32, q0); break;
            case 4: rc = launch_stream_pass<T, 4, R>(c, a, q32, q0); break;
            case 8: rc = launch_stream_pass<T, 8, R>(c, a, q32, q0); break;
            default: rc = launch_stream_pass<T, 16, R / 2>(c, a, q32, q0); qb = 16; break;
        }
        if (rc) return rc;
        q0 += qb;
    }
    return FIP_OK;
}

template <int kFmt, int kSplit, int kCtas, int kStage = 0>
int launch_tc_kernel(SearchCtx& c, const ScanArgs& a, const CUtensorMap& tmap_q, int n_qtiles) {
    fip_index* idx = c.idx;
    auto kern = tc::tc_scan_kernel<kFmt, kSplit, kCtas, kStage>;
    constexpr int smem = tc::Cfg<kFmt, kSplit, kCtas, kStage>::SMEM_BYTES;
    FIP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int64_t n_ctiles = ((int64_t)(a.row_end - a.row_begin) + tc::BN - 1) / tc::BN;
    const int64_t group_tiles = n_ctiles * (n_qtiles / kCtas);
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(tc::NUM_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute attr[1];
    int groups = idx->num_sms / kCtas;
    if (kCtas == 2) {
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int& cached = idx->max_clusters[kStage > 0 ? 1 : 0];
        if (cached < 0) {
            cfg.gridDim = dim3(idx->num_sms / 2 * 2);
            int n = 0;
            FIP_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
            cached = std::max(1, n);
        }
        groups = std::min(groups, cached);
    }
    groups = (int)std::min<int64_t>(groups, group_tiles);
    cfg.gridDim = dim3((unsigned)(groups * kCtas));
    FIP_CUDA(cudaLaunchKernelEx(&cfg, kern, tmap_q, c.tmap_scan ? *c.tmap_scan : idx->tmap_x, a, n_qtiles));
    c.launches++;
    return FIP_OK;
}

constexpr int kStageSlots = 8;   // survivor slots per epilogue thread and tile (deferred epilogue)

int launch_tc_phase(SearchCtx& c, const ScanArgs& a, const CUtensorMap& tmap_q, int n_qtiles, int split,
                    bool pair, bool deferred) {
    if (deferred && split == 1 && !a.dense) {
        switch (a.dtype) {
            case kF32:
                return pair ? launch_tc_kernel<2, 1, 2, kStageSlots>(c, a, tmap_q, n_qtiles)
                            : launch_tc_kernel<2, 1, 1, kStageSlots>(c, a, tmap_q, n_qtiles);
            case kBF16:
                return pair ? launch_tc_kernel<1, 1, 2, kStageSlots>(c, a, tmap_q, n_qtiles)
                            : launch_tc_kernel<1, 1, 1, kStageSlots>(c, a, tmap_q, n_qtiles);
            default:
                return pair ? launch_tc_kernel<0, 1, 2, kStageSlots>(c, a, tmap_q, n_qtiles)
                            : launch_tc_kernel<0, 1, 1, kStageSlots>(c, a, tmap_q, n_qtiles);
        }
    }
    switch (a.dtype) {
        case kF32:
            return pair ? launch_tc_kernel<2, 1, 2>(c, a, tmap_q, n_qtiles)
                        : launch_tc_kernel<2, 1, 1>(c, a, tmap_q, n_qtiles);
        case kBF16:
            if (split == 2)
                return pair ? launch_tc_kernel<1, 2, 2>(c, a, tmap_q, n_qtiles)
                            : launch_tc_kernel<1, 2, 1>(c, a, tmap_q, n_qtiles);
            return pair ? launch_tc_kernel<1, 1, 2>(c, a, tmap_q, n_qtiles)
                        : launch_tc_kernel<1, 1, 1>(c, a, tmap_q, n_qtiles);
        default:
            if (split == 2)
                return pair ? launch_tc_kernel<0, 2, 2>(c, a, tmap_q, n_qtiles)
                            : launch_tc_kernel<0, 2, 1>(c, a, tmap_q, n_qtiles);
            return pair ? launch_tc_kernel<0, 1, 2>(c, a, tmap_q, n_qtiles)
                        : launch_tc_kernel<0, 1, 1>(c, a, tmap_q, n_qtiles);
    }
}

// How a (sub-)search is run.  rigor 0 = the index's configured variant; 1 = a tighter filter
// (two query terms for 16-bit storage, else the exact streaming kernel); 2 = streaming kernel.
struct Mode {
    bool safe = false;  // phases short enough that a candidate list can never overflow
    int rigor = 0;
};
struct SearchInfo {
    int variant = 0, phases = 0, cap = 0, split = 0;
    bool certified_path = false;  // the run can raise "uncertified" flags
};

int launch_select(SearchCtx& c, fip_index* idx, int nq, int k, int cap, bool final_pass, float* out_d,
                  int64_t* out_i, const float* out_scale, int64_t dense_count = -1,
                  int keep_max = 0 /* > 0: exactness window on (tensor-core variant) */,
                  bool live = false /* seed the live-threshold histogram */,
                  const float* rescore_q32 = nullptr /* final pass: recompute every score exactly first */) {
    SelectArgs s{};
    s.cand = idx->cand.p; s.cnt = idx->cnt.p; s.thr = idx->thr.p; s.cap = (uint32_t)cap; s.k = k;
    s.final_pass = final_pass ? 1 : 0; s.out_d = out_d; s.out_i = out_i; s.id_base = idx->cfg.id_base;
    s.out_scale = out_scale; s.overflow = idx->overflow.p; s.dense_count = (int)dense_count;
    if (keep_max > 0) {
        s.win_eps = idx->q_eps.p; s.win_scale = idx->q_scale.p; s.keep_max = keep_max;
        s.uncert = idx->overflow.p + 1;
    }
    if (live && !final_pass) { s.hist = idx->hist.p; s.hbase = idx->hbase.p; }
    // (a finishing pass also parks the window's candidates, up to keep_max, behind the list)
    size_t smem = ((size_t)cap + (size_t)std::max(next_pow2(std::min(k, cap)), rescore_q32 ? keep_max : 0)) * sizeof(uint64_t);
    if (rescore_q32) {
        s.rs_corpus = idx->corpus; s.rs_q32 = rescore_q32; s.rs_dim = idx->cfg.dim;
        const bool big = select_wants_big(cap);
        auto go = [&](auto kern) -> int {
            FIP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            kern<<<nq, big ? kSelectThreadsBig : kSelectThreads, smem, c.stream>>>(s);
            return FIP_OK;
        };
        int rc;
        switch (idx->cfg.dtype) {
            case kF32: rc = big ? go(select_rescore_kernel_big<float>) : go(select_rescore_kernel<float>); break;
            case kF16: rc = big ? go(select_rescore_kernel_big<__half>) : go(select_rescore_kernel<__half>); break;
            default: rc = big ? go(select_rescore_kernel_big<__nv_bfloat16>) : go(select_rescore_kernel<__nv_bfloat16>); break;
        }
        if (rc) return rc;
    } else if (select_wants_big(cap)) {
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel_big<<<nq, kSelectThreadsBig, smem, c.stream>>>(s);
    } else {
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel<<<nq, kSelectThreads, smem, c.stream>>>(s);
    }
    FIP_CUDA(cudaGetLastError());
    c.launches++;
    return FIP_OK;
}

// One complete search of `nq` device-resident fp32 queries; results to device out_d / out_i.
// fp32 storage: make the fp16 filter copy current (see fip_index::corpus16).  *usable = the copy
// exists and every stored row is inside fp16's range.  Synchronises `stream` once per corpus change.
int ensure_filter_copy(fip_index* idx, cudaStream_t stream, bool* usable) {
    *usable = false;
    if (idx->c16_gen == idx->corpus_gen) { *usable = idx->c16_usable; return FIP_OK; }
    idx->c16_gen = idx->corpus_gen;
    idx->c16_usable = false;
    if (idx->corpus16) { cudaFree(idx->corpus16); idx->corpus16 = nullptr; }
    FIP_CUDA(idx->h_mn2.ensure(1));
    FIP_CUDA(cudaMemcpyAsync(idx->h_mn2.p, idx->max_norm2.p, sizeof(float), cudaMemcpyDeviceToHost, stream));
    FIP_CUDA(cudaStreamSynchronize(stream));
    const float mn2 = idx->h_mn2.p[0];
    if (!(mn2 > 1e-30f && mn2 < 4.0e9f)) return FIP_OK;   // ||x||^2 < 4e9 => every component < 65504
    const size_t elems = (size_t)idx->n_rows * idx->cfg.dim;
    if (cudaMalloc((void**)&idx->corpus16, elems * sizeof(__half)) != cudaSuccess) {
        cudaGetLastError();
        idx->corpus16 = nullptr;
        return FIP_OK;
    }
    convert_rows_kernel<__half><<<idx->num_sms * 16, 256, 0, stream>>>((const float*)idx->corpus, idx->corpus16, (int64_t)elems);
    FIP_CUDA(cudaGetLastError());
    int rc = make_tmap(&idx->tmap_x16, idx->corpus16, kF16, idx->cfg.dim, (uint64_t)idx->n_rows);
    if (rc) return rc;
    idx->c16_xmax = sqrtf(mn2);
    idx->c16_usable = true;
    *usable = true;
    return FIP_OK;
}

// Does not synchronise (except once after a corpus change, to build the fp16 filter copy).
int run_search(SearchCtx& c, const float* q32, int nq, int k, float* out_d, int64_t* out_i,
               Mode mode, SearchInfo* info) {
    fip_index* idx = c.idx;
    const int64_t n = idx->n_rows;
    const int dim = idx->cfg.dim;
    const int dtype = idx->cfg.dtype;
    *info = SearchInfo{};
    if (flag_words(nq) > idx->overflow.n) {
        // growing the flag buffer loses its contents: fold unchecked summary words first
        if (idx->overflow.p && c.level0) {
            fold_sticky_kernel<<<1, 32, 0, c.stream>>>(idx->overflow.p, idx->sticky.p);
            FIP_CUDA(cudaStreamSynchronize(c.stream));
        }
        FIP_CUDA(idx->overflow.ensure(std::max<size_t>(flag_words(nq), flag_words(4096))));
        FIP_CUDA(cudaMemsetAsync(idx->overflow.p, 0, 2 * sizeof(uint32_t), c.stream));
    }
    if (n == 0) {
        int64_t tot = (int64_t)nq * k;
        fill_padding_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, c.stream>>>(out_d, out_i, tot);
        FIP_CUDA(cudaGetLastError());
        if (c.level0) fold_sticky_kernel<<<1, 32, 0, c.stream>>>(idx->overflow.p, idx->sticky.p);
        FIP_CUDA(cudaMemsetAsync(idx->overflow.p, 0, flag_words(nq) * sizeof(uint32_t), c.stream));
        c.launches++;
        return FIP_OK;
    }
    int variant = idx->variant;
    // fp32 storage: the tensor-core variant filters on the fp16 copy when there is one
    bool copy16 = false;
    if (dtype == kF32 && idx->filter_copy && mode.rigor == 0 && variant != FIP_VARIANT_STREAM &&
        (dim * 2) % tc::BK_BYTES == 0) {
        int rc16 = ensure_filter_copy(idx, c.stream, &copy16);
        if (rc16) return rc16;
    }
    const int fdtype = copy16 ? (int)kF16 : dtype;   // operand type of the tensor-core filter
    const bool tc_ok = (dim * dtype_size(fdtype)) % tc::BK_BYTES == 0;
    if (variant == FIP_VARIANT_AUTO) {
        // Arithmetic intensity of a flat scan is 2*nq / sizeof(storage) flop per corpus byte.  The
        // CUDA-core kernel (FMA + unpack + LDS per element) keeps up with HBM up to ~2 flop/byte
        // (measured: 1.00 of the copy bandwidth at nq = 1 and 0.89-0.92 at nq = 2 for 16-bit storage
        // -- where the tensor-core kernel holds 1.05 from nq = 2 on, so 16-bit storage streams only
        // single queries -- and 0.94-0.98 at nq = 4 for fp32 storage); above it the tensor pipe is needed to stay on the HBM
        // roof, and above ~210 flop/byte (nq > ~210 for 16-bit storage) the tensor pipe itself is
        // the roof.  The streaming kernel maps the 32 lanes of a warp onto the 16-byte vectors of a
        // row, so rows shorter than 1 KB (2 x 512 bytes) leave lanes idle (0.68-0.83 measured at
        // 256 / 768-byte rows): those go to the tensor-core kernel at any batch size.
        // stream_max_batch > 0 overrides the rule.
        const bool long_rows = (size_t)dim * dtype_size(dtype) >= 1024;
        // (with the fp16 filter copy the tensor-core scan reads half the bytes the streaming kernel
        // would: it wins at every batch size)
        const int stream_max = idx->stream_max_batch > 0 ? idx->stream_max_batch
                               : copy16                  ? 0
                                                         : (long_rows ? (dtype == kF32 ? 4 : 1) : 0);
        variant = (tc_ok && nq > stream_max) ? FIP_VARIANT_TENSOR : FIP_VARIANT_STREAM;
    }
    if (variant == FIP_VARIANT_TENSOR && !tc_ok)
        return fail(FIP_ERR_INVALID, "tensor-core variant needs rows that are a multiple of 128 bytes "
                                     "(dim % 64 == 0 for fp16/bf16, dim % 32 == 0 for fp32)");
    // Query terms of the tensor-core filter: one term (fp16 / bf16 / tf32) by default; rigor 1
    // adds a second 16-bit term (error 2^-17 instead of 2^-11 / 2^-9).
    int split = dtype == kF32 ? 1 : (idx->query_split ? idx->query_split : 1);
    if (variant == FIP_VARIANT_TENSOR && mode.rigor == 1) {
        if (dtype != kF32 && split == 1) split = 2;
        else variant = FIP_VARIANT_STREAM;
    }
    if (mode.rigor >= 2) variant = FIP_VARIANT_STREAM;
    // The tensor-core scan filters on rounded operands, so its lists keep more than k candidates
    // and the survivors are rescored exactly (fp32 FMA on the stored values) before the final
    // selection.  With the certificate on (default) every candidate inside the query's exactness
    // window stays (select.cuh): k_plan is only the size the phases are planned for and keep_max
    // the truncation limit.  With it off a fixed margin of k_plan - k extra candidates is kept.
    const bool rescore = variant == FIP_VARIANT_TENSOR;
    const bool certify = rescore && idx->certify != 0;
    const int k_out = k;
    // (bf16 single-term filter: eps is ~6x that of fp16, its window holds ~0.65 k extra candidates
    // on Gaussian data by the Cauchy-Schwarz bound)
    const int k_plan = !rescore ? k
                       : (dtype == kBF16 && split == 1) ? k + std::max(48, 3 * k / 4)
                                                        : k + std::max(16, k / 4);
    const bool live = idx->live_thr && !mode.safe;
    const int cap = cap_for_k(k_out, k_plan, certify, live, mode.safe, nq, n);
    const int keep_max = keep_max_for(cap, k_out, k_plan, certify);
    const int k_sel = certify ? k_out : k_plan;   // what the intermediate selects look for
    // Deferred epilogue (tc_scan.cuh): pays off when most tiles hold survivors, i.e. for large k
    // (survivors per query ~ k log(N / cap)); "defer_epilogue" 1 / 2 forces it off / on.
    const bool deferred = idx->defer_epilogue == 2 || (idx->defer_epilogue == 0 && k_sel >= 256);
    idx->st_deferred = (variant == FIP_VARIANT_TENSOR && deferred && split == 1) ? 1 : 0;
    FIP_CUDA(idx->thr.ensure(nq));
    FIP_CUDA(idx->cnt.ensure(nq));
    FIP_CUDA(idx->cand.ensure((size_t)nq * cap));

    const ListInit linit{idx->thr.p, idx->cnt.p, idx->overflow.p, c.level0 ? idx->sticky.p : nullptr, nq};
    if (variant != FIP_VARIANT_TENSOR) {   // (the tensor variant resets the lists inside prepare_queries_kernel)
        init_lists_kernel<<<(unsigned)((flag_words(nq) + 255) / 256), 256, 0, c.stream>>>(
            linit.thr, linit.cnt, linit.flags, linit.sticky, nq);
        FIP_CUDA(cudaGetLastError());
        c.launches++;
    }

    CUtensorMap tmap_q{};
    int nq_pad = 0, n_qtiles = 0;
    bool pair = false;
    const float* out_scale = nullptr;
    if (variant == FIP_VARIANT_TENSOR) {
        n_qtiles = (nq + tc::BM - 1) / tc::BM;
        // CTA pairs (cta_group::2) take two query tiles per corpus tile: worth it when that does
        // not pad the batch by more than ~15 %.
        pair = idx->cta_pair == 2 || (idx->cta_pair == 0 && n_qtiles >= 2 &&
                                      (n_qtiles % 2 == 0 || n_qtiles >= 7));
        if (pair) n_qtiles = (n_qtiles + 1) / 2 * 2;
        nq_pad = n_qtiles * tc::BM;
        const size_t esize = dtype_size(fdtype);
        FIP_CUDA(idx->qprep.ensure((size_t)split * nq_pad * dim * esize));
        FIP_CUDA(idx->q_scale.ensure(nq));
        FIP_CUDA(idx->q_eps.ensure(nq));
        const unsigned pgrid = (nq_pad + 7) / 8;
        // what software cannot see: fp32 accumulation in the tensor core, tf32 operand truncation
        // ... and, for the fp16 filter copy, the rounding of the stored rows: 2^-11 relative per
        // component, 2^-25 absolute for components in fp16's subnormal range
        const float rel = (float)dim * 1.1920929e-7f + (fdtype == kF32 ? (1.0f / 512 + 1.0f / 1048576) : 0.f) +
                          (copy16 ? 1.001f / 2048 + sqrtf((float)dim) * 2.9802322e-8f / idx->c16_xmax : 0.f);
        const float* mn2 = idx->max_norm2.p;
        float* sc = idx->q_scale.p;
        float* ep = idx->q_eps.p;
        if (fdtype == kBF16) {
            auto* dst = reinterpret_cast<__nv_bfloat16*>(idx->qprep.p);
            if (split == 2) tc::prepare_queries_kernel<__nv_bfloat16, 2><<<pgrid, 256, 0, c.stream>>>(q32, dst, sc, ep, mn2, rel, nq, nq_pad, dim, linit);
            else tc::prepare_queries_kernel<__nv_bfloat16, 1><<<pgrid, 256, 0, c.stream>>>(q32, dst, sc, ep, mn2, rel, nq, nq_pad, dim, linit);
        } else if (fdtype == kF16) {
            auto* dst = reinterpret_cast<__half*>(idx->qprep.p);
            if (split == 2) tc::prepare_queries_kernel<__half, 2><<<pgrid, 256, 0, c.stream>>>(q32, dst, sc, ep, mn2, rel, nq, nq_pad, dim, linit);
            else tc::prepare_queries_kernel<__half, 1><<<pgrid, 256, 0, c.stream>>>(q32, dst, sc, ep, mn2, rel, nq, nq_pad, dim, linit);
        } else {
            tc::prepare_queries_kernel<float, 1><<<pgrid, 256, 0, c.stream>>>(
                q32, reinterpret_cast<float*>(idx->qprep.p), sc, ep, mn2, rel, nq, nq_pad, dim, linit);
        }
        FIP_CUDA(cudaGetLastError());
        c.launches++;
        int rc = make_tmap(&tmap_q, idx->qprep.p, fdtype, dim, (uint64_t)split * nq_pad);
        if (rc) return rc;
        c.tmap_scan = copy16 ? &idx->tmap_x16 : &idx->tmap_x;
        if (!copy16 && !idx->tmap_x_valid) {
            rc = make_tmap(&idx->tmap_x, idx->corpus, dtype, dim, (uint64_t)n);
            if (rc) return rc;
            idx->tmap_x_valid = true;
        }
        out_scale = idx->q_scale.p;
        idx->st_pair = pair ? 2 : 1;
    }

    ScanArgs a{};
    a.corpus = idx->corpus; a.n_rows = n; a.dim = dim; a.nq = nq;
    a.dtype = variant == FIP_VARIANT_TENSOR ? fdtype : dtype;   // (the streaming kernel reads the stored rows)
    a.thr = idx->thr.p; a.cnt = idx->cnt.p; a.cand = idx->cand.p; a.cap = (uint32_t)cap;

    auto scan_launch = [&]() -> int {
        if (variant == FIP_VARIANT_TENSOR) return launch_tc_phase(c, a, tmap_q, n_qtiles, split, pair, deferred);
        switch (dtype) {
            case kF32: return launch_stream_phase<float>(c, a, q32);
            case kF16: return launch_stream_phase<__half>(c, a, q32);
            default: return launch_stream_phase<__nv_bfloat16>(c, a, q32);
        }
    };
    auto scan = [&](int64_t r0, int64_t r1, bool dense) -> int {
        a.row_begin = (uint32_t)r0; a.row_end = (uint32_t)r1; a.dense = dense ? 1 : 0;
        const bool timed = idx->profile && c.level0 && !idx->ev_sets.empty();
        const size_t nev = idx->ev.size();
        if (timed) FIP_CUDA(cudaEventRecord(idx->ev[idx->ev_head % nev], c.stream));
        const int before = c.launches;
        int r = scan_launch();
        if (r) return r;
        if (timed) {
            FIP_CUDA(cudaEventRecord(idx->ev[(idx->ev_head + 1) % nev], c.stream));
            idx->ev_head += 2;
            auto& set = idx->ev_sets.back();
            set.count += 2;
            set.launches += c.launches - before;
            set.rows.push_back(r1 - r0);
        }
        return FIP_OK;
    };

    const double ratio = phase_ratio(cap, k_plan, keep_max, live);
    if (live) {
        FIP_CUDA(idx->hist.ensure((size_t)nq * kHistBins));
        FIP_CUDA(idx->hbase.ensure(nq));
        a.hist = idx->hist.p; a.hbase = idx->hbase.p; a.k_thr = k_sel;
        if (keep_max > 0) { a.win_eps = idx->q_eps.p; a.win_scale = idx->q_scale.p; }
    }
    int phases = 0;
    // phase 0: no threshold yet, every score of the first rows is written (slot = row).  For small
    // k a sample smaller than the list capacity gives a good enough first threshold -- the filter
    // launch raises it live -- and the first select pass sorts that many keys per query instead of
    // a whole list: at batch 32 that pass is a visible part of a 0.3 ms search.
    // Lower bound: every CTA filters with the threshold it fetched up to two tiles earlier, so about
    // 2 x SMs x 256 rows are scanned against the FIRST threshold whatever the histogram does; at a
    // pass rate of k' / sample those must fit the list with room to spare (factor 2), or every query
    // overflows and is redone in safe mode (what tests/test_gpu_parity.py caught with k = 100).
    const int64_t first_wave = 2ll * idx->num_sms * tc::BN;
    const int64_t need = ((int64_t)k_plan * first_wave * 2 / cap + 255) / 256 * 256;
    const int64_t dense_rows = (!mode.safe && live) ? std::min<int64_t>(std::max<int64_t>(1024, need), cap) : cap;
    int64_t m = std::min<int64_t>(n, dense_rows);
    int rc = scan(0, m, true);
    if (rc) return rc;
    int64_t pos = m;
    phases++;
    // After the LAST scan of a certified tensor-core search ONE pass finishes the query: window on the
    // approximate scores, exact rescoring of what is inside, the k best sorted and emitted.
    const bool finish = rescore && certify && idx->fused_rescore && keep_max > 0;
    if (pos == n && finish) rc = launch_select(c, idx, nq, k_out, cap, true, out_d, out_i, nullptr, m, keep_max, false, q32);
    else rc = launch_select(c, idx, nq, k_sel, cap, pos == n && !rescore, out_d, out_i, out_scale, m, keep_max, live);
    if (rc) return rc;
    while (pos < n) {
        if (mode.safe) m = std::max<int64_t>(256, (int64_t)(cap - std::max(k_plan, keep_max)) / 256 * 256);
        else m = phase_rows(pos, ratio);
        m = std::min<int64_t>(m, n - pos);
        rc = scan(pos, pos + m, false);
        if (rc) return rc;
        if (idx->debug_counts) {   // survivor statistics (synthetic-distribution study)
            std::vector<uint32_t> h((size_t)nq);
            FIP_CUDA(cudaMemcpyAsync(h.data(), idx->cnt.p, sizeof(uint32_t) * (size_t)nq, cudaMemcpyDeviceToHost, c.stream));
            FIP_CUDA(cudaStreamSynchronize(c.stream));
            for (uint32_t v : h) { idx->st_list_entries += v; idx->st_list_max = std::max<int64_t>(idx->st_list_max, v); }
        }
        pos += m;
        phases++;
        if (pos == n && finish) rc = launch_select(c, idx, nq, k_out, cap, true, out_d, out_i, nullptr, -1, keep_max, false, q32);
        else rc = launch_select(c, idx, nq, k_sel, cap, pos == n && !rescore, out_d, out_i, out_scale, -1, keep_max, live);
        if (rc) return rc;
    }
    if (finish) {
        // done by the finishing pass above
    } else if (rescore && idx->fused_rescore) {
        // lists now hold the k' best by approximate score: the final pass recomputes them exactly
        // (fused, select.cuh) and keeps the k best.
        rc = launch_select(c, idx, nq, k_out, cap, true, out_d, out_i, nullptr, -1, 0, false, q32);
        if (rc) return rc;
    } else if (rescore) {
        // the same in two launches (A/B and bit-equality tests: "fused_rescore" = 0)
        const int k_res = certify ? keep_max : k_plan;   // list slots that may hold a candidate
        const int64_t warps = (int64_t)nq * ((k_res + 3) / 4);   // a warp rescoring 4 candidates at a time
        const unsigned grid = (unsigned)std::min<int64_t>((warps + 7) / 8, (int64_t)idx->num_sms * 8);
        if (dtype == kBF16)
            rescore_kernel<__nv_bfloat16><<<grid, 256, 0, c.stream>>>(
                idx->cand.p, idx->cnt.p, (uint32_t)cap, nq, k_res, (const __nv_bfloat16*)idx->corpus, q32, dim);
        else if (dtype == kF32)
            rescore_kernel<float><<<grid, 256, 0, c.stream>>>(
                idx->cand.p, idx->cnt.p, (uint32_t)cap, nq, k_res, (const float*)idx->corpus, q32, dim);
        else
            rescore_kernel<__half><<<grid, 256, 0, c.stream>>>(
                idx->cand.p, idx->cnt.p, (uint32_t)cap, nq, k_res, (const __half*)idx->corpus, q32, dim);
        FIP_CUDA(cudaGetLastError());
        c.launches++;
        rc = launch_select(c, idx, nq, k_out, cap, true, out_d, out_i, nullptr);
        if (rc) return rc;
    }
    idx->st_filter16 = (variant == FIP_VARIANT_TENSOR && copy16) ? 1 : 0;
    info->variant = variant; info->phases = phases; info->cap = cap; info->split = split;
    info->certified_path = certify;
    return FIP_OK;
}

// K1 in DENSE mode only: keys[q * cap + r] = raw key (score bits, row r) of <q, x_r> for every stored
// row r < n_rows <= cap, nothing else (no thresholds, no selection).  The operands are rounded to
// the index's 16-bit storage type by the tensor-core path, so this is for callers that feed values
// already representable there -- the IVF coarse quantizer's bf16 x 3 split (ivf_api.cuh), which gets
// fp32-accurate inner products out of six bf16 products laid side by side along K.  Enqueue only.
int dense_scores(fip_index* idx, const float* q32, int nq, uint64_t* keys, uint32_t cap, cudaStream_t stream,
                 int* launches) {
    const int dim = idx->cfg.dim, dtype = idx->cfg.dtype;
    const int64_t n = idx->n_rows;
    if (dtype == kF32 || (dim * 2) % tc::BK_BYTES != 0 || n > (int64_t)cap || n == 0)
        return fail(FIP_ERR_INVALID, "dense_scores: needs a 16-bit index with 128-byte-multiple rows that fits the key rows");
    SearchCtx c{idx, stream};
    int n_qtiles = (nq + tc::BM - 1) / tc::BM;
    const bool pair = n_qtiles >= 2 && (n_qtiles % 2 == 0 || n_qtiles >= 7);
    if (pair) n_qtiles = (n_qtiles + 1) / 2 * 2;
    const int nq_pad = n_qtiles * tc::BM;
    FIP_CUDA(idx->qprep.ensure((size_t)nq_pad * dim * 2));
    FIP_CUDA(idx->q_scale.ensure(nq));
    FIP_CUDA(idx->q_eps.ensure(nq));
    FIP_CUDA(idx->thr.ensure(nq));
    FIP_CUDA(idx->cnt.ensure(nq));
    FIP_CUDA(idx->overflow.ensure(std::max<size_t>(flag_words(nq), flag_words(4096))));
    const ListInit linit{idx->thr.p, idx->cnt.p, idx->overflow.p, nullptr, nq};
    const unsigned pgrid = (nq_pad + 7) / 8;
    if (dtype == kBF16)
        tc::prepare_queries_kernel<__nv_bfloat16, 1><<<pgrid, 256, 0, stream>>>(
            q32, reinterpret_cast<__nv_bfloat16*>(idx->qprep.p), idx->q_scale.p, idx->q_eps.p, idx->max_norm2.p, 0.f, nq, nq_pad, dim, linit);
    else
        return fail(FIP_ERR_INVALID, "dense_scores: bf16 storage only (fp16 queries are rescaled per query)");
    FIP_CUDA(cudaGetLastError());
    CUtensorMap tmap_q{};
    int rc = make_tmap(&tmap_q, idx->qprep.p, dtype, dim, (uint64_t)nq_pad);
    if (rc) return rc;
    if (!idx->tmap_x_valid) {
        rc = make_tmap(&idx->tmap_x, idx->corpus, dtype, dim, (uint64_t)n);
        if (rc) return rc;
        idx->tmap_x_valid = true;
    }
    c.tmap_scan = &idx->tmap_x;
    ScanArgs a{};
    a.corpus = idx->corpus; a.n_rows = n; a.dim = dim; a.nq = nq; a.dtype = dtype;
    a.thr = idx->thr.p; a.cnt = idx->cnt.p; a.cand = keys; a.cap = cap;
    a.row_begin = 0u; a.row_end = (uint32_t)n; a.dense = 1;
    rc = launch_tc_phase(c, a, tmap_q, n_qtiles, 1, pair, false);
    if (rc) return rc;
    if (launches) *launches += 1 + c.launches;
    return FIP_OK;
}

// Flags of the search just enqueued: copy_flags_async enqueues the D2H copy, parse_flags (after the
// stream has been synchronised) appends the flagged queries, translated through `ids` (null =
// identity), to `over` / `uncert`.  collect_flags = both with the synchronisation in between.
int copy_flags_async(fip_index* idx, cudaStream_t stream, int nq) {
    const size_t nflags = flag_words(nq);
    FIP_CUDA(idx->h_flags.ensure(std::max<size_t>(nflags, flag_words(4096))));
    FIP_CUDA(cudaMemcpyAsync(idx->h_flags.p, idx->overflow.p, sizeof(uint32_t) * nflags,
                             cudaMemcpyDeviceToHost, stream));
    return FIP_OK;
}
void parse_flags(const fip_index* idx, int nq, const int* ids, std::vector<int>* over, std::vector<int>* uncert) {
    const uint32_t* f = idx->h_flags.p;
    if (!f[0] && !f[1]) return;
    for (int q = 0; q < nq; ++q) {
        const int id = ids ? ids[q] : q;
        if (f[flag_slot(q)]) over->push_back(id);
        else if (f[flag_slot(q) + 1]) uncert->push_back(id);
    }
}
int collect_flags(fip_index* idx, cudaStream_t stream, int nq, const int* ids, std::vector<int>* over,
                  std::vector<int>* uncert) {
    int rc = copy_flags_async(idx, stream, nq);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(stream));
    parse_flags(idx, nq, ids, over, uncert);
    return FIP_OK;
}

// Redo the queries `sel` (indices into q32 / out_*) under `mode`; flagged ones are appended to
// over / uncert (as indices of the ORIGINAL batch).  Synchronises.
int redo_subset(SearchCtx& c, const float* q32, const std::vector<int>& sel, int k, float* out_d,
                int64_t* out_i, Mode mode, std::vector<int>* over, std::vector<int>* uncert, int* phases) {
    fip_index* idx = c.idx;
    const int ns = (int)sel.size();
    if (ns == 0) return FIP_OK;
    const int dim = idx->cfg.dim;
    FIP_CUDA(idx->sel.ensure(sel.size()));
    FIP_CUDA(idx->sub_q.ensure(sel.size() * (size_t)dim));
    FIP_CUDA(idx->sub_d.ensure(sel.size() * (size_t)k));
    FIP_CUDA(idx->sub_i.ensure(sel.size() * (size_t)k));
    FIP_CUDA(cudaMemcpyAsync(idx->sel.p, sel.data(), sel.size() * sizeof(int), cudaMemcpyHostToDevice,
                             c.stream));
    int64_t tot = (int64_t)ns * dim;
    gather_rows_f32_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, c.stream>>>(q32, idx->sel.p,
                                                                              idx->sub_q.p, ns, dim);
    FIP_CUDA(cudaGetLastError());
    c.launches++;
    SearchInfo info;
    int rc = run_search(c, idx->sub_q.p, ns, k, idx->sub_d.p, idx->sub_i.p, mode, &info);
    if (rc) return rc;
    tot = (int64_t)ns * k;
    scatter_results_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, c.stream>>>(
        idx->sub_d.p, idx->sub_i.p, idx->sel.p, out_d, out_i, ns, k);
    FIP_CUDA(cudaGetLastError());
    c.launches++;
    *phases += info.phases;
    return collect_flags(idx, c.stream, ns, sel.data(), over, uncert);  // also keeps `sel` alive
}

// A search is ENQUEUED (search_enqueue: level 0, the configured variant, no synchronisation) and
// later FINISHED (search_finish_synced, once the stream has been synchronised and the flags are on
// the host): queries whose candidate list overflowed are redone in safe mode, queries whose result
// could not be CERTIFIED exact (tensor-core filter, see select.cuh) are redone with a tighter
// filter; whatever is still flagged after that goes to the exact streaming kernel in safe mode,
// which cannot fail.  Between the two the caller may enqueue whatever consumes the result (the K4
// exchange, the D2H copies): the flags come back with the results, one synchronisation per search.
int search_enqueue(fip_index* idx, const float* q32, int nq, int k, float* out_d, int64_t* out_i,
                   cudaStream_t stream) {
    if (idx->pend.active && idx->pend.stream != stream) FIP_CUDA(cudaStreamSynchronize(idx->pend.stream));
    SearchCtx c{idx, stream};
    c.level0 = true;
    if (idx->profile && idx->ev.empty()) {
        idx->ev.resize(1024);
        for (auto& e : idx->ev) FIP_CUDA(cudaEventCreate(&e));
    }
    if (idx->profile) {
        if (idx->ev_sets.size() > 4096) idx->ev_sets.erase(idx->ev_sets.begin(), idx->ev_sets.begin() + 2048);
        idx->ev_sets.push_back(fip_index::EvSet{idx->ev_head, 0, {}, 0});
    }
    idx->st_list_entries = 0; idx->st_list_max = 0;
    Mode m0;
    m0.safe = idx->force_safe != 0;
    SearchInfo info;
    int rc = run_search(c, q32, nq, k, out_d, out_i, m0, &info);
    if (rc) return rc;
    idx->pend = fip_index::PendingSearch{q32, nq, k, out_d, out_i, stream, m0.safe, true, info.variant,
                                         info.phases, info.cap, info.split, c.launches};
    idx->st_launches = c.launches; idx->st_variant = info.variant; idx->st_phases = info.phases;
    idx->st_cap = info.cap; idx->st_split = info.split;
    idx->st_overflow = 0; idx->st_uncert = 0; idx->st_stream_redo = 0;
    idx->st_total_launches += c.launches; idx->st_total_searches++;
    return FIP_OK;
}

void collect_event_times(fip_index* idx) {
    idx->st_scan_ms = 0.0;
    idx->st_scan_launches = 0;
    idx->st_scan_searches = 0;
    idx->st_scan_each.clear();
    idx->st_gap_each.clear();
    idx->st_scan_rows.clear();
    const size_t nev = idx->ev.size();
    for (const auto& set : idx->ev_sets) {
        if (!set.count || idx->ev_head - set.start > nev) continue;   // nothing timed / ring wrapped over it
        bool ok = true;
        std::vector<float> each, gaps;
        for (int i = 0; i + 1 < set.count && ok; i += 2) {
            float ms = 0.f;
            ok = cudaEventElapsedTime(&ms, idx->ev[(set.start + i) % nev], idx->ev[(set.start + i + 1) % nev]) == cudaSuccess;
            each.push_back(ms);
            if (ok && i + 2 < set.count) {
                ok = cudaEventElapsedTime(&ms, idx->ev[(set.start + i + 1) % nev], idx->ev[(set.start + i + 2) % nev]) == cudaSuccess;
                gaps.push_back(ms);
            }
        }
        if (!ok) { cudaGetLastError(); continue; }
        if (idx->st_scan_each.size() < each.size()) idx->st_scan_each.resize(each.size(), 0.f);
        if (idx->st_gap_each.size() < gaps.size()) idx->st_gap_each.resize(gaps.size(), 0.f);
        for (size_t i = 0; i < each.size(); ++i) { idx->st_scan_each[i] += each[i]; idx->st_scan_ms += each[i]; }
        for (size_t i = 0; i < gaps.size(); ++i) idx->st_gap_each[i] += gaps[i];
        idx->st_scan_rows = set.rows;
        idx->st_scan_launches += set.launches;
        idx->st_scan_searches++;
    }
    idx->ev_sets.clear();
}

// The pending search's stream is synchronised and idx->h_flags holds its flags.  *redone = number of
// queries that were searched again (their rows of out_d / out_i were rewritten on the device).
int search_finish_synced(fip_index* idx, int* redone) {
    *redone = 0;
    if (!idx->pend.active) return FIP_OK;
    const fip_index::PendingSearch p = idx->pend;
    idx->pend.active = false;
    collect_event_times(idx);
    std::vector<int> over, uncert, over2, uncert2;
    parse_flags(idx, p.nq, nullptr, &over, &uncert);
    if (over.empty() && uncert.empty()) return FIP_OK;
    SearchCtx c{idx, p.stream};
    int phases = p.phases;
    const size_t n_over = over.size(), n_uncert = uncert.size();
    int rc;
    if (!over.empty()) {
        Mode m; m.safe = true; m.rigor = 0;
        rc = redo_subset(c, p.q32, over, p.k, p.out_d, p.out_i, m, &over2, &uncert2, &phases);
        if (rc) return rc;
    }
    if (!uncert.empty()) {
        Mode m; m.safe = p.safe; m.rigor = 1;
        rc = redo_subset(c, p.q32, uncert, p.k, p.out_d, p.out_i, m, &over2, &uncert2, &phases);
        if (rc) return rc;
    }
    over2.insert(over2.end(), uncert2.begin(), uncert2.end());
    if (!over2.empty()) {
        std::sort(over2.begin(), over2.end());
        Mode m; m.safe = true; m.rigor = 2;
        std::vector<int> o3, u3;
        rc = redo_subset(c, p.q32, over2, p.k, p.out_d, p.out_i, m, &o3, &u3, &phases);
        if (rc) return rc;
        if (!o3.empty() || !u3.empty())
            return fail(FIP_ERR_STATE, "internal error: the exact fallback path raised a flag");
    }
    // everything flagged has been dealt with: the summary words must not reach `sticky`
    FIP_CUDA(cudaMemsetAsync(idx->overflow.p, 0, 2 * sizeof(uint32_t), p.stream));
    FIP_CUDA(cudaStreamSynchronize(p.stream));
    *redone = (int)(n_over + n_uncert);
    idx->st_launches = p.launches + c.launches; idx->st_phases = phases;
    idx->st_overflow = (int64_t)n_over; idx->st_uncert = (int64_t)n_uncert;
    idx->st_stream_redo = (int64_t)over2.size();
    idx->st_total_launches += c.launches;
    return FIP_OK;
}

// Blocking search: enqueue, fetch the flags, finish.
int search_checked(fip_index* idx, const float* q32, int nq, int k, float* out_d, int64_t* out_i,
                   cudaStream_t stream) {
    int rc = search_enqueue(idx, q32, nq, k, out_d, out_i, stream);
    if (rc) return rc;
    rc = copy_flags_async(idx, stream, nq);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(stream));
    int redone = 0;
    return search_finish_synced(idx, &redone);
}

int check_search_args(const fip_index* idx, const void* q, int64_t nq, int32_t k, const void* d,
                      const void* i) {
    if (!idx) return fail(FIP_ERR_INVALID, "null index");
    if (nq < 0 || nq > (1 << 20)) return fail(FIP_ERR_INVALID, "nq out of range");
    if (k < 1 || k > kMaxK) return fail(FIP_ERR_INVALID, "k must be in [1, 2048]");
    if (nq > 0 && (!q || !d || !i)) return fail(FIP_ERR_INVALID, "null buffer");
    return FIP_OK;
}

}  // namespace

// ================================================================================= C ABI
extern "C" {

int fip_create(const fip_config* cfg, fip_index** out) {
    if (!cfg || !out) return fail(FIP_ERR_INVALID, "null argument");
    *out = nullptr;
    if (cfg->dim <= 0) return fail(FIP_ERR_INVALID, "dim must be positive");
    if (cfg->dtype != FIP_F32 && cfg->dtype != FIP_F16 && cfg->dtype != FIP_BF16)
        return fail(FIP_ERR_INVALID, "unknown dtype");
    if ((cfg->dim * dtype_size(cfg->dtype)) % 16 != 0)
        return fail(FIP_ERR_INVALID, "row size must be a multiple of 16 bytes "
                                     "(dim % 4 == 0 for fp32, dim % 8 == 0 for fp16/bf16)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(FIP_ERR_NODEVICE, "no CUDA device visible; this library has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(FIP_ERR_INVALID, "bad device ordinal");
    cudaDeviceProp prop;
    FIP_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return fail(FIP_ERR_NODEVICE, std::string("device '") + prop.name +
                                          "' is not sm_100 (B200); kernels are built for sm_100a only");
    FIP_CUDA(cudaSetDevice(cfg->device));
    fip_index* idx = new (std::nothrow) fip_index();
    if (!idx) return fail(FIP_ERR_NOMEM, "host allocation failed");
    idx->cfg = *cfg;
    idx->num_sms = prop.multiProcessorCount;
    cudaError_t e = cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete idx; FIP_CUDA(e); }
    e = idx->max_norm2.ensure(1);
    if (e == cudaSuccess) e = cudaMemset(idx->max_norm2.p, 0, sizeof(float));
    if (e == cudaSuccess) e = idx->sticky.ensure(1);
    if (e == cudaSuccess) e = cudaMemset(idx->sticky.p, 0, sizeof(uint32_t));
    if (e != cudaSuccess) { fip_destroy(idx); FIP_CUDA(e); }
    if (cfg->capacity_rows > 0) {
        int rc = reserve_rows(idx, cfg->capacity_rows);
        if (rc) { fip_destroy(idx); return rc; }
    }
    *out = idx;
    return FIP_OK;
}

void fip_destroy(fip_index* idx) {
    if (!idx) return;
    cudaSetDevice(idx->cfg.device);
    if (idx->stream) cudaStreamSynchronize(idx->stream);
    if (idx->corpus) cudaFree(idx->corpus);
    if (idx->corpus16) cudaFree(idx->corpus16);
    idx->sticky.release(); idx->thr.release(); idx->q_scale.release(); idx->q_eps.release(); idx->max_norm2.release(); idx->q32.release(); idx->out_d.release();
    idx->cnt.release(); idx->hist.release(); idx->hbase.release(); idx->overflow.release(); idx->cand.release(); idx->qprep.release();
    idx->out_i.release(); idx->sel.release(); idx->sub_q.release(); idx->sub_d.release();
    idx->sub_i.release(); idx->h_q.release(); idx->h_d.release(); idx->h_i.release();
    idx->h_flags.release(); idx->add_stage.release(); idx->h_add[0].release(); idx->h_add[1].release();
    idx->dd_hash.release(); idx->dd_rep.release(); idx->dd_pos.release(); idx->dd_sel.release(); idx->dd_map.release();
    idx->dd_count.release(); idx->h_dd_count.release(); idx->dd_q.release(); idx->dd_d.release(); idx->dd_i.release();
    for (auto& e : idx->ev) cudaEventDestroy(e);
    if (idx->stream) cudaStreamDestroy(idx->stream);
    delete idx;
}

// Folds the squared norms of stored rows [row0, row0 + n) into idx->max_norm2 (enqueue only).
static int update_row_norms(fip_index* idx, int64_t row0, int64_t n) {
    const char* src = (const char*)idx->corpus + (size_t)row0 * idx->row_bytes();
    const unsigned grid = (unsigned)std::min<int64_t>((n + 7) / 8, (int64_t)idx->num_sms * 16);
    switch (idx->cfg.dtype) {
        case kF32: row_norm_max_kernel<float><<<grid, 256, 0, idx->stream>>>((const float*)src, n, idx->cfg.dim, idx->max_norm2.p); break;
        case kF16: row_norm_max_kernel<__half><<<grid, 256, 0, idx->stream>>>((const __half*)src, n, idx->cfg.dim, idx->max_norm2.p); break;
        default: row_norm_max_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>((const __nv_bfloat16*)src, n, idx->cfg.dim, idx->max_norm2.p); break;
    }
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

static int convert_append(fip_index* idx, const float* src_dev, int64_t n) {
    const int64_t tot = n * idx->cfg.dim;
    char* dst = (char*)idx->corpus + (size_t)idx->n_rows * idx->row_bytes();
    unsigned grid = (unsigned)std::min<int64_t>((tot + 255) / 256, 148 * 32);
    switch (idx->cfg.dtype) {
        case kF32:
            FIP_CUDA(cudaMemcpyAsync(dst, src_dev, (size_t)tot * 4, cudaMemcpyDeviceToDevice, idx->stream));
            break;
        case kF16:
            convert_rows_kernel<__half><<<grid, 256, 0, idx->stream>>>(src_dev, (__half*)dst, tot);
            break;
        default:
            convert_rows_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>(src_dev, (__nv_bfloat16*)dst, tot);
            break;
    }
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

// Host rows in `dtype` -> the index.  Rows travel in their OWN width (a 16-bit corpus file is
// uploaded as 16 bits, not widened to fp32 on the host) through two pinned staging buffers, so the
// CPU copy of chunk c+1 overlaps the H2D transfer and conversion of chunk c.
static int add_host_rows(fip_index* idx, const void* rows, int64_t n, int32_t dtype) {
    const int sdt = idx->cfg.dtype;
    if (dtype != FIP_F32 && dtype != FIP_F16 && dtype != FIP_BF16) return fail(FIP_ERR_INVALID, "unknown dtype");
    if (dtype != sdt && dtype != FIP_F32 && sdt != FIP_F32)
        return fail(FIP_ERR_INVALID, "fip_add_host: fp16 <-> bf16 conversion is not supported; pass fp32 "
                                     "rows or rows in the index's storage dtype");
    if (idx->n_rows + n >= (int64_t)0xFFFFFF00ll) return fail(FIP_ERR_INVALID, "shard limited to 2^32-256 rows");
    int rc = reserve_rows(idx, idx->n_rows + n);
    if (rc) return rc;
    const int dim = idx->cfg.dim;
    const size_t src_row = (size_t)dim * dtype_size(dtype);
    const int64_t chunk_rows = std::max<int64_t>(1, (int64_t)(32 << 20) / (int64_t)src_row);
    const size_t chunk_bytes = (size_t)std::min(chunk_rows, n) * src_row;
    FIP_CUDA(idx->h_add[0].ensure(chunk_bytes));
    FIP_CUDA(idx->h_add[1].ensure(chunk_bytes));
    const bool direct = dtype == sdt;
    if (!direct) FIP_CUDA(idx->add_stage.ensure(2 * ((chunk_bytes + 3) / 4)));
    cudaEvent_t ev[2] = {nullptr, nullptr};
    for (auto& e : ev) FIP_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    auto cleanup = [&]() { for (auto& e : ev) if (e) cudaEventDestroy(e); };
    int64_t done = 0;
    for (int64_t r = 0, c = 0; r < n; r += chunk_rows, ++c) {
        const int64_t m = std::min(chunk_rows, n - r);
        const int b = (int)(c & 1);
        cudaError_t e = c >= 2 ? cudaEventSynchronize(ev[b]) : cudaSuccess;   // staging buffer b is free again
        if (e == cudaSuccess) {
            memcpy(idx->h_add[b].p, (const char*)rows + (size_t)r * src_row, (size_t)m * src_row);
            char* dst = (char*)idx->corpus + (size_t)(idx->n_rows + done) * idx->row_bytes();
            void* stage = direct ? (void*)dst : (void*)((char*)idx->add_stage.p + (size_t)b * ((chunk_bytes + 3) / 4) * 4);
            e = cudaMemcpyAsync(stage, idx->h_add[b].p, (size_t)m * src_row, cudaMemcpyHostToDevice, idx->stream);
            if (e == cudaSuccess && !direct) {
                const int64_t tot = m * dim;
                const unsigned grid = (unsigned)std::min<int64_t>((tot + 255) / 256, 148 * 32);
                if (dtype == FIP_F32 && sdt == kF16) convert_rows_kernel<__half><<<grid, 256, 0, idx->stream>>>((const float*)stage, (__half*)dst, tot);
                else if (dtype == FIP_F32) convert_rows_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>((const float*)stage, (__nv_bfloat16*)dst, tot);
                else if (dtype == FIP_F16) widen_rows_kernel<__half><<<grid, 256, 0, idx->stream>>>((const __half*)stage, (float*)dst, tot);
                else widen_rows_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>((const __nv_bfloat16*)stage, (float*)dst, tot);
                e = cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaEventRecord(ev[b], idx->stream);
        }
        if (e != cudaSuccess) { cudaStreamSynchronize(idx->stream); cleanup(); FIP_CUDA(e); }
        rc = update_row_norms(idx, idx->n_rows + done, m);
        if (rc) { cudaStreamSynchronize(idx->stream); cleanup(); return rc; }
        done += m;
    }
    cudaError_t e = cudaStreamSynchronize(idx->stream);
    cleanup();
    FIP_CUDA(e);
    // rows become visible only once every chunk has landed; whatever reads the corpus through a
    // cached view (tensor map, fp16 filter copy) is invalidated together with the row count
    idx->n_rows += n;
    idx->tmap_x_valid = false; idx->corpus_gen++;
    return FIP_OK;
}

int fip_add_host(fip_index* idx, const void* rows, int64_t n, int32_t dtype) {
    if (!idx || (n > 0 && !rows) || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    return add_host_rows(idx, rows, n, dtype);
}

int fip_add_host_f32(fip_index* idx, const float* rows, int64_t n) {
    return fip_add_host(idx, rows, n, FIP_F32);
}

int fip_add_device(fip_index* idx, const void* rows_dev, int64_t n, int32_t dtype) {
    if (!idx || (n > 0 && !rows_dev) || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    if (idx->n_rows + n >= (int64_t)0xFFFFFF00ll) return fail(FIP_ERR_INVALID, "shard limited to 2^32-256 rows");
    rc = reserve_rows(idx, idx->n_rows + n);
    if (rc) return rc;
    if (dtype == idx->cfg.dtype) {
        char* dst = (char*)idx->corpus + (size_t)idx->n_rows * idx->row_bytes();
        FIP_CUDA(cudaMemcpyAsync(dst, rows_dev, (size_t)n * idx->row_bytes(), cudaMemcpyDeviceToDevice, idx->stream));
    } else if (dtype == FIP_F32) {
        rc = convert_append(idx, (const float*)rows_dev, n);
        if (rc) return rc;
    } else {
        return fail(FIP_ERR_INVALID, "fip_add_device: rows must be fp32 or the index's storage dtype");
    }
    rc = update_row_norms(idx, idx->n_rows, n);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(idx->stream));
    idx->n_rows += n;
    idx->tmap_x_valid = false; idx->corpus_gen++;
    return FIP_OK;
}

int fip_add_synthetic(fip_index* idx, int64_t n, uint64_t seed, int64_t global_row0) {
    if (!idx || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    if (idx->n_rows + n >= (int64_t)0xFFFFFF00ll) return fail(FIP_ERR_INVALID, "shard limited to 2^32-256 rows");
    rc = reserve_rows(idx, idx->n_rows + n);
    if (rc) return rc;
    char* dst = (char*)idx->corpus + (size_t)idx->n_rows * idx->row_bytes();
    const int grid = idx->num_sms * 16;
    switch (idx->cfg.dtype) {
        case kF32: synthetic_fill_kernel<float><<<grid, 256, 0, idx->stream>>>((float*)dst, n, idx->cfg.dim, global_row0, seed, 0u); break;
        case kF16: synthetic_fill_kernel<__half><<<grid, 256, 0, idx->stream>>>((__half*)dst, n, idx->cfg.dim, global_row0, seed, 0u); break;
        default: synthetic_fill_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>((__nv_bfloat16*)dst, n, idx->cfg.dim, global_row0, seed, 0u); break;
    }
    FIP_CUDA(cudaGetLastError());
    rc = update_row_norms(idx, idx->n_rows, n);
    if (rc) return rc;
    FIP_CUDA(cudaStreamSynchronize(idx->stream));
    idx->n_rows += n;
    idx->tmap_x_valid = false; idx->corpus_gen++;
    return FIP_OK;
}

int fip_reset(fip_index* idx) {
    if (!idx) return fail(FIP_ERR_INVALID, "null index");
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    FIP_CUDA(cudaMemsetAsync(idx->max_norm2.p, 0, sizeof(float), idx->stream));
    FIP_CUDA(cudaStreamSynchronize(idx->stream));
    idx->n_rows = 0;
    idx->tmap_x_valid = false; idx->corpus_gen++;
    return FIP_OK;
}

int fip_reconstruct(fip_index* idx, int64_t row0, int64_t n, float* out_host) {
    if (!idx || !out_host || row0 < 0 || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    std::lock_guard<std::mutex> lock(idx->mu);
    if (row0 + n > idx->n_rows) return fail(FIP_ERR_INVALID, "row range out of bounds");
    if (n == 0) return FIP_OK;
    int rc = use_device(idx);
    if (rc) return rc;
    const int64_t tot = n * idx->cfg.dim;
    const char* src = (const char*)idx->corpus + (size_t)row0 * idx->row_bytes();
    if (idx->cfg.dtype == kF32) {
        FIP_CUDA(cudaMemcpyAsync(out_host, src, (size_t)tot * 4, cudaMemcpyDeviceToHost, idx->stream));
    } else {
        FIP_CUDA(idx->add_stage.ensure((size_t)tot));
        unsigned grid = (unsigned)std::min<int64_t>((tot + 255) / 256, 148 * 32);
        if (idx->cfg.dtype == kF16)
            widen_rows_kernel<__half><<<grid, 256, 0, idx->stream>>>((const __half*)src, idx->add_stage.p, tot);
        else
            widen_rows_kernel<__nv_bfloat16><<<grid, 256, 0, idx->stream>>>((const __nv_bfloat16*)src, idx->add_stage.p, tot);
        FIP_CUDA(cudaGetLastError());
        FIP_CUDA(cudaMemcpyAsync(out_host, idx->add_stage.p, (size_t)tot * 4, cudaMemcpyDeviceToHost, idx->stream));
    }
    FIP_CUDA(cudaStreamSynchronize(idx->stream));
    return FIP_OK;
}

int64_t fip_ntotal(const fip_index* idx) { return idx ? idx->n_rows : 0; }
int32_t fip_dim(const fip_index* idx) { return idx ? idx->cfg.dim : 0; }
int32_t fip_dtype_of(const fip_index* idx) { return idx ? idx->cfg.dtype : -1; }

// One host-buffer search over the concatenation of `batch` (all with the same k): pinned staging,
// H2D, search, D2H, scatter back to every caller's buffers.
// Duplicate elimination of a device-resident query batch (dedupe.cuh).  Enqueues the kernels, reads
// the number of distinct rows back (one synchronisation) and, if there are duplicates, gathers the
// distinct rows into idx->dd_q: *nd < nq then, and idx->dd_map maps every row to its representative.
static int dedupe_queries(fip_index* idx, const float* q32, int nq, cudaStream_t stream, int* nd) {
    *nd = nq;
    if (nq < 2) return FIP_OK;
    const int dim = idx->cfg.dim;
    FIP_CUDA(idx->dd_hash.ensure(nq));
    FIP_CUDA(idx->dd_rep.ensure(nq));
    FIP_CUDA(idx->dd_pos.ensure(nq));
    FIP_CUDA(idx->dd_sel.ensure(nq));
    FIP_CUDA(idx->dd_map.ensure(nq));
    FIP_CUDA(idx->dd_count.ensure(1));
    FIP_CUDA(idx->h_dd_count.ensure(1));
    const unsigned grid = (unsigned)((nq + 7) / 8);
    query_hash_kernel<<<grid, 256, 0, stream>>>(q32, nq, dim, idx->dd_hash.p);
    query_rep_kernel<<<grid, 256, 0, stream>>>(q32, idx->dd_hash.p, nq, dim, idx->dd_rep.p);
    query_compact_kernel<<<1, 1024, 0, stream>>>(idx->dd_rep.p, nq, idx->dd_pos.p, idx->dd_sel.p, idx->dd_map.p,
                                                 idx->dd_count.p);
    FIP_CUDA(cudaGetLastError());
    FIP_CUDA(cudaMemcpyAsync(idx->h_dd_count.p, idx->dd_count.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    FIP_CUDA(cudaStreamSynchronize(stream));
    *nd = idx->h_dd_count.p[0];
    if (*nd >= nq) { *nd = nq; return FIP_OK; }
    FIP_CUDA(idx->dd_q.ensure((size_t)*nd * dim));
    const int64_t tot = (int64_t)*nd * dim;
    gather_rows_f32_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(q32, idx->dd_sel.p, idx->dd_q.p, *nd, dim);
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}
static int expand_results(fip_index* idx, const float* sd, const int64_t* si, float* dd, int64_t* di, int nq,
                          int k, cudaStream_t stream) {
    const int64_t tot = (int64_t)nq * k;
    expand_results_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, stream>>>(sd, si, idx->dd_map.p, dd, di, nq, k);
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

static int run_host_batch(fip_index* idx, const std::vector<HostRequest*>& batch) {
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    const int k = batch[0]->k;
    int64_t nq = 0;
    for (auto* b : batch) nq += b->nq;
    const size_t dim = (size_t)idx->cfg.dim;
    const size_t qn = (size_t)nq * dim, on = (size_t)nq * k;
    FIP_CUDA(idx->h_q.ensure(qn));
    FIP_CUDA(idx->h_d.ensure(on));
    FIP_CUDA(idx->h_i.ensure(on));
    FIP_CUDA(idx->q32.ensure(qn));
    FIP_CUDA(idx->out_d.ensure(on));
    FIP_CUDA(idx->out_i.ensure(on));
    size_t off = 0;
    for (auto* b : batch) {
        memcpy(idx->h_q.p + off * dim, b->q, (size_t)b->nq * dim * sizeof(float));
        off += (size_t)b->nq;
    }
    FIP_CUDA(cudaMemcpyAsync(idx->q32.p, idx->h_q.p, qn * sizeof(float), cudaMemcpyHostToDevice, idx->stream));
    // bit-identical queries (of different callers, mostly) are searched once
    int nd = (int)nq;
    if (idx->dedupe == 2 || (idx->dedupe == 1 && batch.size() > 1)) {
        rc = dedupe_queries(idx, idx->q32.p, (int)nq, idx->stream, &nd);
        if (rc) return rc;
    }
    const bool dup = nd < (int)nq;
    idx->co.st_queries += nq; idx->co.st_distinct += nd;
    const float* q_use = dup ? idx->dd_q.p : idx->q32.p;
    float* d_use = idx->out_d.p;
    int64_t* i_use = idx->out_i.p;
    if (dup) {
        FIP_CUDA(idx->dd_d.ensure((size_t)nd * k));
        FIP_CUDA(idx->dd_i.ensure((size_t)nd * k));
        d_use = idx->dd_d.p; i_use = idx->dd_i.p;
    }
    // results and flags come back together: one synchronisation per search; only a flagged search
    // (rare: overflow / uncertified queries) is finished on the host and copied again
    rc = search_enqueue(idx, q_use, nd, k, d_use, i_use, idx->stream);
    if (rc) return rc;
    for (int pass = 0; pass < 2; ++pass) {
        if (dup) { rc = expand_results(idx, d_use, i_use, idx->out_d.p, idx->out_i.p, (int)nq, k, idx->stream); if (rc) return rc; }
        FIP_CUDA(cudaMemcpyAsync(idx->h_d.p, idx->out_d.p, on * sizeof(float), cudaMemcpyDeviceToHost, idx->stream));
        FIP_CUDA(cudaMemcpyAsync(idx->h_i.p, idx->out_i.p, on * sizeof(int64_t), cudaMemcpyDeviceToHost, idx->stream));
        if (pass == 0) { rc = copy_flags_async(idx, idx->stream, nd); if (rc) return rc; }
        FIP_CUDA(cudaStreamSynchronize(idx->stream));
        if (pass == 1) break;
        int redone = 0;
        rc = search_finish_synced(idx, &redone);
        if (rc) return rc;
        if (!redone) break;
    }
    off = 0;
    for (auto* b : batch) {
        memcpy(b->d, idx->h_d.p + off * k, (size_t)b->nq * k * sizeof(float));
        memcpy(b->i, idx->h_i.p + off * k, (size_t)b->nq * k * sizeof(int64_t));
        off += (size_t)b->nq;
    }
    idx->co.count(batch.size());
    return FIP_OK;
}

int fip_search(fip_index* idx, const float* queries, int64_t nq, int32_t k, float* distances,
               int64_t* ids) {
    int rc = check_search_args(idx, queries, nq, k, distances, ids);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    HostRequest me{queries, nq, k, distances, ids};
    return idx->co.submit(me, [idx](const std::vector<HostRequest*>& batch) { return run_host_batch(idx, batch); });
}

int fip_search_device(fip_index* idx, const float* queries_dev, int64_t nq, int32_t k,
                      float* distances_dev, int64_t* ids_dev, void* cuda_stream) {
    int rc = check_search_args(idx, queries_dev, nq, k, distances_dev, ids_dev);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(idx->mu);
    rc = use_device(idx);
    if (rc) return rc;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : idx->stream;
    return search_checked(idx, queries_dev, (int)nq, k, distances_dev, ids_dev, s);
}

int fip_search_device_async(fip_index* idx, const float* queries_dev, int64_t nq, int32_t k,
                            float* distances_dev, int64_t* ids_dev, void* cuda_stream) {
    int rc = check_search_args(idx, queries_dev, nq, k, distances_dev, ids_dev);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(idx->mu);
    rc = use_device(idx);
    if (rc) return rc;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : idx->stream;
    return search_enqueue(idx, queries_dev, (int)nq, k, distances_dev, ids_dev, s);
}

int fip_search_status(fip_index* idx, int32_t* flagged) {
    if (!idx || !flagged) return fail(FIP_ERR_INVALID, "null argument");
    *flagged = 0;
    std::lock_guard<std::mutex> lock(idx->mu);
    int rc = use_device(idx);
    if (rc) return rc;
    cudaStream_t s = idx->pend.stream ? idx->pend.stream : idx->stream;
    if (!idx->overflow.p) return FIP_OK;   // never searched
    FIP_CUDA(idx->h_flags.ensure(flag_words(4096)));
    FIP_CUDA(cudaMemcpyAsync(idx->h_flags.p, idx->overflow.p, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    FIP_CUDA(cudaMemcpyAsync(idx->h_flags.p + 2, idx->sticky.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    FIP_CUDA(cudaStreamSynchronize(s));
    const uint32_t* f = idx->h_flags.p;
    *flagged = (f[0] ? 1 : 0) | (f[1] ? 2 : 0) | (int32_t)f[2];
    if (*flagged) {
        FIP_CUDA(cudaMemsetAsync(idx->overflow.p, 0, 2 * sizeof(uint32_t), s));
        FIP_CUDA(cudaMemsetAsync(idx->sticky.p, 0, sizeof(uint32_t), s));
        FIP_CUDA(cudaStreamSynchronize(s));
    }
    idx->pend.active = false;
    collect_event_times(idx);
    return FIP_OK;
}

const uint32_t* fip_search_flags_device(const fip_index* idx) { return idx ? idx->overflow.p : nullptr; }

int fip_merge_strided_device(int32_t device, const float* in_d, const int64_t* in_i, int64_t stride_d,
                             int64_t stride_i, int32_t g, int64_t nq, int32_t k_in, int32_t k_out,
                             float* out_d, int64_t* out_i, void* cuda_stream) {
    if (!in_d || !in_i || !out_d || !out_i || g < 1 || nq < 0 || k_in < 1 || k_out < 1)
        return fail(FIP_ERR_INVALID, "bad argument");
    if ((int64_t)g * k_in > kMaxCap) return fail(FIP_ERR_INVALID, "g * k_in must be <= 16384");
    if (stride_d < nq * k_in || stride_i < nq * k_in)
        return fail(FIP_ERR_INVALID, "shard stride smaller than one shard's result");
    if (nq == 0) return FIP_OK;
    FIP_CUDA(cudaSetDevice(device));
    MergeArgs m{in_d, in_i, g, (int)nq, k_in, k_out, stride_d, stride_i, out_d, out_i};
    int n = next_pow2(g * k_in);
    size_t smem = (size_t)n * sizeof(uint64_t);
    FIP_CUDA(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 1024)));
    int threads = std::min(1024, std::max(32, n / 2));
    merge_topk_kernel<<<(unsigned)nq, threads, smem, (cudaStream_t)cuda_stream>>>(m);
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

int fip_merge_device(int32_t device, const float* in_d, const int64_t* in_i, int32_t g, int64_t nq,
                     int32_t k_in, int32_t k_out, float* out_d, int64_t* out_i, void* cuda_stream) {
    return fip_merge_strided_device(device, in_d, in_i, nq * (int64_t)k_in, nq * (int64_t)k_in, g, nq, k_in,
                                    k_out, out_d, out_i, cuda_stream);
}

// ------------------------------------------------------------------ fused exchange + merge (K4)
struct fip_exchange {
    int device = 0, world = 1, rank = 0, nq_max = 0, k_max = 0;
    int64_t slot_bytes = 0, flags_off = 0, status_off = 0, total_bytes = 0;
    int resident_ctas[16] = {};   // per log2(merged list length): co-resident CTAs of the kernel (0 = not queried)
    int share = 1;                // ranks of the group living on THIS device (tests): they split the residency
    uint8_t* local = nullptr;
    uint8_t* peer[kMaxRanks] = {};
    bool ipc_opened[kMaxRanks] = {};
    bool connected = false;
    uint32_t epoch = 0;
    std::mutex mu;
};

static int64_t round_up16(int64_t v) { return (v + 15) / 16 * 16; }

int fip_exchange_create(int32_t device, int32_t world, int32_t rank, int64_t nq_max, int32_t k_max,
                        fip_exchange** out) {
    if (!out) return fail(FIP_ERR_INVALID, "null argument");
    *out = nullptr;
    if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world)
        return fail(FIP_ERR_INVALID, "world must be in [1, 16] and 0 <= rank < world");
    if (nq_max < 1 || nq_max > (1 << 20) || k_max < 1 || k_max > kMaxK)
        return fail(FIP_ERR_INVALID, "nq_max / k_max out of range");
    if ((int64_t)world * k_max > kMaxCap) return fail(FIP_ERR_INVALID, "world * k_max must be <= 16384");
    FIP_CUDA(cudaSetDevice(device));
    fip_exchange* x = new (std::nothrow) fip_exchange();
    if (!x) return fail(FIP_ERR_NOMEM, "host allocation failed");
    x->device = device; x->world = world; x->rank = rank; x->nq_max = (int)nq_max; x->k_max = k_max;
    x->slot_bytes = round_up16(round_up16(nq_max * k_max * 4) + nq_max * k_max * 8);
    x->flags_off = 2 * (int64_t)world * x->slot_bytes;
    x->status_off = x->flags_off + round_up16(2 * (int64_t)world * nq_max * 4);
    x->total_bytes = x->status_off + round_up16(2 * (int64_t)world * 4);
    cudaError_t e = cudaMalloc((void**)&x->local, (size_t)x->total_bytes);
    if (e == cudaSuccess) e = cudaMemset(x->local, 0, (size_t)x->total_bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { if (x->local) cudaFree(x->local); delete x; FIP_CUDA(e); }
    x->peer[rank] = x->local;
    *out = x;
    return FIP_OK;
}

int fip_exchange_ipc_handle(fip_exchange* x, void* handle_out) {
    if (!x || !handle_out) return fail(FIP_ERR_INVALID, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == FIP_IPC_HANDLE_BYTES, "handle size");
    FIP_CUDA(cudaSetDevice(x->device));
    cudaIpcMemHandle_t h;
    FIP_CUDA(cudaIpcGetMemHandle(&h, x->local));
    memcpy(handle_out, &h, sizeof(h));
    return FIP_OK;
}

int fip_exchange_connect_ipc(fip_exchange* x, const void* all_handles) {
    if (!x || !all_handles) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lock(x->mu);
    FIP_CUDA(cudaSetDevice(x->device));
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank || x->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)all_handles + (size_t)r * sizeof(h), sizeof(h));
        void* p = nullptr;
        FIP_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        x->peer[r] = (uint8_t*)p;
        x->ipc_opened[r] = true;
    }
    x->connected = true;
    return FIP_OK;
}

int fip_exchange_connect_local(fip_exchange* x, fip_exchange* const* all) {
    if (!x || !all) return fail(FIP_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lock(x->mu);
    FIP_CUDA(cudaSetDevice(x->device));
    x->share = 1;
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) continue;
        const fip_exchange* o = all[r];
        if (!o || o->world != x->world || o->rank != r || o->slot_bytes != x->slot_bytes || o->nq_max != x->nq_max)
            return fail(FIP_ERR_INVALID, "exchange objects do not describe the same group");
        if (o->device != x->device) {
            int can = 0;
            FIP_CUDA(cudaDeviceCanAccessPeer(&can, x->device, o->device));
            if (!can) return fail(FIP_ERR_STATE, "no peer access between the two devices");
            cudaError_t e = cudaDeviceEnablePeerAccess(o->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) FIP_CUDA(e);
            cudaGetLastError();
        }
        x->peer[r] = o->local;
        if (o->device == x->device) x->share++;
    }
    x->connected = true;
    return FIP_OK;
}

int fip_exchange_merge(fip_exchange* x, const float* d_loc, const int64_t* i_loc, int64_t nq, int32_t k,
                       float* out_d, int64_t* out_i, void* cuda_stream) {
    return fip_exchange_merge_status(x, d_loc, i_loc, nq, k, out_d, out_i, nullptr, nullptr, cuda_stream);
}

// phase: ExchangeArgs::phase.  0 and 1 open a new epoch, 2 finishes the one its phase-1 launch opened.
static int exchange_launch(fip_exchange* x, const float* d_loc, const int64_t* i_loc, int64_t nq, int32_t k,
                           float* out_d, int64_t* out_i, const uint32_t* flags_loc_dev,
                           uint32_t* out_status_dev, void* cuda_stream, int phase) {
    if (!x || (nq > 0 && (!d_loc || !i_loc || !out_d || !out_i))) return fail(FIP_ERR_INVALID, "null argument");
    if (nq < 0 || nq > x->nq_max || k < 1 || k > x->k_max)
        return fail(FIP_ERR_INVALID, "nq / k exceed what the exchange was created for");
    if (!x->connected && x->world > 1) return fail(FIP_ERR_STATE, "exchange is not connected to its peers");
    if (nq == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(x->mu);
    // every rank makes the same sequence of calls (like a collective), so the epoch advances in
    // lock-step and consecutive launches alternate buffer parity
    // (committed only once the kernel has been launched: a call that fails before that must not
    // leave this rank one epoch ahead of its peers)
    uint32_t epoch = phase == 2 ? x->epoch : x->epoch + 1;
    if (epoch == 0) epoch = 2;  // wrapped (0xFFFFFFFF -> 2 keeps the parity alternating); 0 = "never written"
    FIP_CUDA(cudaSetDevice(x->device));
    ExchangeArgs a{};
    for (int r = 0; r < x->world; ++r) a.peer[r] = x->peer[r];
    a.g = x->world; a.rank = x->rank; a.nq = (int)nq; a.k = k; a.epoch = epoch;
    a.slot_bytes = x->slot_bytes; a.flags_off = x->flags_off; a.nq_max = x->nq_max;
    a.off_i = (int)round_up16(nq * k * 4);
    a.d_loc = d_loc; a.i_loc = i_loc; a.out_d = out_d; a.out_i = out_i;
    a.status_off = x->status_off; a.flags_loc = flags_loc_dev; a.out_status = out_status_dev;
    a.phase = phase;
    const int n = next_pow2(x->world * k);
    const size_t smem = (size_t)n * sizeof(uint64_t);
    FIP_CUDA(cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)std::max<size_t>(smem, 1024)));
    const int threads = std::min(1024, std::max(64, n / 2));
    // grid bounded by what is resident at once (every CTA loops over its queries)
    // (cached per list length n: n fixes both the CTA size and its shared memory, and with 1024
    // threads the shared memory alone decides between 2 CTAs per SM and 1)
    int lg = 0;
    while ((1 << lg) < n) lg++;
    if (!x->resident_ctas[lg]) {
        int per_sm = 0, sms = 0;
        FIP_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, exchange_merge_kernel, threads, smem));
        FIP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, x->device));
        x->resident_ctas[lg] = std::max(1, per_sm * sms);
    }
    const unsigned grid = (unsigned)std::min<int64_t>(nq, std::max(1, x->resident_ctas[lg] / x->share));
    exchange_merge_kernel<<<grid, threads, smem, (cudaStream_t)cuda_stream>>>(a);
    FIP_CUDA(cudaGetLastError());
    x->epoch = epoch;
    return FIP_OK;
}

int fip_exchange_merge_status(fip_exchange* x, const float* d_loc, const int64_t* i_loc, int64_t nq, int32_t k,
                              float* out_d, int64_t* out_i, const uint32_t* flags_loc_dev,
                              uint32_t* out_status_dev, void* cuda_stream) {
    return exchange_launch(x, d_loc, i_loc, nq, k, out_d, out_i, flags_loc_dev, out_status_dev, cuda_stream, 0);
}

void fip_exchange_destroy(fip_exchange* x) {
    if (!x) return;
    cudaSetDevice(x->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < x->world; ++r)
        if (x->ipc_opened[r] && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
    if (x->local) cudaFree(x->local);
    cudaGetLastError();
    delete x;
}

int fip_set_option(fip_index* idx, const char* key, int64_t value) {
    if (!idx || !key) return fail(FIP_ERR_INVALID, "bad argument");
    std::lock_guard<std::mutex> lock(idx->mu);
    std::string k(key);
    if (k == "variant") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "variant must be 0, 1 or 2");
        idx->variant = (int)value;
    } else if (k == "stream_max_batch") {
        idx->stream_max_batch = (int)std::max<int64_t>(0, value);
    } else if (k == "query_split") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "query_split must be 0 (auto), 1 or 2");
        idx->query_split = (int)value;
    } else if (k == "cta_pair") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "cta_pair must be 0 (auto), 1 or 2");
        idx->cta_pair = (int)value;
    } else if (k == "coalesce") {
        idx->co.enabled = value != 0;
    } else if (k == "coalesce_max_batch") {
        if (value < 1) return fail(FIP_ERR_INVALID, "coalesce_max_batch must be >= 1");
        idx->co.max_batch = value;
    } else if (k == "live_thresholds") {
        idx->live_thr = value != 0;
    } else if (k == "certify") {
        idx->certify = value != 0;
    } else if (k == "filter_copy") {
        idx->filter_copy = value != 0;
        if (!idx->filter_copy && idx->corpus16) {
            if (int rc = use_device(idx)) return rc;
            cudaFree(idx->corpus16);
            idx->corpus16 = nullptr;
            idx->c16_gen = 0;
            idx->c16_usable = false;
        }
    } else if (k == "profile_events") {
        idx->profile = value != 0;
    } else if (k == "fused_rescore") {
        idx->fused_rescore = value != 0;
    } else if (k == "dedupe") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "dedupe must be 0 (off), 1 (coalesced batches) or 2 (always)");
        idx->dedupe = (int)value;
    } else if (k == "defer_epilogue") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "defer_epilogue must be 0 (auto), 1 (off) or 2 (on)");
        idx->defer_epilogue = (int)value;
    } else if (k == "debug_counts") {
        idx->debug_counts = value != 0;
    } else if (k == "force_safe_mode") {
        idx->force_safe = value != 0;
    } else {
        return fail(FIP_ERR_INVALID, "unknown option " + k);
    }
    return FIP_OK;
}

int64_t fip_get_stat(const fip_index* idx, const char* key) {
    if (!idx || !key) return -1;
    std::string k(key);
    if (k == "launches") return idx->st_launches;
    if (k == "variant") return idx->st_variant;
    if (k == "phases") return idx->st_phases;
    if (k == "overflow_queries") return idx->st_overflow;
    if (k == "uncertified_queries") return idx->st_uncert;
    if (k == "stream_redo_queries") return idx->st_stream_redo;
    if (k == "query_split_used") return idx->st_split;
    if (k == "filter_f16") return idx->st_filter16;
    if (k == "deferred_epilogue") return idx->st_deferred;
    if (k == "candidate_capacity") return idx->st_cap;
    if (k == "cta_group") return idx->st_pair;
    if (k == "scan_time_ns") return (int64_t)(idx->st_scan_ms * 1e6);
    if (k == "scan_launches") return idx->st_scan_launches;
    if (k == "scan_searches") return idx->st_scan_searches;
    if (k.rfind("scan_rows_", 0) == 0) {
        const size_t i = (size_t)atoi(k.c_str() + 10);
        return i < idx->st_scan_rows.size() ? idx->st_scan_rows[i] : -1;
    }
    if (k.rfind("scan_ns_", 0) == 0 || k.rfind("gap_ns_", 0) == 0) {
        const bool gap = k[0] == 'g';
        const auto& v = gap ? idx->st_gap_each : idx->st_scan_each;
        const size_t i = (size_t)atoi(k.c_str() + (gap ? 7 : 8));
        return i < v.size() ? (int64_t)(v[i] * 1e6) : -1;
    }
    if (k == "list_entries_after_filter") return idx->st_list_entries;
    if (k == "list_max_after_filter") return idx->st_list_max;
    if (k == "host_batches") return idx->co.st_batches;
    if (k == "host_calls") return idx->co.st_calls;
    if (k == "coalesced_calls") return idx->co.st_coalesced_calls;
    if (k == "last_batch_calls") return idx->co.st_last_batch_calls;
    if (k == "host_queries") return idx->co.st_queries;
    if (k == "distinct_queries") return idx->co.st_distinct;
    if (k == "total_launches") return idx->st_total_launches;
    if (k == "total_searches") return idx->st_total_searches;
    return -1;
}

const char* fip_last_error(void) { return g_last_error.c_str(); }
int32_t fip_abi_version(void) { return 1; }
int32_t fip_compiled_sm(void) { return 100; }

}  // extern "C"

// IVF-Flat (L2) operator: include/ivfflat_b200.h
#include "ivf_api.cuh"
// Multi-GPU group behind one search call: include/flatip_b200.h (fip_group_*)
#include "group_api.cuh"
// Search with the workspace in a static device array (no allocator calls): fip_search_rows_static
#include "static_scratch.cuh"
