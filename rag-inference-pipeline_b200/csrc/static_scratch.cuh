// fip_search_rows_static: exact flat inner-product search that never asks the CUDA allocator for
// anything.  (Included at the end of flatip_api.cu: it reuses the file's kernels and phase rules.)
//
// Who this is for.  Every other search entry point of this library grows a per-index workspace with
// cudaMalloc on first use (candidate lists, thresholds, flag words).  Some hosts of a retrieval
// operator cannot afford that call at search time:
//   * a retriever that shares its GPU with an LLM server which has already claimed nearly all of
//     HBM for its KV cache -- a cudaMalloc of a few MB then simply fails, at the worst moment;
//   * a service whose caching allocator owns the device and which treats any raw cudaMalloc /
//     cudaFree as forbidden because both synchronise the whole device and stall the decode loop
//     running beside the search;
//   * latency-bound callers that are only allowed to run code with a fixed memory footprint after
//     start-up.
// For them the workspace lives in ONE file-scope __device__ array below.  It is part of the library
// image: each GPU that loads libflatip_b200.so reserves kStaticScratchBytes (about 8.1 MiB) for the
// life of the process, whether or not this entry point is ever called.
//
// What runs: the streaming scan (K2, stream_scan.cuh) and the select pass (K3, select.cuh) -- the
// same kernels and the same phase schedule fip_search_device uses with "variant" = STREAM, so the
// two return identical bits.  The only difference is where thr / cnt / cand / hist / flags live.
//
// Limits: at most kStaticMaxQueries (64) queries per call and k <= 2048; larger requests are
// refused with FIP_ERR_INVALID (use fip_search_device, which allocates).  Rows must be a multiple of
// 16 bytes and 16-byte aligned, fewer than 2^31 of them.
//
// Sharing: all calls on one GPU use the same array, and it still holds whatever the previous call
// left there.  Each call resets the words that must start clean (list lengths, thresholds, flag
// words; the histogram is rewritten by the first select pass before any scan reads it).  Two calls
// whose kernels were in flight on one GPU at the same time would trample each other's lists; the
// entry point therefore holds a per-device lock until its own work has drained from the stream, so
// threads of one process queue up instead of racing.  (A stream of your own that still has kernels
// of an EARLIER call in flight cannot exist either: the call returns only once its work is done.)
//
// Set-up: the first call on a device looks up the array's device address and the SM count; later
// calls reuse them.  Make that first call before you begin any stream capture.
#pragma once

namespace {

constexpr int kStaticMaxQueries = 64;
constexpr size_t kStaticCandBytes = (size_t)kStaticMaxQueries * kMaxCap * sizeof(uint64_t);   // 8 MiB
constexpr size_t kStaticHistBytes = (size_t)kStaticMaxQueries * kHistBins * sizeof(uint32_t); // 64 KiB
constexpr size_t kStaticWordBytes = 256;   // one [kStaticMaxQueries] array of 4-byte words
constexpr size_t kStaticFlagBytes = 768;   // flag_words(64) = 130 words, rounded up to 256-byte units
constexpr size_t kStaticOffHist = kStaticCandBytes;
constexpr size_t kStaticOffThr = kStaticOffHist + kStaticHistBytes;
constexpr size_t kStaticOffCnt = kStaticOffThr + kStaticWordBytes;
constexpr size_t kStaticOffHbase = kStaticOffCnt + kStaticWordBytes;
constexpr size_t kStaticOffFlags = kStaticOffHbase + kStaticWordBytes;
constexpr size_t kStaticScratchBytes = kStaticOffFlags + kStaticFlagBytes;
static_assert(kStaticMaxQueries * sizeof(uint32_t) <= kStaticWordBytes, "word arrays");
static_assert((2 + 2 * kStaticMaxQueries) * sizeof(uint32_t) <= kStaticFlagBytes, "flag words");

}  // namespace

// The workspace itself.  In HOST code this name is only a handle for cudaGetSymbolAddress; it is
// not a device pointer and must never be handed to a kernel as one (static_scratch_for does the
// translation, once per device).
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {

constexpr int kStaticMaxDevices = 64;
struct StaticDevice {
    std::mutex mu;            // held for a whole call: one user of the array per GPU at a time
    unsigned char* base = nullptr;   // device address of g_static_scratch on this GPU
    int num_sms = 0;
};
StaticDevice g_static_dev[kStaticMaxDevices];

// Device address of the array on `device` (which must be the current device) -- looked up on the
// first call, cached afterwards.  Caller holds sd.mu.
int static_scratch_for(StaticDevice& sd, int device) {
    if (sd.base) return FIP_OK;
    int major = 0, sms = 0;
    FIP_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10)
        return fail(FIP_ERR_NODEVICE, "device is not sm_100 (B200); kernels are built for sm_100a only");
    FIP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    unsigned char* p = nullptr;
    FIP_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
    if (!p) return fail(FIP_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static workspace");
    sd.num_sms = sms;
    sd.base = p;
    return FIP_OK;
}

struct StaticWork {
    uint64_t* cand;
    uint32_t *hist, *cnt, *hbase, *flags;
    float* thr;
    int num_sms;
    cudaStream_t stream;
};

template <typename T, int QB, int R>
int static_stream_pass(const StaticWork& w, const ScanArgs& a, const float* q32, int q0) {
    auto kern = stream_scan_kernel<T, QB, R>;
    const StreamPlan plan = plan_stream(QB, R, (size_t)a.dim * sizeof(T), a.dim);
    if (plan.warps == 0 || plan.slots < 2)
        return fail(FIP_ERR_INVALID, "dim too large for the streaming kernel (a row block does not fit "
                                     "in shared memory twice)");
    FIP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.smem));
    const int64_t blocks = ((int64_t)(a.row_end - a.row_begin) + R - 1) / R;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(w.num_sms, (blocks + plan.warps - 1) / plan.warps));
    kern<<<grid, plan.warps * 32, plan.smem, w.stream>>>(a, q32, q0, plan.slots);
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

// Same split of a batch into 16 / 8 / 4 / 2 / 1-query passes as launch_stream_phase.
template <typename T>
int static_stream_phase(const StaticWork& w, const ScanArgs& a, const float* q32) {
    constexpr int R = sizeof(T) == 4 ? 2 : 4;
    for (int q0 = 0; q0 < a.nq;) {
        const int rem = a.nq - q0;
        int qb = rem >= 16 ? 16 : next_pow2(rem);
        while (qb > 1 && plan_stream(qb, qb == 16 ? R / 2 : R, (size_t)a.dim * sizeof(T), a.dim).slots < 2) qb >>= 1;
        int rc;
        switch (qb) {
            case 1: rc = static_stream_pass<T, 1, R>(w, a, q32, q0); break;
            case 2: rc = static_stream_pass<T, 2, R>(w, a, q32, q0); break;
            case 4: rc = static_stream_pass<T, 4, R>(w, a, q32, q0); break;
            case 8: rc = static_stream_pass<T, 8, R>(w, a, q32, q0); break;
            default: rc = static_stream_pass<T, 16, R / 2>(w, a, q32, q0); qb = 16; break;
        }
        if (rc) return rc;
        q0 += qb;
    }
    return FIP_OK;
}

int static_select(const StaticWork& w, int nq, int k, int cap, bool final_pass, float* out_d, int64_t* out_i,
                  int64_t id_base, int64_t dense_count, bool live) {
    SelectArgs s{};
    s.cand = w.cand; s.cnt = w.cnt; s.thr = w.thr; s.cap = (uint32_t)cap; s.k = k;
    s.final_pass = final_pass ? 1 : 0; s.out_d = out_d; s.out_i = out_i; s.id_base = id_base;
    s.overflow = w.flags; s.dense_count = (int)dense_count;
    if (live && !final_pass) { s.hist = w.hist; s.hbase = w.hbase; }
    const size_t smem = ((size_t)cap + (size_t)next_pow2(std::min(k, cap))) * sizeof(uint64_t);
    if (select_wants_big(cap)) {
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel_big<<<nq, kSelectThreadsBig, smem, w.stream>>>(s);
    } else {
        FIP_CUDA(cudaFuncSetAttribute(select_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        select_topk_kernel<<<nq, kSelectThreads, smem, w.stream>>>(s);
    }
    FIP_CUDA(cudaGetLastError());
    return FIP_OK;
}

// One pass over the rows.  safe = false: the geometric phases with live thresholds run_search uses
// for the streaming variant (a list can overflow on adversarial row orders; the flag words say so).
// safe = true: phases short enough that no list can overflow.  Enqueue only.
int static_search_once(const StaticWork& w, const void* rows, int64_t n, int dim, int dtype, int64_t id_base,
                       const float* q32, int nq, int k, float* out_d, int64_t* out_i, bool safe) {
    const bool live = !safe;
    const int cap = cap_for_k(k, k, false, live, safe, nq, n);
    init_lists_kernel<<<(unsigned)((flag_words(nq) + 255) / 256), 256, 0, w.stream>>>(w.thr, w.cnt, w.flags, nullptr, nq);
    FIP_CUDA(cudaGetLastError());
    ScanArgs a{};
    a.corpus = rows; a.n_rows = n; a.dim = dim; a.nq = nq; a.dtype = dtype;
    a.thr = w.thr; a.cnt = w.cnt; a.cand = w.cand; a.cap = (uint32_t)cap;
    if (live) { a.hist = w.hist; a.hbase = w.hbase; a.k_thr = k; }
    auto scan = [&](int64_t r0, int64_t r1, bool dense) -> int {
        a.row_begin = (uint32_t)r0; a.row_end = (uint32_t)r1; a.dense = dense ? 1 : 0;
        switch (dtype) {
            case kF32: return static_stream_phase<float>(w, a, q32);
            case kF16: return static_stream_phase<__half>(w, a, q32);
            default: return static_stream_phase<__nv_bfloat16>(w, a, q32);
        }
    };
    const double ratio = phase_ratio(cap, k, 0, live);
    const int64_t first_wave = 2ll * w.num_sms * tc::BN;
    const int64_t need = ((int64_t)k * first_wave * 2 / cap + 255) / 256 * 256;
    const int64_t dense_rows = live ? std::min<int64_t>(std::max<int64_t>(1024, need), cap) : cap;
    int64_t m = std::min<int64_t>(n, dense_rows);
    int rc = scan(0, m, true);
    if (rc) return rc;
    int64_t pos = m;
    rc = static_select(w, nq, k, cap, pos == n, out_d, out_i, id_base, m, live);
    if (rc) return rc;
    while (pos < n) {
        m = safe ? std::max<int64_t>(256, (int64_t)(cap - k) / 256 * 256) : phase_rows(pos, ratio);
        m = std::min<int64_t>(m, n - pos);
        rc = scan(pos, pos + m, false);
        if (rc) return rc;
        pos += m;
        rc = static_select(w, nq, k, cap, pos == n, out_d, out_i, id_base, -1, live);
        if (rc) return rc;
    }
    return FIP_OK;
}

}  // namespace

extern "C" int fip_search_rows_static(int32_t device, const void* rows_dev, int64_t n_rows, int32_t dim,
                                      int32_t dtype, int64_t id_base, const float* queries_dev, int64_t nq,
                                      int32_t k, float* distances_dev, int64_t* ids_dev, void* cuda_stream) {
    if (dim <= 0) return fail(FIP_ERR_INVALID, "dim must be positive");
    if (dtype != FIP_F32 && dtype != FIP_F16 && dtype != FIP_BF16) return fail(FIP_ERR_INVALID, "unknown dtype");
    if (((int64_t)dim * dtype_size(dtype)) % 16 != 0)
        return fail(FIP_ERR_INVALID, "row size must be a multiple of 16 bytes "
                                     "(dim % 4 == 0 for fp32, dim % 8 == 0 for fp16/bf16)");
    if (n_rows < 0 || n_rows >= (1ll << 31)) return fail(FIP_ERR_INVALID, "n_rows must be in [0, 2^31)");
    if (nq < 0) return fail(FIP_ERR_INVALID, "nq out of range");
    if (nq > kStaticMaxQueries)
        return fail(FIP_ERR_INVALID, "fip_search_rows_static serves at most 64 queries per call (its workspace is a "
                                     "fixed " + std::to_string(kStaticScratchBytes) + "-byte static array); split the "
                                     "batch or use fip_search_device");
    if (k < 1 || k > kMaxK)
        return fail(FIP_ERR_INVALID, "fip_search_rows_static: k must be in [1, 2048] (fixed-size static workspace)");
    if (nq == 0) return FIP_OK;
    if (!queries_dev || !distances_dev || !ids_dev || (n_rows > 0 && !rows_dev)) return fail(FIP_ERR_INVALID, "null buffer");
    if (((uintptr_t)rows_dev & 15u) != 0) return fail(FIP_ERR_INVALID, "rows_dev must be 16-byte aligned");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(FIP_ERR_NODEVICE, "no CUDA device visible; this library has no CPU fallback");
    }
    if (device < 0 || device >= ndev || device >= kStaticMaxDevices) return fail(FIP_ERR_INVALID, "bad device ordinal");
    FIP_CUDA(cudaSetDevice(device));
    StaticDevice& sd = g_static_dev[device];
    std::lock_guard<std::mutex> lock(sd.mu);   // released only after the stream has drained (below)
    int rc = static_scratch_for(sd, device);
    if (rc) return rc;
    StaticWork w{};
    w.cand = reinterpret_cast<uint64_t*>(sd.base);
    w.hist = reinterpret_cast<uint32_t*>(sd.base + kStaticOffHist);
    w.thr = reinterpret_cast<float*>(sd.base + kStaticOffThr);
    w.cnt = reinterpret_cast<uint32_t*>(sd.base + kStaticOffCnt);
    w.hbase = reinterpret_cast<uint32_t*>(sd.base + kStaticOffHbase);
    w.flags = reinterpret_cast<uint32_t*>(sd.base + kStaticOffFlags);
    w.num_sms = sd.num_sms;
    w.stream = (cudaStream_t)cuda_stream;
    if (n_rows == 0) {
        const int64_t tot = nq * (int64_t)k;
        fill_padding_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, w.stream>>>(distances_dev, ids_dev, tot);
        FIP_CUDA(cudaGetLastError());
        FIP_CUDA(cudaStreamSynchronize(w.stream));
        return FIP_OK;
    }
    uint32_t summary[2] = {0u, 0u};   // flag words [0] = some list overflowed, [1] unused here
    for (int attempt = 0; attempt < 2; ++attempt) {
        rc = static_search_once(w, rows_dev, n_rows, dim, dtype, id_base, queries_dev, (int)nq, k, distances_dev,
                                ids_dev, attempt == 1);
        if (rc == FIP_OK) {
            cudaError_t e = cudaMemcpyAsync(summary, w.flags, sizeof(summary), cudaMemcpyDeviceToHost, w.stream);
            if (e != cudaSuccess) { cudaGetLastError(); rc = fail(FIP_ERR_CUDA, std::string("flag read-back: ") + cudaGetErrorString(e)); }
        }
        // drain the stream whatever happened: the lock must not be released with kernels of this
        // call still able to touch the array
        cudaError_t es = cudaStreamSynchronize(w.stream);
        if (rc) return rc;
        if (es != cudaSuccess) { cudaGetLastError(); return fail(FIP_ERR_CUDA, std::string("fip_search_rows_static: ") + cudaGetErrorString(es)); }
        if (!summary[0]) return FIP_OK;   // no list overflowed: the result is exact
    }
    return fail(FIP_ERR_CUDA, "fip_search_rows_static: a candidate list overflowed in safe mode (internal error)");
}
