// fip_group: ONE corpus row-sharded over several GPUs of THIS process, behind one blocking search
// call -- the process model SURVEY 8(e) names ("one Python process, g devices, one host thread +
// stream per device inside fip_search") and what lets the reference's single in-process call site
//     faiss_store = self.registry.get("faiss_store"); faiss_store.search(embeddings, k)
//     (/root/reference/src/pipeline/services/retrieval/api.py:383-389)
// use all 8 GPUs of a box.  (Included at the end of flatip_api.cu: it uses the file's internals.)
//
// Layout: shard r lives on devices[r] and owns the contiguous global rows [bound[r], bound[r + 1])
// (uniform: bound[r] = r * ceil(planned_rows / g); the last shard also takes whatever is added
// beyond the plan), so "shard, then position" == "id ascending" and the fixed tie-break survives
// the merge.  WEIGHTED shards (fip_group_create_weighted): a search ends with the slowest GPU, and
// under the ~1000 W cap the B200s of one box differ by up to 10 % in the clock they sustain
// (measured: 59.4 vs 65.6 ms for equal halves of the headline corpus on a 2-GPU box), so bound[]
// may follow measured per-device throughput (fip_calibrate_devices) instead: every GPU then
// finishes at the same time.
//
// A search (run_group_batch):
//   every rank, on its own persistent host thread and stream, in parallel:
//     H2D of the (shared, pinned) query batch -> local search ENQUEUED (search_enqueue, no host
//     round trip) -> K4 fused exchange + merge over NVLink peer memory (fip_exchange, wired with
//     connect_local), which also ORs the shards' status words -> rank 0: D2H of the global result;
//     every rank: D2H of the merged status -> ONE stream synchronisation.
//   Only if some shard flagged a query (overflow / uncertified: rare) do the ranks finish their
//   local searches on the host (redo of the flagged queries) and exchange + merge once more.
#pragma once

#include <atomic>
#include <chrono>
#include <functional>
#include <thread>

struct fip_group {
    int world = 0;
    int dim = 0, dtype = 0;
    int64_t planned_rows = 0, per = 0;
    std::vector<int64_t> bound;   // [world + 1] global row boundaries of the shards (relative to row 0)
    std::vector<int> devices;
    std::vector<fip_index*> shard;
    std::vector<fip_exchange*> xchg;
    int xchg_nq = 0, xchg_k = 0;
    bool shared_device = false;   // test configuration: several shards on one GPU (see fip_group_create)

    struct Rank {
        cudaStream_t stream = nullptr;
        DevBuf<float> q32, loc_d, out_d, fin_d;
        DevBuf<int64_t> loc_i, out_i, fin_i;
        int nd = 0;                       // distinct queries of the current search (dedupe.cuh)
        DevBuf<uint32_t> status;          // [1] merged status of the exchange (device)
        PinnedBuf<uint32_t> h_status;     // [1]
        int rc = 0;
        std::string err;
    };
    std::vector<Rank> ranks;
    PinnedBuf<float> h_q, h_d;
    PinnedBuf<int64_t> h_i;

    // persistent workers: one host thread per device
    std::vector<std::thread> workers;
    std::mutex wmu;
    std::condition_variable cv_job, cv_done;
    uint64_t job_gen = 0;
    int outstanding = 0;
    bool stop = false;
    std::function<int(int)> job;

    std::mutex mu;      // serialises group operations
    Coalescer co;
    int dedupe = 1;     // 0 off, 1 coalesced batches, 2 always (every rank dedupes its own copy of the batch)
    std::atomic<int> debug_fail_rank{0};   // testing: r + 1 = rank r fails its next local part (once)
    int64_t st_flagged_searches = 0, st_exchange_launches = 0, st_searches = 0;
};

namespace {

void group_worker(fip_group* g, int r) {
    cudaSetDevice(g->devices[r]);
    uint64_t seen = 0;
    for (;;) {
        std::function<int(int)> fn;
        {
            std::unique_lock<std::mutex> lk(g->wmu);
            g->cv_job.wait(lk, [&] { return g->stop || g->job_gen != seen; });
            if (g->stop) return;
            seen = g->job_gen;
            fn = g->job;
        }
        g_last_error.clear();
        const int rc = fn(r);
        g->ranks[r].rc = rc;
        g->ranks[r].err = rc ? g_last_error : std::string();
        {
            std::lock_guard<std::mutex> lk(g->wmu);
            if (--g->outstanding == 0) g->cv_done.notify_all();
        }
    }
}

// Runs fn(r) on every rank's worker thread at once; returns the first failure (its message becomes
// the calling thread's fip_last_error).
int group_run(fip_group* g, std::function<int(int)> fn) {
    {
        std::lock_guard<std::mutex> lk(g->wmu);
        g->job = std::move(fn);
        g->outstanding = g->world;
        g->job_gen++;
    }
    g->cv_job.notify_all();
    {
        std::unique_lock<std::mutex> lk(g->wmu);
        g->cv_done.wait(lk, [&] { return g->outstanding == 0; });
    }
    for (int r = 0; r < g->world; ++r)
        if (g->ranks[r].rc) return fail(g->ranks[r].rc, "device " + std::to_string(g->devices[r]) + ": " + g->ranks[r].err);
    return FIP_OK;
}

// Global rows [g0, g0 + n) -> (shard, local count, offset into the caller's rows) pieces.
struct GroupPiece { int r; int64_t first, count; };
std::vector<GroupPiece> group_route(const fip_group* g, int64_t g0, int64_t n) {
    std::vector<GroupPiece> out;
    for (int r = 0; r < g->world && n > 0; ++r) {
        const int64_t lo = g->bound[r];
        const int64_t hi = r + 1 == g->world ? INT64_MAX : g->bound[r + 1];
        const int64_t b = std::max(g0, lo), e = std::min(g0 + n, hi);
        if (b < e) out.push_back({r, b - g0, e - b});
    }
    return out;
}

int64_t group_ntotal(const fip_group* g) {
    int64_t t = 0;
    for (auto* s : g->shard) t += s->n_rows;
    return t;
}

int group_ensure_exchange(fip_group* g, int nq, int k) {
    if (g->world == 1) return FIP_OK;
    if (!g->xchg.empty() && nq <= g->xchg_nq && k <= g->xchg_k) return FIP_OK;
    if ((int64_t)g->world * k > kMaxCap)
        return fail(FIP_ERR_INVALID, "devices * k must be <= 16384 (capacity of the merge kernel)");
    for (auto* x : g->xchg) fip_exchange_destroy(x);
    g->xchg.clear();
    const int nq_max = std::max(1024, next_pow2(nq));
    const int k_max = std::min(std::max(128, next_pow2(k)), kMaxCap / g->world);
    g->xchg.assign(g->world, nullptr);
    for (int r = 0; r < g->world; ++r) {
        int rc = fip_exchange_create(g->devices[r], g->world, r, nq_max, k_max, &g->xchg[r]);
        if (rc) return rc;
    }
    for (int r = 0; r < g->world; ++r) {
        int rc = fip_exchange_connect_local(g->xchg[r], g->xchg.data());
        if (rc) return rc;
    }
    g->xchg_nq = nq_max; g->xchg_k = k_max;
    return FIP_OK;
}

// One rank's share of a group search (runs on that rank's worker thread).  second = the exchange
// is repeated after flagged queries were redone on the host.  stage: 0 = everything (single
// device); 1 = the local part: every allocation this search needs, H2D, dedupe, the search
// enqueued (synchronised only in the shared-device test configuration, where a spinning exchange
// kernel must not hold the SMs another shard's search still needs); 2 = exchange onwards;
// 3 = publish launch of the exchange only, synchronised, and 4 = its wait + merge launch onwards:
// 1, 3, 4 is the shared-device sequence, in which no kernel ever waits for a kernel that has not
// finished yet (so it also completes when a profiler serialises every launch).
// Several devices always run 1 then 2 as two rounds over the workers (run_group_batch): the K4
// kernels wait for one another on the GPUs, so a rank may launch its own only once EVERY rank is
// known to have got through the part that can fail (out of memory, a bad launch), and no
// cudaMalloc / cudaFree happens on any rank while a peer's exchange kernel may be resident.
int group_rank_search(fip_group* g, int r, int nq, int k, bool second, bool dedupe, int stage = 0) {
    fip_index* idx = g->shard[r];
    fip_group::Rank& R = g->ranks[r];
    std::lock_guard<std::mutex> lock(idx->mu);
    FIP_CUDA(cudaSetDevice(g->devices[r]));
    const size_t qn = (size_t)nq * g->dim, on = (size_t)nq * k;
    const bool multi = g->world > 1;
    if (stage >= 2) {
        // stage 1 already ran the local part
    } else if (!second) {
        int want = r + 1;
        if (g->debug_fail_rank.compare_exchange_strong(want, 0))
            return fail(FIP_ERR_CUDA, "injected failure of the local part (option debug_fail_rank)");
        FIP_CUDA(R.q32.ensure(qn));
        FIP_CUDA(R.loc_d.ensure(on));
        FIP_CUDA(R.loc_i.ensure(on));
        if (multi) { FIP_CUDA(R.out_d.ensure(on)); FIP_CUDA(R.out_i.ensure(on)); }
        FIP_CUDA(R.status.ensure(1));
        FIP_CUDA(R.h_status.ensure(1));
        if (r == 0 && dedupe) { FIP_CUDA(R.fin_d.ensure(on)); FIP_CUDA(R.fin_i.ensure(on)); }
        FIP_CUDA(cudaMemcpyAsync(R.q32.p, g->h_q.p, qn * sizeof(float), cudaMemcpyHostToDevice, R.stream));
        // bit-identical queries are searched once: every rank reduces its own copy of the batch
        // (same deterministic result everywhere, nothing to communicate)
        R.nd = nq;
        if (dedupe) {
            int rc = dedupe_queries(idx, R.q32.p, nq, R.stream, &R.nd);
            if (rc) return rc;
        }
        int rc = search_enqueue(idx, R.nd < nq ? idx->dd_q.p : R.q32.p, R.nd, k, R.loc_d.p, R.loc_i.p, R.stream);
        if (rc) return rc;
    } else {
        // finish the local search on the host: redo whatever this shard flagged
        int rc = copy_flags_async(idx, R.stream, R.nd);
        if (rc) return rc;
        FIP_CUDA(cudaStreamSynchronize(R.stream));
        int redone = 0;
        rc = search_finish_synced(idx, &redone);
        if (rc) return rc;
    }
    if (stage == 1) {
        if (g->shared_device) FIP_CUDA(cudaStreamSynchronize(R.stream));
        return FIP_OK;
    }
    const int nd = R.nd;
    if (stage == 3) {
        int rc = exchange_launch(g->xchg[r], R.loc_d.p, R.loc_i.p, nd, k, R.out_d.p, R.out_i.p, idx->overflow.p,
                                 R.status.p, R.stream, 1);
        if (rc) return rc;
        FIP_CUDA(cudaStreamSynchronize(R.stream));
        return FIP_OK;
    }
    FIP_CUDA(cudaMemsetAsync(R.status.p, 0, sizeof(uint32_t), R.stream));
    const float* fin_d = R.loc_d.p;
    const int64_t* fin_i = R.loc_i.p;
    if (multi) {
        int rc = exchange_launch(g->xchg[r], R.loc_d.p, R.loc_i.p, nd, k, R.out_d.p, R.out_i.p, idx->overflow.p,
                                 R.status.p, R.stream, stage == 4 ? 2 : 0);
        if (rc) return rc;
        fin_d = R.out_d.p; fin_i = R.out_i.p;
        FIP_CUDA(cudaMemcpyAsync(R.h_status.p, R.status.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, R.stream));
    } else {
        // single device: the status is the shard's own summary words (fetched with all the flags)
        int rc = copy_flags_async(idx, R.stream, nd);
        if (rc) return rc;
    }
    if (r == 0 && nd < nq) {   // every caller's row from its representative's result
        FIP_CUDA(R.fin_d.ensure(on));
        FIP_CUDA(R.fin_i.ensure(on));
        int rc = expand_results(idx, fin_d, fin_i, R.fin_d.p, R.fin_i.p, nq, k, R.stream);
        if (rc) return rc;
        fin_d = R.fin_d.p; fin_i = R.fin_i.p;
    }
    if (r == 0) {
        FIP_CUDA(cudaMemcpyAsync(g->h_d.p, fin_d, on * sizeof(float), cudaMemcpyDeviceToHost, R.stream));
        FIP_CUDA(cudaMemcpyAsync(g->h_i.p, fin_i, on * sizeof(int64_t), cudaMemcpyDeviceToHost, R.stream));
    }
    FIP_CUDA(cudaStreamSynchronize(R.stream));
    if (!multi) R.h_status.p[0] = (idx->h_flags.p[0] ? 1u : 0u) | (idx->h_flags.p[1] ? 2u : 0u);
    return FIP_OK;
}

int run_group_batch(fip_group* g, const std::vector<HostRequest*>& batch) {
    std::lock_guard<std::mutex> lock(g->mu);
    const int k = batch[0]->k;
    int64_t nq64 = 0;
    for (auto* b : batch) nq64 += b->nq;
    const int nq = (int)nq64;
    const size_t dim = (size_t)g->dim;
    FIP_CUDA(cudaSetDevice(g->devices[0]));
    FIP_CUDA(g->h_q.ensure((size_t)nq * dim));
    FIP_CUDA(g->h_d.ensure((size_t)nq * k));
    FIP_CUDA(g->h_i.ensure((size_t)nq * k));
    size_t off = 0;
    for (auto* b : batch) {
        memcpy(g->h_q.p + off * dim, b->q, (size_t)b->nq * dim * sizeof(float));
        off += (size_t)b->nq;
    }
    int rc = group_ensure_exchange(g, nq, k);
    if (rc) return rc;
    const bool dd = g->dedupe == 2 || (g->dedupe == 1 && batch.size() > 1);
    auto all_ranks = [&](bool second) -> int {
        if (g->world == 1) return group_run(g, [g, nq, k, second, dd](int r) { return group_rank_search(g, r, nq, k, second, dd); });
        const int rc1 = group_run(g, [g, nq, k, second, dd](int r) { return group_rank_search(g, r, nq, k, second, dd, 1); });
        if (rc1) {
            // some rank failed before the exchange: NO rank launches it (the epochs stay in step,
            // nothing spins).  Drain what the others enqueued: their H2D copies read g->h_q.
            const std::string why = g_last_error;
            group_run(g, [g](int r) { cudaStreamSynchronize(g->ranks[r].stream); cudaGetLastError(); return 0; });
            return fail(rc1, why);
        }
        if (!g->shared_device) return group_run(g, [g, nq, k, second, dd](int r) { return group_rank_search(g, r, nq, k, second, dd, 2); });
        const int rc3 = group_run(g, [g, nq, k, second, dd](int r) { return group_rank_search(g, r, nq, k, second, dd, 3); });
        if (rc3) return rc3;
        return group_run(g, [g, nq, k, second, dd](int r) { return group_rank_search(g, r, nq, k, second, dd, 4); });
    };
    rc = all_ranks(false);
    if (rc) return rc;
    g->st_searches++;
    g->st_exchange_launches += g->world > 1 ? g->world : 0;
    g->co.st_queries += nq; g->co.st_distinct += g->ranks[0].nd;
    uint32_t status = 0;
    for (auto& R : g->ranks) status |= R.h_status.p[0];   // identical on every rank when world > 1
    if (status) {
        g->st_flagged_searches++;
        if (g->world > 1) {
            rc = all_ranks(true);
            g->st_exchange_launches += g->world;
        } else {
            // the flags are already on the host (copied with the results)
            int redone = 0;
            std::lock_guard<std::mutex> l0(g->shard[0]->mu);
            rc = search_finish_synced(g->shard[0], &redone);
            if (!rc && redone) {
                auto& R0 = g->ranks[0];
                cudaStream_t s = R0.stream;
                const float* fd = R0.loc_d.p;
                const int64_t* fi = R0.loc_i.p;
                if (R0.nd < nq) {
                    rc = expand_results(g->shard[0], fd, fi, R0.fin_d.p, R0.fin_i.p, nq, k, s);
                    if (rc) return rc;
                    fd = R0.fin_d.p; fi = R0.fin_i.p;
                }
                FIP_CUDA(cudaMemcpyAsync(g->h_d.p, fd, (size_t)nq * k * sizeof(float), cudaMemcpyDeviceToHost, s));
                FIP_CUDA(cudaMemcpyAsync(g->h_i.p, fi, (size_t)nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
                FIP_CUDA(cudaStreamSynchronize(s));
            }
        }
        if (rc) return rc;
    } else {
        // nothing flagged: the pending searches are complete as they are
        for (int r = 0; r < g->world; ++r) {
            fip_index* idx = g->shard[r];
            std::lock_guard<std::mutex> l(idx->mu);
            if (idx->pend.active) { idx->pend.active = false; collect_event_times(idx); }
        }
    }
    off = 0;
    for (auto* b : batch) {
        memcpy(b->d, g->h_d.p + off * k, (size_t)b->nq * k * sizeof(float));
        memcpy(b->i, g->h_i.p + off * k, (size_t)b->nq * k * sizeof(int64_t));
        off += (size_t)b->nq;
    }
    g->co.count(batch.size());
    return FIP_OK;
}

}  // namespace

extern "C" {

int fip_group_create(const fip_config* cfg, const int32_t* devices, int32_t n_devices,
                     int64_t planned_rows, fip_group** out) {
    return fip_group_create_weighted(cfg, devices, n_devices, planned_rows, nullptr, out);
}

int fip_group_create_weighted(const fip_config* cfg, const int32_t* devices, int32_t n_devices,
                              int64_t planned_rows, const double* weights, fip_group** out) {
    if (!cfg || !devices || !out) return fail(FIP_ERR_INVALID, "null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > kMaxRanks) return fail(FIP_ERR_INVALID, "1 to 16 devices");
    if (planned_rows < 0) return fail(FIP_ERR_INVALID, "planned_rows must be >= 0");
    // A device listed more than once puts several shards on one GPU: only useful to exercise the
    // whole group path (routing, per-shard id bases, K4, status words) on a single-GPU box.
    bool shared = false;
    for (int a = 0; a < n_devices; ++a)
        for (int b = a + 1; b < n_devices; ++b)
            if (devices[a] == devices[b]) shared = true;
    if (shared && !getenv("FLATIP_B200_ALLOW_SHARED_DEVICE"))
        return fail(FIP_ERR_INVALID, "a device may appear only once in a group "
                                     "(FLATIP_B200_ALLOW_SHARED_DEVICE=1 lifts this for tests)");
    fip_group* g = new (std::nothrow) fip_group();
    if (!g) return fail(FIP_ERR_NOMEM, "host allocation failed");
    g->shared_device = shared;
    g->world = n_devices; g->dim = cfg->dim; g->dtype = cfg->dtype;
    g->planned_rows = planned_rows;
    g->per = std::max<int64_t>(1, (planned_rows + n_devices - 1) / n_devices);
    g->bound.assign(n_devices + 1, 0);
    double wsum = 0.0;
    if (weights)
        for (int r = 0; r < n_devices; ++r) {
            if (!(weights[r] > 0.0)) { delete g; return fail(FIP_ERR_INVALID, "shard weights must be positive"); }
            wsum += weights[r];
        }
    double acc = 0.0;
    for (int r = 0; r < n_devices; ++r) {
        if (weights) {
            // proportional boundaries, on whole 256-row corpus tiles
            acc += weights[r];
            int64_t b = (int64_t)((double)planned_rows * (acc / wsum) / 256.0 + 0.5) * 256;
            g->bound[r + 1] = std::max(g->bound[r], std::min(planned_rows, b));
        } else {
            g->bound[r + 1] = std::min(planned_rows, (int64_t)(r + 1) * g->per);
        }
    }
    g->bound[n_devices] = planned_rows;
    g->devices.assign(devices, devices + n_devices);
    g->ranks.resize(n_devices);
    for (int r = 0; r < n_devices; ++r) {
        fip_config c = *cfg;
        c.device = devices[r];
        c.id_base = cfg->id_base + g->bound[r];
        c.capacity_rows = g->bound[r + 1] - g->bound[r];
        fip_index* idx = nullptr;
        int rc = fip_create(&c, &idx);
        if (rc == FIP_OK && cudaStreamCreateWithFlags(&g->ranks[r].stream, cudaStreamNonBlocking) != cudaSuccess)
            rc = fail(FIP_ERR_CUDA, "cudaStreamCreate failed");
        if (rc) {
            const std::string msg = g_last_error;
            if (idx) fip_destroy(idx);
            fip_group_destroy(g);
            return fail(rc, msg);
        }
        idx->co.enabled = 0;   // coalescing happens once, at the group level
        g->shard.push_back(idx);
    }
    for (int r = 0; r < n_devices; ++r) g->workers.emplace_back(group_worker, g, r);
    *out = g;
    return FIP_OK;
}

void fip_group_destroy(fip_group* g) {
    if (!g) return;
    {
        std::lock_guard<std::mutex> lk(g->wmu);
        g->stop = true;
    }
    g->cv_job.notify_all();
    for (auto& t : g->workers) if (t.joinable()) t.join();
    for (auto* x : g->xchg) fip_exchange_destroy(x);
    for (size_t r = 0; r < g->ranks.size(); ++r) {
        cudaSetDevice(g->devices[r]);
        auto& R = g->ranks[r];
        if (R.stream) { cudaStreamSynchronize(R.stream); cudaStreamDestroy(R.stream); }
        R.q32.release(); R.loc_d.release(); R.out_d.release(); R.loc_i.release(); R.out_i.release();
        R.fin_d.release(); R.fin_i.release();
        R.status.release(); R.h_status.release();
    }
    for (auto* s : g->shard) fip_destroy(s);
    g->h_q.release(); g->h_d.release(); g->h_i.release();
    cudaGetLastError();
    delete g;
}

int32_t fip_group_size(const fip_group* g) { return g ? g->world : 0; }
fip_index* fip_group_shard(fip_group* g, int32_t r) { return (g && r >= 0 && r < g->world) ? g->shard[r] : nullptr; }
int64_t fip_group_ntotal(const fip_group* g) { return g ? group_ntotal(g) : 0; }

namespace {
// An add that failed on one shard after others had succeeded would leave the sum of the shards'
// rows out of step with bound[] (later rows routed to the wrong shard, wrong ids): put every shard
// back to the row count it had before the call.  The appended rows stay in the allocation, unused.
int group_add_rollback(fip_group* g, const std::vector<int64_t>& before, int rc) {
    const std::string why = g_last_error;
    for (int r = 0; r < g->world; ++r) {
        fip_index* idx = g->shard[r];
        std::lock_guard<std::mutex> l(idx->mu);
        if (idx->n_rows != before[r]) {
            idx->n_rows = before[r];
            idx->tmap_x_valid = false;
            idx->corpus_gen++;
        }
    }
    return fail(rc, why + " (group unchanged: rows other shards had already taken were dropped again)");
}
std::vector<int64_t> group_rows_now(const fip_group* g) {
    std::vector<int64_t> v;
    for (auto* s : g->shard) v.push_back(s->n_rows);
    return v;
}
}  // namespace

int fip_group_add_synthetic(fip_group* g, int64_t n, uint64_t seed) {
    if (!g || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(g->mu);
    const int64_t g0 = group_ntotal(g);
    const int64_t base = g->shard[0]->cfg.id_base;
    const auto pieces = group_route(g, g0, n);
    const auto before = group_rows_now(g);
    const int rc = group_run(g, [g, pieces, seed, g0, base](int r) {
        int want = r + 1;
        if (g->debug_fail_rank.compare_exchange_strong(want, 0))
            return fail(FIP_ERR_NOMEM, "injected failure of the add (option debug_fail_rank)");
        for (const auto& p : pieces)
            if (p.r == r) return fip_add_synthetic(g->shard[r], p.count, seed, base + g0 + p.first);
        return (int)FIP_OK;
    });
    return rc ? group_add_rollback(g, before, rc) : FIP_OK;
}

int fip_group_add_host(fip_group* g, const void* rows, int64_t n, int32_t dtype) {
    if (!g || (n > 0 && !rows) || n < 0) return fail(FIP_ERR_INVALID, "bad argument");
    if (n == 0) return FIP_OK;
    std::lock_guard<std::mutex> lock(g->mu);
    const int64_t g0 = group_ntotal(g);
    const auto pieces = group_route(g, g0, n);
    const size_t src_row = (size_t)g->dim * dtype_size(dtype);
    const auto before = group_rows_now(g);
    const int rc = group_run(g, [g, pieces, rows, dtype, src_row](int r) {
        for (const auto& p : pieces)
            if (p.r == r)
                return fip_add_host(g->shard[r], (const char*)rows + (size_t)p.first * src_row, p.count, dtype);
        return (int)FIP_OK;
    });
    return rc ? group_add_rollback(g, before, rc) : FIP_OK;
}

int fip_group_reset(fip_group* g) {
    if (!g) return fail(FIP_ERR_INVALID, "null group");
    std::lock_guard<std::mutex> lock(g->mu);
    for (auto* s : g->shard) { int rc = fip_reset(s); if (rc) return rc; }
    return FIP_OK;
}

int fip_group_search(fip_group* g, const float* queries, int64_t nq, int32_t k, float* distances,
                     int64_t* ids) {
    if (!g) return fail(FIP_ERR_INVALID, "null group");
    int rc = check_search_args(g->shard[0], queries, nq, k, distances, ids);
    if (rc) return rc;
    if (nq == 0) return FIP_OK;
    HostRequest me{queries, nq, k, distances, ids};
    return g->co.submit(me, [g](const std::vector<HostRequest*>& batch) { return run_group_batch(g, batch); });
}

int fip_group_set_option(fip_group* g, const char* key, int64_t value) {
    if (!g || !key) return fail(FIP_ERR_INVALID, "bad argument");
    const std::string k(key);
    if (k == "coalesce") { g->co.enabled = value != 0; return FIP_OK; }
    if (k == "dedupe") {
        if (value < 0 || value > 2) return fail(FIP_ERR_INVALID, "dedupe must be 0, 1 or 2");
        g->dedupe = (int)value;
        return FIP_OK;
    }
    if (k == "coalesce_max_batch") {
        if (value < 1) return fail(FIP_ERR_INVALID, "coalesce_max_batch must be >= 1");
        g->co.max_batch = value;
        return FIP_OK;
    }
    if (k == "debug_fail_rank") {   // tests: shard `value - 1` fails the local part of its next search
        if (value < 0 || value > g->world) return fail(FIP_ERR_INVALID, "debug_fail_rank must be in [0, devices]");
        g->debug_fail_rank = (int)value;
        return FIP_OK;
    }
    for (auto* s : g->shard) { int rc = fip_set_option(s, key, value); if (rc) return rc; }
    return FIP_OK;
}

// Sustained search throughput of every listed device, measured at the same time (the power cap is
// what differs between the GPUs of a box, so they must all be under load): each device searches its
// own `rows`-row synthetic corpus with `batch` queries, k results, for about `seconds`; rates[r] =
// corpus rows scanned per second of scan-kernel time.
int fip_calibrate_devices(const int32_t* devices, int32_t n_devices, int32_t dim, int32_t dtype,
                          int64_t rows, int32_t batch, int32_t k, double seconds, double* rates) {
    if (!devices || !rates || n_devices < 1 || n_devices > kMaxRanks || rows < 1024 || batch < 1 || k < 1)
        return fail(FIP_ERR_INVALID, "bad argument");
    std::vector<int> rcs(n_devices, 0);
    std::vector<std::string> errs(n_devices);
    std::vector<std::thread> th;
    for (int r = 0; r < n_devices; ++r)
        th.emplace_back([&, r]() {
            auto body = [&]() -> int {
                fip_config c{};
                c.dim = dim; c.dtype = dtype; c.device = devices[r]; c.capacity_rows = rows;
                fip_index* idx = nullptr;
                int rc = fip_create(&c, &idx);
                if (rc) return rc;
                struct Guard { fip_index* i; ~Guard() { fip_destroy(i); } } guard{idx};
                rc = fip_add_synthetic(idx, rows, 99, 0);
                if (rc) return rc;
                rc = fip_set_option(idx, "profile_events", 1);
                if (rc) return rc;
                DevBuf<float> q, d;
                DevBuf<int64_t> i;
                struct Scratch {   // released on every way out of this body
                    DevBuf<float>&q, &d; DevBuf<int64_t>& i;
                    ~Scratch() { q.release(); d.release(); i.release(); }
                } scratch{q, d, i};
                FIP_CUDA(q.ensure((size_t)batch * dim));
                FIP_CUDA(d.ensure((size_t)batch * k));
                FIP_CUDA(i.ensure((size_t)batch * k));
                synthetic_fill_kernel<float><<<64, 256, 0, idx->stream>>>(q.p, batch, dim, 0, 4321, 1u);
                FIP_CUDA(cudaGetLastError());
                double scan_ms = 0.0;
                int64_t searches = 0;
                const auto t0 = std::chrono::steady_clock::now();
                bool first = true;
                while (true) {
                    for (int it = 0; it < 8; ++it) {
                        rc = fip_search_device_async(idx, q.p, batch, k, d.p, i.p, idx->stream);
                        if (rc) break;
                    }
                    int32_t flagged = 0;
                    if (!rc) rc = fip_search_status(idx, &flagged);
                    if (rc) break;
                    const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                    if (!first && el > 0.3 * seconds) {   // the first 30 % only warms the chip up
                        scan_ms += idx->st_scan_ms;
                        searches += idx->st_scan_searches;
                    }
                    first = false;
                    if (el > seconds) break;
                }
                if (rc) return rc;
                rates[r] = scan_ms > 0.0 ? (double)rows * (double)searches / (scan_ms * 1e-3) : 0.0;
                return FIP_OK;
            };
            cudaSetDevice(devices[r]);
            g_last_error.clear();
            rcs[r] = body();
            errs[r] = g_last_error;
        });
    for (auto& t : th) t.join();
    for (int r = 0; r < n_devices; ++r)
        if (rcs[r]) return fail(rcs[r], "device " + std::to_string(devices[r]) + ": " + errs[r]);
    return FIP_OK;
}

int64_t fip_group_get_stat(const fip_group* g, const char* key) {
    if (!g || !key) return -1;
    const std::string k(key);
    if (k == "devices") return g->world;
    if (k == "rows_per_shard") return g->per;
    if (k.rfind("shard_begin_", 0) == 0) {
        const int r = atoi(k.c_str() + 12);
        return (r >= 0 && r <= g->world) ? g->bound[r] : -1;
    }
    if (k == "searches") return g->st_searches;
    if (k == "flagged_searches") return g->st_flagged_searches;
    if (k == "exchange_launches") return g->st_exchange_launches;
    if (k == "host_batches") return g->co.st_batches;
    if (k == "host_calls") return g->co.st_calls;
    if (k == "coalesced_calls") return g->co.st_coalesced_calls;
    if (k == "last_batch_calls") return g->co.st_last_batch_calls;
    if (k == "host_queries") return g->co.st_queries;
    if (k == "distinct_queries") return g->co.st_distinct;
    // everything else: summed over the shards' last searches (launches, overflow_queries, ...),
    // except "variant"-like values which are taken from shard 0
    if (k == "variant" || k == "phases" || k == "candidate_capacity" || k == "cta_group" || k == "filter_f16" ||
        k == "query_split_used")
        return fip_get_stat(g->shard[0], key);
    int64_t sum = 0;
    for (auto* s : g->shard) {
        const int64_t v = fip_get_stat(s, key);
        if (v < 0) return v;
        sum += v;
    }
    if (k == "launches") sum += g->world > 1 ? g->world : 0;   // + one K4 launch per device
    return sum;
}

}  // extern "C"
