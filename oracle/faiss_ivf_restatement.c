/*
 * CPU ORACLE / TIMED CPU BASELINE for the IVF-Flat (L2) operator -- TEST INFRASTRUCTURE ONLY
 * (never linked into the product).
 *
 * PARITY UNPINNED (see oracle/ivf_oracle.py header): faiss-cpu 1.13.1 holds the reference's
 * arithmetic (/root/reference/uv.lock:959-960; index built at
 * /root/reference/scripts/create_test_docs.py:83-104, searched at
 * /root/reference/src/pipeline/components/faiss_store.py:152 with nprobe from :84-93); its sources
 * are not under /root/reference.  This file restates faiss's published CPU algorithm for
 * IndexIVFFlat(METRIC_L2).search:
 *
 *   IndexIVF::search(n, x, k):
 *     quantizer->search(n, x, nprobe, coarse_dis, idx)         -- IndexFlatL2 over the centroids
 *     search_preassigned(...), parallel_mode 0: "#pragma omp parallel for" OVER QUERIES; per query
 *       a k-max-heap (CMax: root = current worst, neutral FLT_MAX / id -1), and for each probed
 *       list in coarse order  IVFFlatScannerL2::scan_codes:
 *           dis = fvec_L2sqr(x, code, d);  if (dis < heap_top) heap_replace_top(dis, id)
 *       then heap_reorder -> ascending distances.
 *
 * The deliberate difference: ties are ordered (distance asc, position asc) like the CUDA operator
 * (faiss's order among exact ties is an artefact of its heap).
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* fvec_L2sqr: sum (x - y)^2, vectorised by the compiler (omp simd allows the reassociation that
 * faiss's hand-written AVX kernels do explicitly). */
static inline float l2sqr(const float* x, const float* y, int64_t d) {
    float s = 0.f;
#pragma omp simd reduction(+ : s)
    for (int64_t i = 0; i < d; ++i) {
        const float t = x[i] - y[i];
        s += t * t;
    }
    return s;
}

/* "a is worse than b" under (distance asc, position asc) */
static inline int worse(float da, int64_t pa, float db, int64_t pb) {
    return da > db || (da == db && pa > pb);
}

typedef struct { float* d; int64_t* p; int64_t n, k; } maxheap_t;

static void sift_down(maxheap_t* h, int64_t i) {
    const float d = h->d[i];
    const int64_t p = h->p[i];
    for (;;) {
        int64_t c = 2 * i + 1;
        if (c >= h->n) break;
        if (c + 1 < h->n && worse(h->d[c + 1], h->p[c + 1], h->d[c], h->p[c])) c++;
        if (!worse(h->d[c], h->p[c], d, p)) break;
        h->d[i] = h->d[c]; h->p[i] = h->p[c]; i = c;
    }
    h->d[i] = d; h->p[i] = p;
}
static void sift_up(maxheap_t* h, int64_t i) {
    const float d = h->d[i];
    const int64_t p = h->p[i];
    while (i > 0) {
        const int64_t par = (i - 1) / 2;
        if (!worse(d, p, h->d[par], h->p[par])) break;
        h->d[i] = h->d[par]; h->p[i] = h->p[par]; i = par;
    }
    h->d[i] = d; h->p[i] = p;
}
static inline void heap_offer(maxheap_t* h, float d, int64_t p) {
    if (h->n < h->k) {
        h->d[h->n] = d; h->p[h->n] = p; h->n++;
        sift_up(h, h->n - 1);
    } else if (worse(h->d[0], h->p[0], d, p)) {
        h->d[0] = d; h->p[0] = p;
        sift_down(h, 0);
    }
}

/* nprobe nearest centroids of one query, nearest first (distance asc, list number asc). */
static void coarse_one(const float* q, const float* cent, int64_t nlist, int64_t d, int64_t nprobe,
                       float* hd, int64_t* hp, int64_t* out) {
    maxheap_t h = {hd, hp, 0, nprobe};
    for (int64_t c = 0; c < nlist; ++c) heap_offer(&h, l2sqr(q, cent + c * d, d), c);
    for (int64_t i = h.n; i < nprobe; ++i) out[i] = -1;
    while (h.n > 0) {   /* pop worst-first into the tail */
        out[h.n - 1] = h.p[0];
        h.n--;
        if (h.n > 0) { h.d[0] = h.d[h.n]; h.p[0] = h.p[h.n]; sift_down(&h, 0); }
    }
}

/*
 * queries [nq, d]; centroids [nlist, d]; off [nlist + 1] list boundaries in the concatenated
 * codes [ntotal, d] / ids [ntotal].  Outputs: D [nq, k] ascending squared L2 (FLT_MAX padding),
 * I [nq, k] ids (-1 padding).  nthreads <= 0: all OpenMP threads.  Returns 0.
 */
int oracle_ivf_search_l2(const float* queries, int64_t nq, int64_t d, const float* centroids,
                         int64_t nlist, const int64_t* off, const float* codes, const int64_t* ids,
                         int64_t k, int64_t nprobe, float* D, int64_t* I, int nthreads) {
    if (nprobe > nlist) nprobe = nlist;
    if (nq <= 0 || k <= 0 || nprobe <= 0) return nq < 0 || k <= 0 || nprobe <= 0 ? -1 : 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
    int failed = 0;
#pragma omp parallel
    {
        const int64_t hk = k > nprobe ? k : nprobe;
        float* hd = (float*)malloc(sizeof(float) * (size_t)hk);
        int64_t* hp = (int64_t*)malloc(sizeof(int64_t) * (size_t)hk);
        int64_t* probes = (int64_t*)malloc(sizeof(int64_t) * (size_t)nprobe);
        if (!hd || !hp || !probes) {
#pragma omp atomic write
            failed = 1;
        } else {
#pragma omp for schedule(dynamic, 1)
            for (int64_t qi = 0; qi < nq; ++qi) {
                const float* q = queries + qi * d;
                coarse_one(q, centroids, nlist, d, nprobe, hd, hp, probes);
                maxheap_t h = {hd, hp, 0, k};
                for (int64_t j = 0; j < nprobe; ++j) {
                    const int64_t l = probes[j];
                    if (l < 0) continue;
                    for (int64_t pos = off[l]; pos < off[l + 1]; ++pos)
                        heap_offer(&h, l2sqr(q, codes + pos * d, d), pos);
                }
                float* Dq = D + qi * k;
                int64_t* Iq = I + qi * k;
                for (int64_t i = h.n; i < k; ++i) { Dq[i] = FLT_MAX; Iq[i] = -1; }
                while (h.n > 0) {
                    Dq[h.n - 1] = h.d[0];
                    Iq[h.n - 1] = ids[h.p[0]];
                    h.n--;
                    if (h.n > 0) { h.d[0] = h.d[h.n]; h.p[0] = h.p[h.n]; sift_down(&h, 0); }
                }
            }
        }
        free(hd); free(hp); free(probes);
    }
    return failed ? -2 : 0;
}
