/*
 * CPU ORACLE / TIMED CPU BASELINE -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * PARITY UNPINNED (see oracle/flatip_oracle.py header): faiss-cpu 1.13.1 is the third-party
 * dependency that holds the reference's arithmetic for this path (/root/reference/uv.lock:959-960,
 * called at /root/reference/src/pipeline/components/faiss_store.py:152); its sources are not under
 * /root/reference and the wheel cannot be installed here.  This file restates faiss's *published*
 * CPU algorithm for IndexFlatIP.search:
 *
 *   knn_inner_product(x[nx,d], y[ny,d], k):
 *     nx <  20 : "seq"  path -- OpenMP over queries, every thread scans all rows with a SIMD dot
 *     nx >= 20 : "blas" path -- for query blocks of 4096 x corpus blocks of 1024:
 *                 sgemm("T","N", nyi, nxi, d, 1, y_blk, d, x_blk, d, 0, ip, nyi); add_results()
 *     result handler: k < 100 -> per-query min-heap of the k largest;
 *                     k >= 100 -> reservoir of capacity 2k with a running threshold, shrunk to k
 *                                 by partitioning when it fills up.
 *     end: each row sorted best-first.  Unfilled slots: id -1, score -FLT_MAX.
 *
 * The only deliberate difference: the final order among equal scores is (score desc, id asc), the
 * tie-break BASELINE.json's north_star fixes (faiss's own order there is an implementation detail).
 *
 * The sgemm is supplied by the caller as a Fortran-ABI ILP64 function pointer (the OpenBLAS that
 * numpy bundles: scipy_sgemm_64_), so the timed baseline uses a real BLAS like faiss does; a plain
 * OpenMP/SIMD fallback is used when the pointer is NULL.
 */
#include <float.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef void (*sgemm_ilp64_fn)(const char*, const char*, const int64_t*, const int64_t*,
                               const int64_t*, const float*, const float*, const int64_t*,
                               const float*, const int64_t*, const float*, float*, const int64_t*);

enum { BLAS_THRESHOLD = 20, QUERY_BS = 4096, DATABASE_BS = 1024, MIN_K_RESERVOIR = 100 };

/* ---- ordering: "a is worse than b" under (score desc, id asc) ------------------------------ */
static inline int worse(float sa, int64_t ia, float sb, int64_t ib) {
    return sa < sb || (sa == sb && ia > ib);
}

/* ---- min-heap of the k best (root = current worst) ---------------------------------------- */
typedef struct { float* s; int64_t* id; int64_t k; int64_t n; } heap_t;

static void heap_sift_down(heap_t* h, int64_t i) {
    float s = h->s[i]; int64_t id = h->id[i];
    for (;;) {
        int64_t c = 2 * i + 1;
        if (c >= h->n) break;
        if (c + 1 < h->n && worse(h->s[c + 1], h->id[c + 1], h->s[c], h->id[c])) c++;
        if (!worse(h->s[c], h->id[c], s, id)) break;
        h->s[i] = h->s[c]; h->id[i] = h->id[c]; i = c;
    }
    h->s[i] = s; h->id[i] = id;
}
static void heap_sift_up(heap_t* h, int64_t i) {
    float s = h->s[i]; int64_t id = h->id[i];
    while (i > 0) {
        int64_t p = (i - 1) / 2;
        if (!worse(s, id, h->s[p], h->id[p])) break;
        h->s[i] = h->s[p]; h->id[i] = h->id[p]; i = p;
    }
    h->s[i] = s; h->id[i] = id;
}
static inline void heap_add(heap_t* h, float s, int64_t id) {
    if (h->n < h->k) {
        h->s[h->n] = s; h->id[h->n] = id; h->n++;
        heap_sift_up(h, h->n - 1);
    } else if (s > h->s[0]) { /* strict: an equal score met later (higher id) does not displace */
        h->s[0] = s; h->id[0] = id;
        heap_sift_down(h, 0);
    }
}

/* ---- reservoir (capacity 2k, threshold) ---------------------------------------------------- */
typedef struct { float* s; int64_t* id; int64_t k, cap, n; float thr; } reservoir_t;

static void swap_e(float* s, int64_t* id, int64_t a, int64_t b) {
    float ts = s[a]; s[a] = s[b]; s[b] = ts;
    int64_t ti = id[a]; id[a] = id[b]; id[b] = ti;
}
/* quickselect: move the k best of [0,n) to the front (unordered) */
static void select_k_best(float* s, int64_t* id, int64_t n, int64_t k) {
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
        int64_t mid = lo + (hi - lo) / 2;
        float ps = s[mid]; int64_t pi = id[mid];
        int64_t i = lo, j = hi;
        while (i <= j) {
            while (worse(ps, pi, s[i], id[i])) i++;   /* s[i] better than pivot */
            while (worse(s[j], id[j], ps, pi)) j--;   /* s[j] worse than pivot  */
            if (i <= j) { swap_e(s, id, i, j); i++; j--; }
        }
        if (k - 1 <= j) hi = j; else if (k - 1 >= i) lo = i; else break;
    }
}
static void reservoir_shrink(reservoir_t* r) {
    select_k_best(r->s, r->id, r->n, r->k);
    r->n = r->k;
    float t = r->s[0];
    for (int64_t i = 1; i < r->k; i++) if (r->s[i] < t) t = r->s[i];
    r->thr = t;
}
static inline void reservoir_add(reservoir_t* r, float s, int64_t id) {
    if (s > r->thr) {
        r->s[r->n] = s; r->id[r->n] = id; r->n++;
        if (r->n == r->cap) reservoir_shrink(r);
    }
}

/* ---- final ordering ------------------------------------------------------------------------ */
typedef struct { float s; int64_t id; } pair_t;
static int cmp_pair(const void* a, const void* b) {
    const pair_t* x = (const pair_t*)a; const pair_t* y = (const pair_t*)b;
    if (x->s > y->s) return -1;
    if (x->s < y->s) return 1;
    return (x->id > y->id) - (x->id < y->id);
}
static void finalize_row(const float* s, const int64_t* id, int64_t n, int64_t k, int64_t id_base,
                         float* D, int64_t* I) {
    pair_t* p = (pair_t*)malloc(sizeof(pair_t) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; i++) { p[i].s = s[i]; p[i].id = id[i]; }
    qsort(p, (size_t)n, sizeof(pair_t), cmp_pair);
    int64_t m = n < k ? n : k;
    for (int64_t i = 0; i < m; i++) { D[i] = p[i].s; I[i] = p[i].id + id_base; }
    for (int64_t i = m; i < k; i++) { D[i] = -FLT_MAX; I[i] = -1; }
    free(p);
}

static inline float dot_f32(const float* a, const float* b, int64_t d) {
    float acc = 0.f;
#pragma omp simd reduction(+ : acc)
    for (int64_t i = 0; i < d; i++) acc += a[i] * b[i];
    return acc;
}

static void fallback_gemm(int64_t nyi, int64_t nxi, int64_t d, const float* y, const float* x,
                          float* ip) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nxi; i++)
        for (int64_t j = 0; j < nyi; j++) ip[i * nyi + j] = dot_f32(x + i * d, y + j * d, d);
}

/*
 * x: queries [nx, d]; y: corpus [ny, d]; outputs D [nx, k] float, I [nx, k] int64.
 * sgemm may be NULL.  nthreads <= 0 keeps the OpenMP default.  Returns 0 on success.
 */
int oracle_knn_inner_product(const float* x, const float* y, int64_t d, int64_t nx, int64_t ny,
                             int64_t k, int64_t id_base, float* D, int64_t* I,
                             sgemm_ilp64_fn sgemm, int nthreads) {
    if (!x || !y || !D || !I || d <= 0 || nx < 0 || ny < 0 || k <= 0) return -1;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    const int use_reservoir = k >= MIN_K_RESERVOIR;
    const int64_t cap = use_reservoir ? 2 * k : k;
    float* hs = (float*)malloc(sizeof(float) * (size_t)(nx * cap + 1));
    int64_t* hi = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nx * cap + 1));
    int64_t* hn = (int64_t*)calloc((size_t)nx + 1, sizeof(int64_t));
    float* hthr = (float*)malloc(sizeof(float) * (size_t)(nx + 1));
    if (!hs || !hi || !hn || !hthr) { free(hs); free(hi); free(hn); free(hthr); return -2; }
    for (int64_t i = 0; i < nx; i++) hthr[i] = -FLT_MAX;

    if (nx < BLAS_THRESHOLD) {
        /* "seq" path: parallel over queries only (nx == 1 -> one thread), like faiss. */
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < nx; i++) {
            heap_t h = { hs + i * cap, hi + i * cap, k, 0 };
            reservoir_t r = { hs + i * cap, hi + i * cap, k, cap, 0, -FLT_MAX };
            const float* xi = x + i * d;
            for (int64_t j = 0; j < ny; j++) {
                float s = dot_f32(xi, y + j * d, d);
                if (use_reservoir) reservoir_add(&r, s, j); else heap_add(&h, s, j);
            }
            hn[i] = use_reservoir ? r.n : h.n;
        }
    } else {
        float* ip = (float*)malloc(sizeof(float) * (size_t)QUERY_BS * DATABASE_BS);
        if (!ip) { free(hs); free(hi); free(hn); free(hthr); return -2; }
        const float one = 1.f, zero = 0.f;
        for (int64_t i0 = 0; i0 < nx; i0 += QUERY_BS) {
            int64_t i1 = i0 + QUERY_BS < nx ? i0 + QUERY_BS : nx;
            int64_t nxi = i1 - i0;
            for (int64_t j0 = 0; j0 < ny; j0 += DATABASE_BS) {
                int64_t j1 = j0 + DATABASE_BS < ny ? j0 + DATABASE_BS : ny;
                int64_t nyi = j1 - j0;
                if (sgemm) sgemm("T", "N", &nyi, &nxi, &d, &one, y + j0 * d, &d, x + i0 * d, &d,
                                 &zero, ip, &nyi);
                else fallback_gemm(nyi, nxi, d, y + j0 * d, x + i0 * d, ip);
#pragma omp parallel for schedule(static)
                for (int64_t i = i0; i < i1; i++) {
                    const float* row = ip + (i - i0) * nyi;
                    if (use_reservoir) {
                        reservoir_t r = { hs + i * cap, hi + i * cap, k, cap, hn[i], hthr[i] };
                        for (int64_t j = 0; j < nyi; j++) reservoir_add(&r, row[j], j0 + j);
                        hn[i] = r.n; hthr[i] = r.thr;
                    } else {
                        heap_t h = { hs + i * cap, hi + i * cap, k, hn[i] };
                        for (int64_t j = 0; j < nyi; j++) heap_add(&h, row[j], j0 + j);
                        hn[i] = h.n;
                    }
                }
            }
        }
        free(ip);
    }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nx; i++)
        finalize_row(hs + i * cap, hi + i * cap, hn[i], k, id_base, D + i * k, I + i * k);
    free(hs); free(hi); free(hn); free(hthr);
    return 0;
}

/*
 * Synthetic corpus generator, C restatement of rag-inference-pipeline_b200/synthetic.py (and of the
 * device generator): Philox4x32-10, counter (row_lo, row_hi, col/2, stream), key = seed; element =
 * (sum of four 16-bit halves - 131070) * scale.  Writes fp32 rows [row0, row0+nrows) x dim.
 * storage: 0 = keep fp32, 1 = round through fp16 (RNE), 2 = round through bf16 (RNE).
 */
static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
static inline float round_bf16(float x) {
    uint32_t u; memcpy(&u, &x, 4);
    u = (u + 0x7FFFu + ((u >> 16) & 1u)) & 0xFFFF0000u;
    memcpy(&x, &u, 4); return x;
}
static inline float round_f16(float x) { return (float)(_Float16)x; }

int oracle_synthetic_rows(float* out, int64_t row0, int64_t nrows, int64_t dim, uint64_t seed,
                          uint32_t stream, int storage) {
    if (!out || nrows < 0 || dim <= 0) return -1;
    const float scale = (float)(1.0 / 37837.22723876);
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < nrows; r++) {
        uint64_t row = (uint64_t)(row0 + r);
        float* o = out + r * dim;
        for (int64_t p = 0; p < (dim + 1) / 2; p++) {
            uint32_t c[4] = { (uint32_t)row, (uint32_t)(row >> 32), (uint32_t)p, stream };
            philox4x32_10(c, k0, k1);
            int32_t s0 = (int32_t)((c[0] & 0xFFFF) + (c[0] >> 16) + (c[1] & 0xFFFF) + (c[1] >> 16)) - 131070;
            int32_t s1 = (int32_t)((c[2] & 0xFFFF) + (c[2] >> 16) + (c[3] & 0xFFFF) + (c[3] >> 16)) - 131070;
            float e0 = (float)s0 * scale, e1 = (float)s1 * scale;
            if (storage == 1) { e0 = round_f16(e0); e1 = round_f16(e1); }
            else if (storage == 2) { e0 = round_bf16(e0); e1 = round_bf16(e1); }
            o[2 * p] = e0;
            if (2 * p + 1 < dim) o[2 * p + 1] = e1;
        }
    }
    return 0;
}

/* Pins the OpenMP team size of this library (independent of OMP_NUM_THREADS). */
void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
