/* C ABI of the B200-native IVF-Flat (L2) operator -- second operator behind the retrieval store's
 * `search(queries, k) -> (distances, ids)` boundary.
 *
 * What it replaces in the reference (third-party faiss-cpu 1.13.1, /root/reference/uv.lock:959-960):
 *   faiss.IndexIVFFlat(faiss.IndexFlatL2(d), d, nlist, METRIC_L2)   scripts/create_test_docs.py:83-85
 *   index.train(x) / index.add(x) / index.nprobe = 64               scripts/create_test_docs.py:88-101
 *   faiss.read_index(path); index.nprobe = settings.faiss_nprobe    src/pipeline/components/faiss_store.py:66-93
 *   index.search(embeddings, k)                                      src/pipeline/components/faiss_store.py:152
 *
 * Semantics (faiss IndexIVFFlat, METRIC_L2): a query is compared with the vectors of the `nprobe`
 * inverted lists whose centroids are nearest to it (squared L2); the k smallest squared L2
 * distances are returned ascending with their ids; missing results are padded with id -1 and
 * distance FLT_MAX.  Fixed tie-break: (distance ascending, list position ascending), which is id
 * ascending for rows added in id order.
 *
 * Same conventions as flatip_b200.h: plain pointers and sizes, 0 = success, negative FIP_ERR_*
 * codes, message via fip_last_error() (thread-local), no exceptions, no CPU fallback (FIP_ERR_NODEVICE
 * without an sm_100 GPU).  All host buffers are caller-owned.  Calls on one handle are serialised.
 */
#ifndef IVFFLAT_B200_H
#define IVFFLAT_B200_H

#include <stdint.h>

#include "flatip_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fivf_index fivf_index;

/* IndexIVFFlat(quantizer = IndexFlatL2(dim), dim, nlist, METRIC_L2) on GPU `device`.
 * dim % 32 == 0, 32 <= dim <= 4096, 1 <= nlist <= 16384. */
int fivf_create(int32_t dim, int32_t nlist, int32_t device, fivf_index** out);
void fivf_destroy(fivf_index* h);

int64_t fivf_ntotal(const fivf_index* h);
int32_t fivf_nlist(const fivf_index* h);
int32_t fivf_dim(const fivf_index* h);
int32_t fivf_is_trained(const fivf_index* h);

/* index.train(x): k-means of the coarse quantizer on the GPU (niter < 0: faiss's IVF default, 10).
 * Fails like faiss when n < nlist.  Centroids are deterministic in (x, seed) but NOT those faiss
 * would produce (different random stream). */
int fivf_train(fivf_index* h, const float* x, int64_t n, int32_t niter, uint64_t seed);
/* Import / export of the coarse quantizer: centroids [nlist, dim] fp32 (host). */
int fivf_set_centroids(fivf_index* h, const float* centroids);
int fivf_get_centroids(fivf_index* h, float* out);

/* index.add(x) (ids == NULL: sequential from ntotal) / index.add_with_ids(x, ids). */
int fivf_add(fivf_index* h, const float* x, int64_t n, const int64_t* ids);
/* Append rows to GIVEN lists (an index file keeps the lists it was built with). */
int fivf_add_preassigned(fivf_index* h, const float* x, int64_t n, const int64_t* ids, const int32_t* list_numbers);
/* quantizer.assign(x): nearest centroid of each row. */
int fivf_assign(fivf_index* h, const float* x, int64_t n, int32_t* out);

/* invlists: sizes [nlist]; codes [ntotal, dim] and ids [ntotal] concatenated in list order
 * (either output may be NULL). */
int fivf_list_sizes(fivf_index* h, int64_t* out);
int fivf_export_lists(fivf_index* h, float* codes, int64_t* ids);

/* index.nprobe = nprobe; D, I = index.search(queries, k).  Host buffers: queries [nq, dim] fp32,
 * distances [nq, k] fp32 (squared L2, ascending), ids [nq, k] int64.  Blocking. */
int fivf_search(fivf_index* h, const float* queries, int64_t nq, int32_t k, int32_t nprobe,
                float* distances, int64_t* ids);
/* Same with device buffers on the index's GPU (co-located embedder, benchmarks).  Blocking: the
 * sizes of the list-major work lists are read back to size the scan launches. */
int fivf_search_device(fivf_index* h, const float* queries_dev, int64_t nq, int32_t k, int32_t nprobe,
                       float* distances_dev, int64_t* ids_dev);

/* options: "force_safe", "live_thresholds", "max_queries_per_tile" (4 or 16), "profile_events"
 * (1: CUDA events around the scan launches, read back as "scan_ns"), "tensor_min_queries" (a list
 * probed by at least this many queries of a batch is scored on the tensor cores; default 1, 0 = never),
 * "tensor_format" (0, default: the tensor-core filter reads an fp16 copy of the lists -- built on first
 * use, +50 % device memory -- whenever every component of the lists and of the batch's queries is
 * inside fp16's range; 1: always tf32 straight from the fp32 rows.  Results are identical either way:
 * survivors are rescored exactly in fp32),
 * "device_plan" (default 1: for batches of >= 1024 (query, probe) pairs the probe table is inverted into
 * work lists by GPU kernels and only their sizes are read back; 2: for every batch; 0: always on the
 * host, which safe mode does anyway).
 * "coarse_tensor" (1 default: batches of >= 512 queries, 2: every batch, nlist >= 256: the coarse quantizer runs on the tensor cores with
 * fp32-accurate scores -- every fp32 value split into three bf16 terms, the six significant products
 * laid side by side along K (K' = 6 dim), K1's dense mode, fp32 accumulation; 0: fp32 FFMA GEMM).
 * stats of the last search: "coarse_tensor" (1: used), "launches", "phases", "tiles", "tiles16", "rows_scanned",
 * "pairs_scored", "distinct_rows" (rows of the lists probed by at least one query), "scan_ns",
 * "tensor_items", "tensor_pairs", "crowded_window" (1: the tf32 filter's exactness window overflowed
 * for some query and the batch was redone with the fp32 filter, then -- clusters of more than 2k + 64
 * duplicates -- in safe mode with a window of half the list capacity), "crowded_window_unresolved"
 * (1: even that window overflowed: over 8192 rows tie within fp32 rounding for some query, whose
 * result is then exact only up to that rounding), "overflow", "candidate_capacity", "plan_us", "scan16_ctas_per_sm",
 * "device_plan" (1: the last search was planned on the device), "tensor_f16" (1: the last search's
 * tensor-core filter ran on the fp16 copy). */
int fivf_set_option(fivf_index* h, const char* key, int64_t value);
int64_t fivf_get_stat(const fivf_index* h, const char* key);

#ifdef __cplusplus
}
#endif
#endif /* IVFFLAT_B200_H */
