/*
 * flatip_b200.h -- C ABI of the B200-native exact flat inner-product search library
 * (libflatip_b200.so).  Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * The library is the drop-in for the ONE native call on the reference's retrieval hot path:
 *
 *     distances, indices = self._index.search(embeddings, k)
 *         /root/reference/src/pipeline/components/faiss_store.py:152   (also the warm-up, :106)
 *
 * i.e. faiss-cpu's SWIG entry IndexFlatIP.search(n, x, k, distances, labels), plus the pieces of
 * the faiss.Index surface the reference touches around it (read_index -> fip_create + fip_add_*,
 * index.ntotal -> fip_ntotal; faiss_store.py:66-69,73,182).  Each entry point below names the
 * reference interface it replaces.  All functions return 0 on success or a negative fip_status;
 * fip_last_error() returns a thread-local message for the last failure on the calling thread.
 * Every function is safe to call from several host threads at once; calls on the same index are
 * serialised internally (the reference calls search from a thread pool,
 * /root/reference/src/pipeline/services/retrieval/api.py:348-349).
 */
#ifndef FLATIP_B200_H_
#define FLATIP_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fip_index fip_index;

enum fip_dtype { FIP_F32 = 0, FIP_F16 = 1, FIP_BF16 = 2 };

enum fip_status {
    FIP_OK = 0,
    FIP_ERR_INVALID = -1,   /* bad argument (shape, k, dtype, null pointer)            */
    FIP_ERR_CUDA = -2,      /* a CUDA call or kernel failed; see fip_last_error()        */
    FIP_ERR_NOMEM = -3,     /* device or host allocation failed                           */
    FIP_ERR_NODEVICE = -4,  /* no usable sm_100 device -- there is NO CPU fallback        */
    FIP_ERR_STATE = -5      /* call not valid in the index's current state                */
};

enum fip_variant {
    FIP_VARIANT_AUTO = 0,    /* chosen from arithmetic intensity (batch, dtype)           */
    FIP_VARIANT_STREAM = 1,  /* K2: 128-bit vectorised CUDA-core streaming scan           */
    FIP_VARIANT_TENSOR = 2   /* K1: TMA + tcgen05/TMEM GEMM with fused threshold filter   */
};

typedef struct fip_config {
    int32_t dim;           /* vector dimension (faiss Index::d; settings.faiss_dim)      */
    int32_t dtype;         /* storage dtype of the corpus: fip_dtype                     */
    int32_t device;        /* CUDA device ordinal                                        */
    int32_t reserved;
    int64_t capacity_rows; /* rows to reserve up front (0 = grow on add)                 */
    int64_t id_base;       /* global id of local row 0 (row-sharded corpora)             */
} fip_config;

/* faiss.read_index / IndexFlatIP(d) constructor  (faiss_store.py:66-69). */
int fip_create(const fip_config* cfg, fip_index** out);
/* FAISSStore.unload -> index destruction (faiss_store.py:160-171). */
void fip_destroy(fip_index* idx);

/* faiss Index::add(n, x): append n fp32 host rows, rounded (RNE) to the storage dtype. */
int fip_add_host_f32(fip_index* idx, const float* rows, int64_t n);
/* Same for host rows in any fip_dtype: fp32 rows are rounded to the storage dtype, rows already in
 * the storage dtype are uploaded as they are (a 16-bit corpus file crosses PCIe as 16 bits), 16-bit
 * rows into an fp32 index are widened on the device.  Pinned, double-buffered staging. */
int fip_add_host(fip_index* idx, const void* rows, int64_t n, int32_t dtype);
/* Same, rows already on the device in `dtype` (fp32 is converted, equal dtypes are copied). */
int fip_add_device(fip_index* idx, const void* rows_dev, int64_t n, int32_t dtype);
/* Append n synthetic rows generated on the device: global rows [global_row0, global_row0+n) of
 * the Philox corpus defined in rag-inference-pipeline_b200/synthetic.py (stands in for
 * /root/reference/scripts/create_test_docs.py:87-96). */
int fip_add_synthetic(fip_index* idx, int64_t n, uint64_t seed, int64_t global_row0);
/* Drop all rows (keeps the allocation). */
int fip_reset(fip_index* idx);
/* faiss Index::reconstruct_n: copy rows [row0, row0+n) to the host, widened to fp32. */
int fip_reconstruct(fip_index* idx, int64_t row0, int64_t n, float* out_host);

/* faiss Index::ntotal (faiss_store.py:73,182). */
int64_t fip_ntotal(const fip_index* idx);
int32_t fip_dim(const fip_index* idx);
int32_t fip_dtype_of(const fip_index* idx);

/*
 * faiss IndexFlatIP::search(n, x, k, distances, labels)  (faiss_store.py:152).
 * queries: host fp32 [nq, dim] row-major; distances: host fp32 [nq, k]; ids: host int64 [nq, k].
 * Each row is best-first under (score descending, id ascending); unfilled slots hold
 * -FLT_MAX / -1 like faiss.  Blocking; host<->device copies are done inside the call.
 */
int fip_search(fip_index* idx, const float* queries, int64_t nq, int32_t k, float* distances,
               int64_t* ids);
/* Same contract with device-resident buffers (fp32 queries, fp32 distances, int64 ids) and a
 * caller-supplied cudaStream_t (0 = the index's own stream).  Blocking on that stream. */
int fip_search_device(fip_index* idx, const float* queries_dev, int64_t nq, int32_t k,
                      float* distances_dev, int64_t* ids_dev, void* cuda_stream);

/*
 * Asynchronous form: enqueues the whole search on `cuda_stream` and returns without waiting.  A
 * search can, rarely (adversarial row order, near-tie clusters tighter than the tensor-core
 * filter's resolution), need some queries redone; the blocking forms do that before they return,
 * this one only records it on the device.  fip_search_status synchronises the stream of the last
 * asynchronous search and reports whether ANY asynchronous search since the previous status call
 * was flagged (bit 0: a candidate list overflowed, bit 1: a query could not be certified exact);
 * the results of a flagged search must be recomputed with fip_search_device.  All asynchronous
 * searches on one index must be ordered on one stream.  fip_search_flags_device returns the device
 * address of the two summary words of the search enqueued last (consumed by fip_exchange_merge, so
 * the status of every shard travels with its rows).
 */
int fip_search_device_async(fip_index* idx, const float* queries_dev, int64_t nq, int32_t k,
                            float* distances_dev, int64_t* ids_dev, void* cuda_stream);
int fip_search_status(fip_index* idx, int32_t* flagged);
const uint32_t* fip_search_flags_device(const fip_index* idx);

/*
 * The same exact search for callers that must not touch the CUDA allocator.  No counterpart in the
 * reference (faiss-cpu owns its own host memory); it exists because of where a drop-in GPU retriever
 * ends up living: next to an LLM server that has already taken almost all of HBM for its KV cache
 * (a few MB of cudaMalloc at search time then fails), or inside a serving process whose pooling
 * allocator owns the device and bans raw cudaMalloc / cudaFree because both stall every stream of
 * the GPU.  Such a caller keeps the corpus and the result buffers in memory it already owns and
 * passes plain device pointers; the library's own workspace (candidate lists, thresholds, flag
 * words) sits in ONE static __device__ array of 8 455 680 bytes (~8.1 MiB) that every GPU loading
 * the library holds until the process exits.  The call makes no cudaMalloc / cudaFree or any other
 * allocating CUDA call.
 *   rows_dev      [n_rows, dim] in `dtype` (fip_dtype), row-major, 16-byte aligned, row size a
 *                 multiple of 16 bytes, n_rows < 2^31
 *   queries_dev   fp32 [nq, dim]; distances_dev fp32 [nq, k]; ids_dev int64 [nq, k] (id = id_base + row)
 *   cuda_stream   cudaStream_t to run on (0 = the default stream); blocking on it
 * Limits: nq <= 64 and k <= 2048 per call -- anything larger is refused with FIP_ERR_INVALID and a
 * message naming the limit (split the batch, or use fip_search_device).  Results: same order, padding
 * and bits as fip_search_device with option "variant" = FIP_VARIANT_STREAM.
 * All calls on one GPU share the array and find the previous call's contents in it; each call
 * resets what has to start clean.  Calls whose kernels overlapped on a GPU would corrupt each
 * other's candidate lists, so the entry point holds a per-device lock until its stream has drained:
 * threads of one process wait for one another.  The first call on a device also looks up the
 * array's device address; make that call before starting any stream capture.
 */
int fip_search_rows_static(int32_t device, const void* rows_dev, int64_t n_rows, int32_t dim,
                           int32_t dtype, int64_t id_base, const float* queries_dev, int64_t nq,
                           int32_t k, float* distances_dev, int64_t* ids_dev, void* cuda_stream);

/*
 * Cross-shard merge (K3): in_d / in_i are device buffers [g, nq, k_in] as an all-gather of the
 * per-shard search results leaves them (shard r holds lower ids than shard r+1); writes the
 * global best k_out per query to out_d / out_i [nq, k_out].  No faiss-cpu equivalent in the
 * reference (single process); it is what IndexShards::search does in upstream faiss.
 */
int fip_merge_device(int32_t device, const float* in_d, const int64_t* in_i, int32_t g, int64_t nq,
                     int32_t k_in, int32_t k_out, float* out_d, int64_t* out_i, void* cuda_stream);
/* Same merge over a PACKED exchange buffer: shard r's distances start at in_d + r * stride_d and its
 * ids at in_i + r * stride_i (strides in elements).  Lets a rank ship distances and ids in ONE
 * all-gather (one buffer per rank: [nq * k_in fp32 | pad to 16 bytes | nq * k_in int64]). */
int fip_merge_strided_device(int32_t device, const float* in_d, const int64_t* in_i, int64_t stride_d,
                             int64_t stride_i, int32_t g, int64_t nq, int32_t k_in, int32_t k_out,
                             float* out_d, int64_t* out_i, void* cuda_stream);

/*
 * Fused exchange + merge over NVLink peer memory (K4, csrc/exchange.cuh): the one-kernel
 * replacement for "all-gather, then fip_merge_device".  One fip_exchange per rank; every rank
 * allocates a buffer its peers can write (CUDA IPC between processes, peer access inside one
 * process), and fip_exchange_merge launches ONE kernel that stores this rank's local [nq, k]
 * result into every rank's buffer, waits per query for the other ranks' rows and merges them.
 * Like a collective, every rank of the group must make the same sequence of fip_exchange_merge
 * calls (same nq, k).  No counterpart in the reference (single process).
 *   create        allocate for up to nq_max queries x k_max results, world <= 16 ranks
 *   ipc_handle    64-byte handle of the local buffer (to be all-gathered by the host runtime)
 *   connect_ipc   map the peers' buffers from `world` concatenated handles (own slot ignored)
 *   connect_local same for exchanges living in THIS process (one per device, or several on one
 *                 device for tests): all[r] is rank r's exchange
 *   merge         asynchronous on `cuda_stream`; d_loc / i_loc / out_* are device buffers [nq, k]
 */
#define FIP_IPC_HANDLE_BYTES 64
typedef struct fip_exchange fip_exchange;
int fip_exchange_create(int32_t device, int32_t world, int32_t rank, int64_t nq_max, int32_t k_max,
                        fip_exchange** out);
int fip_exchange_ipc_handle(fip_exchange* x, void* handle_out);
int fip_exchange_connect_ipc(fip_exchange* x, const void* all_handles);
int fip_exchange_connect_local(fip_exchange* x, fip_exchange* const* all);
int fip_exchange_merge(fip_exchange* x, const float* d_loc, const int64_t* i_loc, int64_t nq, int32_t k,
                       float* out_d, int64_t* out_i, void* cuda_stream);
/* Same, and the shards' search statuses travel with the rows: flags_loc_dev = this rank's
 * fip_search_flags_device() (or null), *out_status_dev (device word, OR-accumulated, may be null)
 * receives bit 0 / bit 1 when ANY rank's local search overflowed / could not be certified. */
int fip_exchange_merge_status(fip_exchange* x, const float* d_loc, const int64_t* i_loc, int64_t nq,
                              int32_t k, float* out_d, int64_t* out_i, const uint32_t* flags_loc_dev,
                              uint32_t* out_status_dev, void* cuda_stream);
void fip_exchange_destroy(fip_exchange* x);

/*
 * fip_group: ONE corpus row-sharded over several GPUs of THIS process behind one blocking search
 * call -- what lets the reference's single in-process call site
 *     registry.get("faiss_store").search(embeddings, k)
 *     (/root/reference/src/pipeline/services/retrieval/api.py:383-389)
 * use every GPU of the box (north_star: "row-sharded across the 8 GPUs", same component API).
 * Shard r lives on devices[r] and owns global rows [r * per, (r + 1) * per), per =
 * ceil(planned_rows / n_devices); rows are added in global order.  fip_group_search: one host
 * thread + stream per device, local search, K4 exchange + merge over NVLink peer memory, the global
 * result is returned from devices[0].  Same contract, coalescing and error behaviour as fip_search.
 * cfg->device is ignored; cfg->id_base is the global id of row 0.
 */
typedef struct fip_group fip_group;
int fip_group_create(const fip_config* cfg, const int32_t* devices, int32_t n_devices,
                     int64_t planned_rows, fip_group** out);
/* Same with shard sizes proportional to `weights` (n_devices positive numbers, e.g. the rates
 * fip_calibrate_devices measured; null = equal shards): a search ends with the slowest GPU, and the
 * GPUs of one box differ by several per cent in the clock their power cap lets them sustain. */
int fip_group_create_weighted(const fip_config* cfg, const int32_t* devices, int32_t n_devices,
                              int64_t planned_rows, const double* weights, fip_group** out);
/* Sustained search throughput (corpus rows per second of scan-kernel time) of every listed device,
 * all measured at the same time on a `rows`-row synthetic corpus, `batch` queries, k results, for
 * about `seconds` seconds (the first 30 % warms the chips up). */
int fip_calibrate_devices(const int32_t* devices, int32_t n_devices, int32_t dim, int32_t dtype,
                          int64_t rows, int32_t batch, int32_t k, double seconds, double* rates);
void fip_group_destroy(fip_group* g);
int32_t fip_group_size(const fip_group* g);
/* Shard r as a plain index (options, statistics, reconstruct; do not add rows through it). */
fip_index* fip_group_shard(fip_group* g, int32_t r);
int64_t fip_group_ntotal(const fip_group* g);
/* Append the next n global rows: generated on the owning devices / uploaded from host rows in
 * `dtype` (every device pulls its own slice through its own pinned staging, in parallel). */
int fip_group_add_synthetic(fip_group* g, int64_t n, uint64_t seed);
int fip_group_add_host(fip_group* g, const void* rows, int64_t n, int32_t dtype);
int fip_group_reset(fip_group* g);
int fip_group_search(fip_group* g, const float* queries, int64_t nq, int32_t k, float* distances,
                     int64_t* ids);
/* Options are applied to every shard ("coalesce*" at the group); statistics are summed over the
 * shards' last searches, plus "devices", "rows_per_shard", "shard_begin_<r>" (first global row of shard r; r = devices: the planned total), "searches", "flagged_searches" (searches
 * in which some shard had to redo a query, costing a second exchange), "exchange_launches". */
int fip_group_set_option(fip_group* g, const char* key, int64_t value);
int64_t fip_group_get_stat(const fip_group* g, const char* key);

/* Options: "variant" (fip_variant), "stream_max_batch" (largest batch the AUTO rule gives to the
 * streaming kernel; 0 = default, by arithmetic intensity: batch <= sizeof(storage element)), "query_split" (0 auto / 1 / 2 sixteen-bit terms per query
 * in the tensor-core variant), "cta_pair" (0 auto / 1 single-CTA MMAs / 2 cta_group::2 CTA pairs),
 * "certify" (1 default / 0: skip the exactness certificate of the tensor-core variant),
 * "filter_copy" (fp32 storage only; 1 default: the tensor-core variant FILTERS on an fp16 copy of the
 * corpus -- built on first search after a change, +50 % device memory, skipped when it does not fit or
 * a row leaves fp16's range -- and rescoring stays exact on the fp32 rows; 0: tf32 on the fp32 rows),
 * "live_thresholds" (1 default: a filter launch raises its own per-query thresholds from a score
 * histogram, so one launch covers the shard; 0: thresholds move only between geometric phases),
 * "coalesce" (1 default / 0) and "coalesce_max_batch" (default 4096): fip_search calls that arrive
 * from other threads while a search is running are queued and served together as ONE batch, "profile_events" (0/1: time every scan launch with CUDA events),
 * "force_safe_mode" (0/1, testing the overflow fallback), "defer_epilogue" (0 auto: k >= 256 / 1 off /
 * 2 on: the tensor-core scan releases an accumulator right after the filter pass and writes the
 * survivors out from a shared-memory staging area afterwards, csrc/tc_scan.cuh), "fused_rescore"
 * (1 default: the final select pass recomputes the survivors' scores exactly itself; 0: separate
 * rescoring kernel, same bits), "dedupe" (1 default:
 * bit-identical queries inside a COALESCED host batch are searched once, on the device; 2: in every
 * host batch; 0: off), "debug_counts" (0/1: read the candidate-list lengths back after every filter
 * launch -> stats "list_entries_after_filter" / "list_max_after_filter"). */
int fip_set_option(fip_index* idx, const char* key, int64_t value);
/* Counters of the last search: "launches", "variant", "phases", "overflow_queries",
 * "uncertified_queries" (queries the tensor-core filter could not certify exact and that were redone
 * with a tighter variant), "stream_redo_queries" (of those and the overflowed ones, how many ended
 * on the exact streaming kernel), "query_split_used", "filter_f16" (1: the last search filtered on the
 * fp16 copy of an fp32 corpus),
 * "candidate_capacity", "cta_group" (1 or 2: MMA CTA group the tensor variant used); with
 * "profile_events" on, over the searches COLLECTED by the last blocking search / fip_search_status
 * (asynchronous searches accumulate until their status is read): "scan_searches" (how many),
 * "scan_time_ns" / "scan_launches" (device time and count of their scan kernels), "scan_ns_<i>" /
 * "gap_ns_<i>" (summed over those searches per scan launch index i: device time, time to the next
 * scan) and "scan_rows_<i>" (corpus rows of launch i); cumulative: "total_launches", "total_searches",
 * "host_calls" / "host_batches" (fip_search calls and the batches they were served in),
 * "coalesced_calls" (calls that shared a batch with another call); "last_batch_calls";
 * "host_queries" / "distinct_queries" (queries received by fip_search and how many were actually
 * searched after duplicate elimination); "deferred_epilogue" (1: the last search used it). */
int64_t fip_get_stat(const fip_index* idx, const char* key);

const char* fip_last_error(void);
/* Library/ABI version, and the SM architecture the kernels were compiled for (100). */
int32_t fip_abi_version(void);
int32_t fip_compiled_sm(void);

#ifdef __cplusplus
}
#endif
#endif /* FLATIP_B200_H_ */
