/*
 * lance_b200.h -- C ABI of the B200-native IVF_PQ query path (liblance_b200.so).
 *
 * This is the drop-in boundary of SURVEY.md 8(b): every entry point replaces one method of
 * Lance's `trait VectorIndex` (rust/lance-index/src/vector.rs:117-217) or the flat-kNN free
 * function `compute_distance` (rust/lance-index/src/vector/flat.rs:28), batched over Q
 * queries.  Signatures carry only plain pointers and sizes (no torch / Arrow types).
 *
 * Conventions
 *  - return 0 = OK, negative = error (LB200_E*); message via lb200_last_error() (thread-local).
 *  - pointers are HOST memory unless the function name ends in `_device`.
 *  - nothing is retained after return except what lb200_index_create copies to the GPU.
 *  - a handle may be used from several host threads at once: every call leases one of the handle's call
 *    contexts (stream + scratch; LB200_CTX_POOL of them, default 16), so concurrent calls overlap on the
 *    GPU -- the exec nodes keep up to get_num_compute_intensive_cpus() search_in_partition calls in flight
 *    per query (rust/lance/src/io/exec/knn.rs:606-634).  GPU work of consecutive users of a context is
 *    ordered by an event, also for the asynchronous *_device entry points on caller streams.
 *  - there is NO CPU fallback: without a usable CUDA device every call fails with
 *    LB200_ECUDA.
 *
 * Limits of this build (LB200_EUNSUPPORTED beyond them): PQ code width 8 or 4 bits; nlist <= 8192;
 * num_sub_vectors <= 192 (the bank-conflict-free scan for multiples of 16 up to 96);
 * k * max(refine_factor,1) <= 512; N < 2^32 rows; Q * nprobes < 2^31 per call;
 * nprobes * k' up to (16384/k')^2 lists (two-level merge above 16384 candidates per query).
 *
 * Numerics: all distances are IEEE f32 computed in the reference's own operation order
 * (see DESIGN.md "Numerics"), so results are bit-identical to the CPU reference wherever the
 * reference itself is deterministic; equal distances are ordered by row id.
 */
#ifndef LANCE_B200_H
#define LANCE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lb200_index lb200_index; /* opaque, device-resident IVF_PQ index */

/* DistanceType, rust/lance-linalg/src/distance.rs:36-43 */
enum { LB200_L2 = 0, LB200_COSINE = 1, LB200_DOT = 2 };

enum {
  LB200_OK = 0,
  LB200_EINVAL = -1,   /* Error::InvalidInput  (lance-core/src/error.rs:19)  */
  LB200_ECUDA = -2,    /* CUDA runtime failure / no device                   */
  LB200_ENOMEM = -3,
  LB200_EUNSUPPORTED = -4 /* Error::Index "not supported" (e.g. num_bits != 8) */
};

/* Replaces IVFIndex::try_new + load_partition + ProductQuantizationStorage::new
 * (rust/lance/src/index/vector/ivf/v2.rs:110-275, rust/lance-index/src/vector/storage.rs:271-298,
 * pq/storage.rs:147-211): eager upload of the whole index.
 *   centroids    [nlist][dim] f32 row-major                (IvfModel, ivf/storage.rs)
 *   codebook     [M][256][dim/M] f32                       (pq/utils.rs:54-71)
 *   part_offsets [nlist+1] rows, prefix sum                (pb::Ivf offsets/lengths)
 *   codes        concat over p of [M][n_p] u8, transposed  (pq/storage.rs:357-377,455-470)
 *   row_ids      [N] u64 in index order                    (`_rowid` column of auxiliary.idx)
 *   raw_vectors  [N][dim] f32 in index order, or NULL      (only for device-side refine)
 * num_bits = 8: as above.
 * num_bits = 4 (pq/distance.rs:138-229): codebook [M][16][dim/M]; codes concat over p of
 *   [M/2][n_p] bytes, low nibble = sub-space 2j, high nibble = 2j+1 (pq.rs:166-170); M even.
 * num_bits = 32: IVF_FLAT (IVFIndex<FlatIndex, FlatQuantizer>, flat/storage.rs:315-375): no
 *   quantizer -- codebook, codes and num_sub_vectors are ignored and raw_vectors is REQUIRED;
 *   search returns exact distances on the raw rows of the probed partitions.
 * 4-bit and IVF_FLAT handles take every search entry point, prefilter bitmasks and the shard contract (merged k' list
 * + exact distances, d_out_exact) included.  With a prefilter a 4-bit handle computes each allowed row from the
 * UNQUANTISED f32 table (PQDistCalculator::distance, pq/storage.rs:602-617), as FlatIndex::search does
 * (flat/index.rs:106-122).  lb200_shard_* / lb200_group_* take all three handle kinds (the cross-GPU pruning bounds are
 * an 8-bit-scan feature; 4-bit and IVF_FLAT shards exchange probe lists and candidate lists only). */
int32_t lb200_index_create(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist,
                           const float* centroids, uint32_t num_sub_vectors, uint32_t num_bits,
                           const float* codebook, const uint64_t* part_offsets,
                           const uint8_t* codes, const uint64_t* row_ids,
                           const float* raw_vectors_or_null, lb200_index** out);
void lb200_index_destroy(lb200_index* ix);

/* Native reader of the reference's legacy IVF_PQ index file (`_indices/<uuid>/index.idx`, Lance file
 * version 0.1).  Replaces open_index_proto + open_vector_index + PQIndex::load
 * (rust/lance/src/index.rs:169-181,774-786, rust/lance/src/index/vector.rs:446-513,
 * rust/lance/src/index/vector/pq.rs:292-330): parses the file on the host into the arrays
 * lb200_index_create takes (codes already transposed per partition).  Pure host code: usable
 * without a GPU.  The current (v3) index -- index.idx + auxiliary.idx -- has its own entry points below. */
typedef struct lb200_index_file lb200_index_file; /* opaque, host memory */
typedef struct {
  int32_t metric;              /* LB200_L2 / COSINE / DOT (pb::VectorMetricType)                 */
  uint32_t dim, nlist, num_sub_vectors, num_bits, spec_version;
  uint64_t num_rows, dataset_version;
  const float* centroids;      /* [nlist][dim]                                                   */
  const float* codebook;       /* [M][256][dim/M]                                                */
  const uint64_t* part_offsets;/* [nlist+1]                                                      */
  const uint8_t* codes;        /* concat over p of [M][n_p]                                      */
  const uint64_t* row_ids;     /* [num_rows]                                                     */
  const char* name;            /* pb::Index.name                                                 */
  const char* column;          /* pb::Index.columns[0]                                           */
} lb200_index_file_info;       /* pointers live until lb200_index_file_close                     */
int32_t lb200_index_file_open(const char* path, lb200_index_file** out);
int32_t lb200_index_file_open_bytes(const void* bytes, uint64_t size, lb200_index_file** out);
int32_t lb200_index_file_get_info(const lb200_index_file* f, lb200_index_file_info* info);
void lb200_index_file_close(lb200_index_file* f);
/* lb200_index_file_open + lb200_index_create in one call (== Dataset::open_vector_index for a
 * legacy IVF_PQ index).  raw_vectors [num_rows][dim] in the FILE's row order, or NULL. */
int32_t lb200_index_open(int32_t device_id, const char* path, const float* raw_vectors_or_null,
                         lb200_index** out);

/* Native reader of the CURRENT IVF_PQ index (IndexFileVersion::V3): `_indices/<uuid>/index.idx` (IVF model with the
 * centroids, `lance:index` / `lance:ivf` schema metadata) + `auxiliary.idx` (`_rowid` and `__pq_code` columns in
 * partition order, the PQ codebook in `storage_metadata`), both Lance-v2 file containers of format 2.0.  Replaces
 * IVFIndex::try_new + IvfQuantizationStorage::try_new / load_partition (rust/lance/src/index/vector/ivf/v2.rs:110-202,
 * rust/lance-index/src/vector/storage.rs:210-298); the writer it mirrors is rust/lance/src/index/vector/builder.rs:
 * 624-770.  Pure host code.  Only the page encodings that writer produces for these columns are accepted (flat,
 * uncompressed, no nulls); anything else is LB200_EUNSUPPORTED.  PARITY UNPINNED: the reference tree holds no v3
 * index file; see csrc/api_file_v3.inl.  lb200_index_file_get_info / _close serve both file kinds. */
int32_t lb200_index_file_open_v3(const char* index_idx_path, const char* auxiliary_idx_path, lb200_index_file** out);
int32_t lb200_index_file_open_v3_bytes(const void* index_idx, uint64_t index_size, const void* auxiliary_idx,
                                       uint64_t auxiliary_size, lb200_index_file** out);
int32_t lb200_index_open_v3(int32_t device_id, const char* index_idx_path, const char* auxiliary_idx_path,
                            const float* raw_vectors_or_null, lb200_index** out);

/* VectorIndex::find_partitions (vector.rs:146; ivf/v2.rs:403-411 ->
 * kmeans_find_partitions, rust/lance-linalg/src/kmeans.rs:624-644), batched.
 * As in the reference, the key is used AS GIVEN: for a cosine index the caller passes the
 * already-normalised key (exec node does that, rust/lance/src/io/exec/knn.rs:366-369).
 * out_part_ids [Q][min(nprobes,nlist)] ascending by (distance, id); out_part_dists optional. */
int32_t lb200_find_partitions(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes,
                              uint32_t* out_part_ids, float* out_part_dists_or_null);

/* VectorIndex::search_in_partition (vector.rs:165; ivf/v2.rs:414-433 -> FlatIndex::search,
 * rust/lance-index/src/vector/flat/index.rs:76-137) for explicit (query, partition) pairs.
 * part_ids [Q][P]; key used as given (already normalised for cosine).
 * lower/upper: distance range [lower, upper), NULL = unbounded (flat/index.rs:97-101).
 * part_bitmask_or_null: per PARTITION pointer (nlist entries, NULL entry = all rows) to a
 *   bitmask over that partition's rows, bit i of word i/64 set = row allowed (PreFilter,
 *   lance-index/src/prefilter.rs:26-51).
 * out_dist/out_rowid [Q][P][k_prime] ascending by (distance,row id); out_count [Q][P]. */
int32_t lb200_search_partitions(lb200_index* ix, const float* queries, uint32_t Q,
                                const uint32_t* part_ids, uint32_t P, uint32_t k_prime,
                                const float* lower_or_null, const float* upper_or_null,
                                const uint64_t* const* part_bitmask_or_null, float* out_dist,
                                uint64_t* out_rowid, uint32_t* out_count);

/* A PreFilter evaluated once (DatasetPreFilter::wait_for_ready, rust/lance/src/index/prefilter.rs:227-290)
 * and shared by all search_in_partition calls of a query, as the Arc<dyn PreFilter> the exec node passes
 * to every partition (knn.rs:621): the per-partition bitmasks are uploaded once and stay on the device.
 * part_bitmask: [nlist] host pointers as for lb200_search_partitions (NULL entry = all rows allowed). */
typedef struct lb200_prefilter lb200_prefilter;
int32_t lb200_prefilter_create(lb200_index* ix, const uint64_t* const* part_bitmask, lb200_prefilter** out);
void lb200_prefilter_destroy(lb200_prefilter* pf);
int32_t lb200_search_partitions_filtered(lb200_index* ix, const float* queries, uint32_t Q,
                                         const uint32_t* part_ids, uint32_t P, uint32_t k_prime,
                                         const float* lower_or_null, const float* upper_or_null,
                                         const lb200_prefilter* prefilter_or_null, float* out_dist,
                                         uint64_t* out_rowid, uint32_t* out_count);

/* VectorIndex::search (vector.rs:133; ivf/v2.rs:368-401): normalise (cosine) ->
 * find_partitions -> search_in_partition per probe -> merge to k' = k*max(refine_factor,1).
 * If refine_factor > 0 and the index holds raw vectors, also runs the scanner's refine
 * (rust/lance/src/dataset/scanner.rs:1706-1721,2066-2136): exact index-metric distance on
 * raw vectors with the original query, bounds re-applied, top-k.
 * Output stride is k (refined) or k' (not refined, refine_factor==0 -> k'=k); out_count [Q]. */
int32_t lb200_search(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes,
                     uint32_t k, uint32_t refine_factor, const float* lower_or_null,
                     const float* upper_or_null, float* out_dist, uint64_t* out_rowid,
                     uint32_t* out_count);

/* Same as lb200_search with DEVICE pointers on the index's device, asynchronous on `stream`
 * (a cudaStream_t; NULL = the handle's own stream, synchronised before return).
 * Shard mode (multi-GPU): when d_out_exact_or_null != NULL the call stops before
 * the final cut and returns the k' merged ANN candidates of THIS shard -- d_out_dist [Q][k']
 * PQ distances, d_out_rowid [Q][k'], d_out_count [Q] -- plus their exact (refine) distances
 * in d_out_exact [Q][k'] (NaN when refine_factor==0 or no raw vectors), so that
 * lb200_merge_topk_device can reproduce the single-index result exactly. */
int32_t lb200_search_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes,
                            uint32_t k, uint32_t refine_factor, const float* lower_or_null,
                            const float* upper_or_null, float* d_out_dist, uint64_t* d_out_rowid,
                            uint32_t* d_out_count, float* d_out_exact_or_null, void* stream);

/* Multi-GPU helpers: the coarse step can be sharded over the ranks by QUERIES
 * (each rank selects probes for Q/G queries, the [Q][P] probe lists are all-gathered) and the
 * partition-sharded scan then runs with the probe lists given.
 * lb200_find_partitions_device == lb200_find_partitions on device pointers, asynchronous on
 *   `stream`; normalize_if_cosine != 0 applies the VectorIndex::search normalisation
 *   (ivf/v2.rs:370-373) first, so raw queries can be passed.
 * lb200_search_probed_device == lb200_search_device with find_partitions replaced by the given
 *   d_part_ids [Q][P] (valid partition ids, as produced by lb200_find_partitions_device). */
int32_t lb200_find_partitions_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes,
                                     int32_t normalize_if_cosine, uint32_t* d_out_part_ids, void* stream);
int32_t lb200_search_probed_device(lb200_index* ix, const float* d_queries, uint32_t Q,
                                   const uint32_t* d_part_ids, uint32_t P, uint32_t k,
                                   uint32_t refine_factor, const float* lower_or_null,
                                   const float* upper_or_null, float* d_out_dist, uint64_t* d_out_rowid,
                                   uint32_t* d_out_count, float* d_out_exact_or_null, void* stream);

/* Shared pruning bounds between the shards of one index (multi-GPU).
 * Inside one index the scan keeps, per query, the running k'-th best distance over everything
 * scanned so far and skips rows above it.  On a partition-sharded index every shard would have
 * to find that bound again on its own (G times the cold starts); instead each shard publishes the
 * bounds it finds into every peer's table with remote atomic-min over NVLink peer memory, from
 * inside the scan kernel.  A bound found on ANY shard is a valid bound for the merged result, so
 * the merged top-k' stays exactly the single-index one (the shard-local lists may be shorter).
 *   export: allocates this shard's table for up to max_queries queries per batch and returns a
 *           64-byte CUDA IPC handle (for peers in other processes) and the device address (for
 *           peers in the same process).
 *   attach: world shards, this one is `self`.  handles = [world][64] bytes as exported (entry
 *           `self` ignored), or NULL with same_process_ptrs = [world] device addresses.
 *   Contract: all shards run the same sequence of shard-mode searches, and a cross-rank
 *   synchronisation (the candidate exchange) separates consecutive batches; tables are
 *   double-buffered by batch parity so a late peer never writes into the next batch's bounds.
 *   Only shard-mode searches (d_out_exact != NULL) with Q <= max_queries use the tables. */
int32_t lb200_peer_thresholds_export(lb200_index* ix, uint32_t max_queries, void* handle_out_64,
                                     uint64_t* dev_ptr_out);
int32_t lb200_peer_thresholds_attach(lb200_index* ix, uint32_t world, uint32_t self, const void* handles_or_null,
                                     const uint64_t* same_process_ptrs_or_null);
int32_t lb200_peer_thresholds_detach(lb200_index* ix);

/* The multi-GPU form of lb200_search: IVF partitions sharded over the GPUs of one box, the
 * whole exchange inside the library -- no collective library, no host round trip per batch.  Per batch a shard
 * selects the probes of its slice of the queries, stores them into every peer's exchange block over NVLink peer
 * memory, scans the probed partitions it owns (sharing pruning bounds with the peers from inside the scan kernel,
 * see lb200_peer_thresholds_*), writes its k' candidates straight into the ROOT shard's gather buffer, and the
 * root merges the G lists (== concat_batches + sort_to_indices(limit), ivf/v2.rs:383-400, then the scanner's
 * refine cut).  Ranks meet through epoch flags in each other's memory; a peer that never arrives makes the call
 * fail after a time-out (LB200_ECUDA) instead of hanging the GPU.  Results are bit-identical to lb200_search on
 * the unsharded index.
 *
 * (a) one process per GPU: every rank builds an lb200_index from ITS partitions (all nlist partitions, the ones
 *     it does not own empty; centroids and codebook replicated), then
 *       lb200_shard_export  -> a 128-byte opaque blob (CUDA IPC handles); capacities are per batch;
 *       the host exchanges the blobs of all ranks any way it likes (MPI, torch.distributed, a socket);
 *       lb200_shard_attach(world, rank, root, blobs [world][128])  (or same_process_ptrs [world][2] as returned
 *       in dev_ptrs_out_2, for shards living in one process);
 *       lb200_shard_search / lb200_shard_search_device on EVERY rank, same arguments, same order.  The result
 *       is written on the root rank only.  lb200_shard_search takes the full batch on every rank but moves only
 *       the rank's own slice over PCIe (the rest arrives from the peers over NVLink).
 * (b) one process driving all GPUs (a host that embeds the library): lb200_group_create shards a whole index by
 *     part_owner [nlist] (NULL = greedy bin-packing on row count, largest first) onto device_ids[G], root =
 *     shard 0; lb200_group_search == lb200_search. */
int32_t lb200_shard_export(lb200_index* ix, uint32_t max_queries, uint32_t max_nprobes, uint32_t max_kprime,
                           void* blob_out_128_or_null, uint64_t* dev_ptrs_out_2_or_null);
int32_t lb200_shard_attach(lb200_index* ix, uint32_t world, uint32_t rank, uint32_t root, const void* blobs_or_null,
                           const uint64_t* same_process_ptrs_or_null);
int32_t lb200_shard_detach(lb200_index* ix);
int32_t lb200_shard_search(lb200_index* ix, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                           uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                           float* out_dist, uint64_t* out_rowid, uint32_t* out_count);
int32_t lb200_shard_search_device(lb200_index* ix, const float* d_queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                                  uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                                  float* d_out_dist, uint64_t* d_out_rowid, uint32_t* d_out_count, void* stream);
int32_t lb200_shard_check(lb200_index* ix); /* error state of the exchange after asynchronous calls */

typedef struct lb200_group lb200_group;
int32_t lb200_group_create(const int32_t* device_ids, uint32_t G, int32_t metric, uint32_t dim, uint32_t nlist,
                           const float* centroids, uint32_t num_sub_vectors, uint32_t num_bits, const float* codebook,
                           const uint64_t* part_offsets, const uint8_t* codes, const uint64_t* row_ids,
                           const float* raw_vectors_or_null, const uint32_t* part_owner_or_null,
                           uint32_t max_queries, uint32_t max_nprobes, uint32_t max_kprime, lb200_group** out);
int32_t lb200_group_search(lb200_group* g, const float* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                           uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                           float* out_dist, uint64_t* out_rowid, uint32_t* out_count);
void lb200_group_destroy(lb200_group* g);

/* compute_distance (flat.rs:28-86; KNNVectorDistanceExec, io/exec/knn.rs:156-183): one
 * distance per (query,row); validity = Arrow validity bitmap (bit set = valid) or NULL;
 * null rows yield NaN.  out [Q][n]. */
int32_t lb200_flat_distance(int32_t device_id, int32_t metric, uint32_t dim, const float* queries,
                            uint32_t Q, const float* vectors, const uint8_t* validity_or_null,
                            uint64_t n, float* out);

/* compute_distance + SortExec(fetch=k) + NOT NULL filter (scanner.rs:2066-2136).
 * out_idx are row positions 0..n-1; out_count [Q] (<k when fewer valid rows). */
int32_t lb200_flat_topk(int32_t device_id, int32_t metric, uint32_t dim, const float* queries,
                        uint32_t Q, const float* vectors, const uint8_t* validity_or_null,
                        uint64_t n, uint32_t k, const float* lower_or_null,
                        const float* upper_or_null, float* out_dist, uint64_t* out_idx,
                        uint32_t* out_count);
/* Device-pointer variant (vectors resident in HBM), asynchronous on `stream`. */
int32_t lb200_flat_topk_device(int32_t device_id, int32_t metric, uint32_t dim,
                               const float* d_queries, uint32_t Q, const float* d_vectors,
                               const uint8_t* d_validity_or_null, uint64_t n, uint32_t k,
                               float* d_out_dist, uint64_t* d_out_idx, uint32_t* d_out_count,
                               void* stream);

/* A vector column kept RESIDENT on the device for flat kNN (compute_distance + SortExec(fetch = k), flat.rs:28-86,
 * scanner.rs:2066-2136; KNNVectorDistanceExec reads the column batch by batch, io/exec/knn.rs:156-183): the rows are
 * uploaded once instead of once per lb200_flat_topk call.  lb200_column_flat_topk[_device] == lb200_flat_topk with the
 * column's rows and validity, results bit-identical.  For L2, cosine and dot on columns of >= 65 536 rows with dim % 64 == 0, dim <= 256
 * and no validity bitmap / bounds the call runs a tensor-core CANDIDATE FILTER first (tcgen05.mma on bf16 copies,
 * operands by TMA, accumulators in TMEM; csrc/flat_tc.cuh): it keeps, per query, every row whose approximate distance is
 * within a proved error bound of an exact upper bound of the k-th distance, and only those rows are scored by the exact
 * fp32 kernels -- same results by construction (csrc/api_column.inl states the bound); if a candidate buffer overflows
 * the exact path answers (and the column's next call gets larger buffers).  Cosine keeps one more bf16 copy of the
 * rows (unit length), made on the first cosine call.  Calls on one column serialise.  The _device form enqueues on
 * `stream` (NULL: the column's own) and RETURNS AFTER THE STREAM HAS DRAINED: it reads the candidate counts back to decide
 * about the fall-back, so it is not an asynchronous call.  lb200_column_last_call reports what the last call did. */
typedef struct lb200_column lb200_column;
int32_t lb200_column_create(int32_t device_id, uint32_t dim, const float* vectors, const uint8_t* validity_or_null,
                            uint64_t n, lb200_column** out);
void lb200_column_destroy(lb200_column* c);
int32_t lb200_column_flat_topk(lb200_column* c, int32_t metric, const float* queries, uint32_t Q, uint32_t k,
                               const float* lower_or_null, const float* upper_or_null, float* out_dist,
                               uint64_t* out_idx, uint32_t* out_count);
int32_t lb200_column_flat_topk_device(lb200_column* c, int32_t metric, const float* d_queries, uint32_t Q, uint32_t k,
                                      const float* lower_or_null, const float* upper_or_null, float* d_out_dist,
                                      uint64_t* d_out_idx, uint32_t* d_out_count, void* stream);
int32_t lb200_column_last_call(const lb200_column* c, uint32_t* used_tc, uint64_t* candidates, uint32_t* fell_back);

/* GPU index BUILD: the steps the reference runs on the CPU (rust/lance-linalg/src/kmeans.rs,
 * lance-index/src/vector/pq.rs) or through torch (python/python/lance/torch/kmeans.py, lance/vector.py:134-700).
 * Given the initial centroids every step is deterministic and bit-identical to the reference's arithmetic
 * (tests/test_gpu_build.py: against the CPU restatement lo_kmeans_fit / lo_pq_quantize and against PQ codes the reference wrote).
 *  lb200_kmeans_train      KMeans::new_with_params / train_kmeans (kmeans.rs:440-561): membership = argmin over
 *                          l2 / dot distances (None when no distance is < +inf), centroid = f32 sum of the members
 *                          in data order / count, empty clusters re-seeded by split_clusters (:135-148), stop when
 *                          |loss - last| / last < tolerance.  metric L2 or DOT (cosine: normalise the rows first, as
 *                          ivf/builder.rs does, and pass L2).  init_centroids_or_null [k][dim]: the reference's
 *                          initialisation is an UNSEEDED random choice of k rows; NULL = a choice seeded by `seed`.
 *                          `redos` (default 1 in the reference) is not offered.
 *  lb200_compute_partitions compute_partitions (kmeans.rs:721-770): out_ids [n], 0xffffffff = None.
 *  lb200_residuals_device  x - centroid[part] (ivf/v2.rs:285-301 at build time).
 *  lb200_pq_train_device   one k-means (k = 256) per sub-space on the sub-vector slices (pq/builder.rs).
 *  lb200_pq_encode         ProductQuantizer::transform (pq.rs:120-190): out_codes [n][M] row-major, code =
 *                          argmin over the sub-space's 256 centroids (first minimum).
 * `_device` variants take device pointers for the large arrays (vectors, ids, codes); centroids / codebooks of the
 * training calls stay host arrays (they are small and the loop reads the loss on the host every round). */
int32_t lb200_kmeans_train(int32_t device_id, int32_t metric, uint32_t dim, const float* vectors, uint64_t n,
                           uint32_t k, uint32_t max_iters, float tolerance, uint64_t seed,
                           const float* init_centroids_or_null, float* out_centroids,
                           uint32_t* out_iters_or_null, double* out_loss_or_null);
int32_t lb200_kmeans_train_device(int32_t device_id, int32_t metric, uint32_t dim, const float* d_vectors,
                                  uint64_t n, uint32_t k, uint32_t max_iters, float tolerance, uint64_t seed,
                                  const float* init_centroids_or_null, float* out_centroids,
                                  uint32_t* out_iters_or_null, double* out_loss_or_null);
int32_t lb200_compute_partitions(int32_t device_id, int32_t metric, uint32_t dim, const float* centroids,
                                 uint32_t k, const float* vectors, uint64_t n, uint32_t* out_ids,
                                 float* out_dists_or_null);
int32_t lb200_compute_partitions_device(int32_t device_id, int32_t metric, uint32_t dim,
                                        const float* d_centroids, uint32_t k, const float* d_vectors,
                                        uint64_t n, uint32_t* d_out_ids, float* d_out_dists_or_null,
                                        void* stream);
int32_t lb200_residuals_device(int32_t device_id, uint32_t dim, const float* d_vectors,
                               const uint32_t* d_part_ids, const float* d_centroids, uint64_t n,
                               float* d_out, void* stream);
int32_t lb200_pq_train_device(int32_t device_id, int32_t metric, uint32_t dim, uint32_t num_sub_vectors,
                              const float* d_vectors, uint64_t n, uint32_t max_iters, float tolerance,
                              uint64_t seed, const float* init_codebook_or_null, float* out_codebook);
int32_t lb200_pq_encode(int32_t device_id, int32_t metric, uint32_t dim, uint32_t num_sub_vectors,
                        const float* codebook, const float* vectors, uint64_t n, uint8_t* out_codes);
int32_t lb200_pq_encode_device(int32_t device_id, int32_t metric, uint32_t dim, uint32_t num_sub_vectors,
                               const float* d_codebook, const float* d_vectors, uint64_t n,
                               uint8_t* d_out_codes, void* stream);

/* lb200_flat_topk_device WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, kernels and results (bit-identical) as lb200_flat_topk_device with
 * no bounds; the device scratch -- the [Q][n] f32 distance block plus the per-segment candidate lists -- lives in
 * a fixed `static __device__` array of lb200_noalloc_scratch_bytes() = 32 MiB that every GPU loading the library
 * holds until exit.  Limit: about Q * n * 4 bytes <= 32 MiB (e.g. 8 queries x 1M rows, 64 x 128k); beyond it the
 * call fails with LB200_EINVAL and a message stating need and capacity.  The call is asynchronous on `stream`,
 * never synchronises and calls no cudaMalloc / cudaFree / allocating routine; the first call on a device resolves
 * the array's device address (cudaGetSymbolAddress), so make one call outside a graph capture.  ALL calls share
 * the one array, which keeps the previous call's contents (the call clears what must start at zero): calls that
 * overlap on the GPU -- different streams or host threads -- RACE; the caller must order them. */
int32_t lb200_flat_topk_device_noalloc(int32_t device_id, int32_t metric, uint32_t dim,
                                       const float* d_queries, uint32_t Q, const float* d_vectors,
                                       const uint8_t* d_validity_or_null, uint64_t n, uint32_t k,
                                       float* d_out_dist, uint64_t* d_out_idx, uint32_t* d_out_count,
                                       void* stream);
uint64_t lb200_noalloc_scratch_bytes(void);

/* IVF_PQ over a Float16 column.  The reference runs its generic code with T = f16:
 * kmeans_find_partitions::<f16> (rust/lance-linalg/src/kmeans.rs:581-586), the residual as a Float16 array
 * (rust/lance/src/index/vector/ivf/v2.rs:285-301), build_distance_table_{l2,dot}::<f16> behind
 * PQDistCalculator::new (rust/lance-index/src/vector/pq/storage.rs:500-512, pq/distance.rs:17-85), normalize::<f16> for
 * cosine keys (rust/lance-linalg/src/kernels.rs:118-121), refine on the Float16 rows.  centroids [nlist][dim],
 * codebook [M][256][dim/M], raw vectors and keys are IEEE binary16 bit patterns; codes are 8-bit, layout as
 * lb200_index_create.  Distances are f32 in the order of the reference's scalar f16 fallbacks (csrc/f16pq.cuh).
 * Limits: num_sub_vectors % 16 == 0; the handle takes the *_f16 entry points only (the f32 search entry points answer
 * LB200_EINVAL); stats / destroy as for any handle.  No bf16 form: the reference's find_partitions and PQ storage
 * dispatch on Float16 / Float32 / Float64 only. */
int32_t lb200_index_create_f16(int32_t device_id, int32_t metric, uint32_t dim, uint32_t nlist, const uint16_t* centroids,
                               uint32_t num_sub_vectors, const uint16_t* codebook, const uint64_t* part_offsets,
                               const uint8_t* codes, const uint64_t* row_ids, const uint16_t* raw_vectors_or_null,
                               lb200_index** out);
int32_t lb200_search_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, uint32_t nprobes, uint32_t k,
                         uint32_t refine_factor, const float* lower_or_null, const float* upper_or_null,
                         float* out_dist, uint64_t* out_rowid, uint32_t* out_count);
/* The two calls the exec nodes make (find_partitions once per query, search_in_partition per probed partition), Float16
 * keys used as given (already normalised for a cosine index), arguments as lb200_find_partitions / lb200_search_partitions
 * (no prefilter form yet). */
int32_t lb200_find_partitions_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, uint32_t nprobes,
                                  uint32_t* out_part_ids, float* out_part_dists_or_null);
int32_t lb200_search_partitions_f16(lb200_index* ix, const uint16_t* queries, uint32_t Q, const uint32_t* part_ids, uint32_t P,
                                    uint32_t k_prime, const float* lower_or_null, const float* upper_or_null, float* out_dist,
                                    uint64_t* out_rowid, uint32_t* out_count);

/* The same two entry points for Float16 columns (the reference's f16 path:
 * `impl L2/Dot/Cosine/Normalize for f16`, rust/lance-linalg/src/distance/{l2.rs:136-166, dot.rs:100-137,
 * cosine.rs:88-121, norm_l2.rs:46-76}, C kernels rust/lance-linalg/src/simd/f16.c).  queries / vectors are
 * IEEE binary16 bit patterns (the key is cast to the column type by the scanner); distances are f32, computed in the
 * order of the reference's deterministic scalar fallbacks (see csrc/api_f16.inl). */
int32_t lb200_flat_distance_f16(int32_t device_id, int32_t metric, uint32_t dim, const uint16_t* queries,
                                uint32_t Q, const uint16_t* vectors, const uint8_t* validity_or_null,
                                uint64_t n, float* out);
int32_t lb200_flat_topk_f16(int32_t device_id, int32_t metric, uint32_t dim, const uint16_t* queries,
                            uint32_t Q, const uint16_t* vectors, const uint8_t* validity_or_null,
                            uint64_t n, uint32_t k, const float* lower_or_null,
                            const float* upper_or_null, float* out_dist, uint64_t* out_idx,
                            uint32_t* out_count);

/* Multi-GPU merge step of SURVEY 8(e): G gathered per-shard candidate lists -> global result.
 * d_dist/d_rowid/d_exact [G][Q][stride], d_count [G][Q]; device pointers.
 * Selects the kprime smallest by (dist,row id) (== the merge of ivf/v2.rs:395-400); if
 * d_exact_or_null != NULL then re-ranks those by (exact,row id) with [lower,upper) applied and
 * keeps k (== the refine of scanner.rs:1706-1721), writing exact distances to d_out_dist;
 * otherwise keeps min(k,kprime).  Output stride k. */
int32_t lb200_merge_topk_device(int32_t device_id, const float* d_dist, const uint64_t* d_rowid,
                                const float* d_exact_or_null, const uint32_t* d_count, uint32_t G,
                                uint32_t Q, uint32_t stride, uint32_t kprime, uint32_t k,
                                const float* lower_or_null, const float* upper_or_null,
                                float* d_out_dist, uint64_t* d_out_rowid, uint32_t* d_out_count,
                                void* stream);

/* Host-pointer form of lb200_merge_topk_device: the union of G result lists of one query batch.
 * == UnionExec over the delta indices of one column + SortExec(_distance).with_fetch(k') of
 * Scanner::ann (rust/lance/src/dataset/scanner.rs:2139-2157; fan-out over deltas in
 * new_knn_exec / ANNIvfSubIndexExec, rust/lance/src/io/exec/knn.rs:405-635).
 * dist/rowid/exact [G][Q][stride], count [G][Q], outputs [Q][k] / [Q]; all host pointers. */
int32_t lb200_merge_topk(int32_t device_id, const float* dist, const uint64_t* rowid,
                         const float* exact_or_null, const uint32_t* count, uint32_t G, uint32_t Q,
                         uint32_t stride, uint32_t kprime, uint32_t k, const float* lower_or_null,
                         const float* upper_or_null, float* out_dist, uint64_t* out_rowid,
                         uint32_t* out_count);

/* flat_knn over a per-query candidate set whose vectors the CALLER fetched (TakeExec,
 * rust/lance/src/io/exec/take.rs:178): compute_distance with the original key and the index's
 * metric (flat.rs:28-86), [lower,upper) filter, SortExec(fetch=k), NOT NULL
 * (scanner.rs:2066-2136).  This is the refine step on the drop-in path (scanner.rs:1706-1721,
 * raw vectors not resident on the device) and the final step of knn_combined
 * (scanner.rs:1755-1835: ANN results of the indexed rows UNION flat kNN of the unindexed
 * fragments, all re-scored exactly).
 * cand_vectors [Q][C][dim], cand_rowids [Q][C], cand_count [Q] (<= C, C <= 4096); outputs
 * [Q][k] ascending by (distance,row id), out_count [Q].  All host pointers. */
int32_t lb200_rerank(int32_t device_id, int32_t metric, uint32_t dim, const float* queries,
                     uint32_t Q, const float* cand_vectors, const uint64_t* cand_rowids,
                     const uint32_t* cand_count, uint32_t C, uint32_t k,
                     const float* lower_or_null, const float* upper_or_null, float* out_dist,
                     uint64_t* out_rowid, uint32_t* out_count);

/* normalize_arrow (rust/lance-linalg/src/kernels.rs:119-142) of a batch of keys: the step the ANN exec
 * nodes apply to the key of a cosine index before find_partitions / search_in_partition
 * (rust/lance/src/io/exec/knn.rs:366-369,616-619), for callers that drive those two entry points
 * themselves (e.g. the prefilter path).  lb200_search does it internally. */
int32_t lb200_normalize(int32_t device_id, uint32_t dim, const float* queries, uint32_t Q, float* out);

/* Instrumentation (not part of the reference surface): per-handle counters of the last
 * search / search_partitions call, used by bench.py for the roofline line. */
typedef struct {
  double scan_ms;          /* CUDA-event time of the PQ scan kernel (0 if profiling off)  */
  double coarse_ms;        /* CUDA-event time of the coarse distance + select kernels      */
  double total_ms;         /* CUDA-event time of the whole device pipeline                 */
  uint64_t scan_code_bytes;/* algorithmic code bytes: sum over (q,p) pairs of n_p * M      */
  uint64_t pairs;          /* (query, partition) pairs scanned                             */
  uint32_t kernel_launches;/* kernels launched by the call                                 */
  uint32_t scan_variant;   /* 1 = generic, 2 = bank-skewed tile16, 3 = skew-stream lane=row  */
  uint32_t scan_ql;        /* queries per CTA of the scan kernel                           */
  uint32_t scan_items;     /* work items of the scan kernel                                */
} lb200_stats;
int32_t lb200_set_profiling(lb200_index* ix, int32_t enabled);
int32_t lb200_get_stats(const lb200_index* ix, lb200_stats* out);
/* Force a scan kernel variant (0 = auto, 1 = generic, 2 = bank-skewed tile16, 3 = skew
 * streams); 1/2 need the tile16 layout (env LB200_LAYOUTS=all at create when M % 16 == 0). */
int32_t lb200_set_scan_variant(lb200_index* ix, int32_t variant);

const char* lb200_last_error(void);
const char* lb200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* LANCE_B200_H */
