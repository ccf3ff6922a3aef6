/*
 * lance_b200.h — C ABI of liblance_b200.so, the B200 (sm_100a) drop-in for the kNN search path of
 * decisiongraph/duckdb-lancedb.
 *
 * Part 1 re-exports, symbol for symbol, the `lance_*` C functions the extension's C++ side binds in
 * /root/reference/src/rust_ffi.cpp:9-41 (implemented there by the Rust crate, rust_lib/src/ffi.rs).
 * Linking the extension against this library instead of liblancedb_rust.a needs no source change for
 * the search path: same names, same argument meaning, same error convention.
 *
 * Part 2 adds the siblings BASELINE.json's north_star asks for (query batches, a prefilter row
 * bitmap instead of the newer revision's predicate string, staging of an externally trained
 * IVF_PQ index) and a device-resident entry used by bench.py and by the multi-GPU merge.
 *
 * Conventions (rust_lib/src/ffi.rs:15-24, src/rust_ffi.cpp:44-206):
 *   - no exception ever crosses this boundary;
 *   - failure = sentinel return (NULL / -1 / non-zero) and a NUL-terminated message truncated to
 *     err_buf_len-1 bytes in err_buf (err_buf may be NULL);
 *   - all in/out buffers are caller-allocated and only borrowed for the call;
 *   - every entry point is thread-safe on one handle (searches run concurrently; add / delete /
 *     create_index are exclusive).
 *   - there is NO CPU fallback: without a CUDA device every compute entry fails with a message.
 */
#ifndef LANCE_B200_H
#define LANCE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ===================================================================== Part 1: reference symbols */

/* rust_ffi.cpp:9-10 / ffi.rs:37-57.  metric: "l2" (default), "cosine", "dot"|"ip"
 * (lance_manager.rs:493-497).  db_path/table_name are recorded; the device store is the index.
 * Environment LANCE_B200_DEVICES="0,1,.." | "all" (more than one device): this call, lance_create_detached_from_arrow and
 * lance_open_detached (for a table saved that way) return a SHARDED handle (lance_b200_create_sharded_at below) — the
 * unmodified extension then uses every listed GPU. */
void *lance_create_detached(const char *db_path, int32_t dimension, const char *metric, const char *table_name,
                            char *err_buf, int err_buf_len);
/* rust_ffi.cpp:11-12 / ffi.rs:62-87.  Multi-column Arrow ingest is outside the kNN path: returns NULL. */
void *lance_create_detached_from_arrow(const char *db_path, void *arrow_schema, const char *metric,
                                       const char *table_name, char *err_buf, int err_buf_len);
/* rust_ffi.cpp:13-14 / ffi.rs:93-111.  Re-opens a store saved by lance_b200_save(); NULL if none. */
void *lance_open_detached(const char *db_path, const char *table_name, const char *metric, char *err_buf,
                          int err_buf_len);
/* rust_ffi.cpp:15 / ffi.rs:137-142 */
void lance_free_detached(void *handle);
/* rust_ffi.cpp:16 / ffi.rs:113-123: always 0 here (vector-only tables) */
int32_t lance_detached_has_extra_columns(void *handle);
/* rust_ffi.cpp:17 / ffi.rs:125-135 */
int32_t lance_detached_dimension(void *handle);
/* rust_ffi.cpp:18 / ffi.rs:228-250: returns the new label or -1 */
int64_t lance_detached_add(void *handle, const float *vector, int32_t dimension, char *err_buf, int err_buf_len);
/* rust_ffi.cpp:19-20 / ffi.rs:253-289: labels are consecutive from an atomic counter
 * (lance_manager.rs:214,232); returns count or -1 */
int32_t lance_detached_add_batch(void *handle, const float *vectors, int32_t num, int32_t dim, int64_t *out_labels,
                                 char *err_buf, int err_buf_len);
/* rust_ffi.cpp:21-22: not on the kNN path, returns -1 */
int32_t lance_detached_add_batch_arrow(void *handle, void *arrow_schema, void *arrow_array, int64_t *out_labels,
                                       char *err_buf, int err_buf_len);
/* rust_ffi.cpp:23-25 / ffi.rs:187-228 / lance_manager.rs:307-393: the listed labels of the source that exist and are
 * not deleted are appended to the target (device-to-device row gather, source storage order) under fresh
 * sequential labels; out_old/out_new (capacity live_count) receive the mapping.  Returns the merged count, -1 on error
 * (null handle, dimension mismatch, source == target). */
int32_t lance_detached_merge(void *target_handle, void *source_handle, const int64_t *live_source_labels,
                             int32_t live_count, int64_t *out_old_labels, int64_t *out_new_labels, char *err_buf,
                             int err_buf_len);
/* rust_ffi.cpp:26-28 / ffi.rs:296-329 — THE hot call.  Writes up to k (label, distance) pairs in
 * ascending distance order (ties: ascending label); returns the count, or -1.
 * dim != index dimension is an error here, as in lance_manager.rs:401-407 (the C++ caller guards it,
 * src/lance_index.cpp:444-446).  nprobes/refine_factor only matter once an IVF_PQ index exists. */
int32_t lance_detached_search(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                              int32_t refine_factor, int64_t *out_labels, float *out_distances, char *err_buf,
                              int err_buf_len);
/* rust_ffi.cpp:29 / ffi.rs:336-353: live rows */
int64_t lance_detached_count(void *handle, char *err_buf, int err_buf_len);
/* rust_ffi.cpp:30 / ffi.rs:356-375 */
int32_t lance_detached_delete(void *handle, int64_t label, char *err_buf, int err_buf_len);
/* rust_ffi.cpp:31-32 / ffi.rs:377-397 */
int32_t lance_detached_delete_batch(void *handle, const int64_t *labels, int32_t count, char *err_buf,
                                    int err_buf_len);
/* rust_ffi.cpp:33-34 / ffi.rs:404-423: trains IVF_PQ on the GPU over the live rows.  0 = "use the
 * default" (lance_manager.rs:483-515): partitions = round(sqrt(n)), sub-vectors = dim/16 (dim/8, 1). */
int32_t lance_detached_create_index(void *handle, int32_t num_partitions, int32_t num_sub_vectors, char *err_buf,
                                    int err_buf_len);
/* rust_ffi.cpp:35-36 / ffi.rs:425-446 -> lance_manager.rs:521-551 (IVF_HNSW_SQ, replace = true).  Accepted, no graph is
 * built: an existing IVF_PQ index is dropped (replaced) and searches are exact scans of the HBM-resident table from then
 * on (recall 1.0).  m / ef_construction <= 0 = defaults, as in the reference. */
int32_t lance_detached_create_hnsw_index(void *handle, int32_t m, int32_t ef_construction, char *err_buf,
                                         int err_buf_len);
/* rust_ffi.cpp:37 / ffi.rs:448-465 -> lance_manager.rs:553-566: tombstoned rows are dropped for good (HBM released), labels kept */
int32_t lance_detached_compact(void *handle, char *err_buf, int err_buf_len);
/* rust_ffi.cpp:38-39 / ffi.rs:472-503: returns dim, or -1 (unknown/deleted label, small buffer) */
int32_t lance_detached_get_vector(void *handle, int64_t label, float *out_vec, int32_t capacity, char *err_buf,
                                  int err_buf_len);
/* rust_ffi.cpp:40-41 / ffi.rs:511-541: NULL out buffers = count only */
int32_t lance_detached_get_all_vectors(void *handle, int64_t *out_labels, float *out_vectors, int64_t *out_count,
                                       char *err_buf, int err_buf_len);

/* ===================================================================== Part 2: new siblings */

/* Replaces the newer revision's `const char *predicate` argument (src/lance_index.cpp:452-453, F3
 * in SURVEY.md): the host evaluates the pushed-down WHERE over its own columns into a bitmap, bit
 * (label & 31) of word (label >> 5) set = the row may be returned.  Prefilter semantics
 * (test/sql/lance_optimizer_filter.test:23-99).  allow_bitmap NULL = no filter.  bitmap_bits =
 * number of labels the bitmap covers (labels beyond it are filtered out). */
int32_t lance_detached_search_filtered(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                       int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                                       int64_t *out_labels, float *out_distances, char *err_buf, int err_buf_len);

/* Query batch: queries[nq][dim]; out_labels/out_distances are [nq][k] (unused tail entries:
 * label -1, distance +inf), out_counts[nq].  Returns nq or -1. */
int32_t lance_detached_search_batch(void *handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                    int32_t nprobes, int32_t refine_factor, const uint32_t *allow_bitmap,
                                    int64_t bitmap_bits, int64_t *out_labels, float *out_distances,
                                    int32_t *out_counts, char *err_buf, int err_buf_len);

/* Same search, everything already in HBM of the handle's device: d_queries[nq][dim] fp32,
 * d_allow (nullable), d_out_labels/d_out_distances [nq][k], d_out_counts[nq].  Enqueued on `stream`
 * (a cudaStream_t; NULL = the legacy default stream); does not synchronise.  Returns 0 or -1. */
int32_t lance_b200_search_device(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                 int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                 int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts, void *stream,
                                 char *err_buf, int err_buf_len);

/* Same final select over a gathered buffer of per-rank packed records: shard g's lists start at
 * d_labels + g*labels_stride, d_dists + g*dists_stride, d_counts + g*counts_stride (strides in elements), so one
 * all_gather of [labels | dists | counts] per rank feeds it directly.  One kernel launch on `stream`. */
int32_t lance_b200_merge_topk_strided_device(int32_t device, const int64_t *d_labels, const float *d_dists,
                                             const int32_t *d_counts, int64_t labels_stride, int64_t dists_stride,
                                             int64_t counts_stride, int32_t G, int32_t nq, int32_t k,
                                             int64_t *d_out_labels, float *d_out_dists, int32_t *d_out_counts,
                                             void *stream, char *err_buf, int err_buf_len);

/* lance_b200_search_device without its one host synchronisation.  The batched flat path ends with a 4-byte
 * read-back ("did any candidate list overflow?", degenerate data only) that lance_b200_search_device waits for;
 * the _async form leaves it pending so that the caller can enqueue the multi-GPU gather and merge behind the
 * search without a gap.  lance_b200_search_finish waits for `stream`, settles the check (re-running overflowed
 * queries with the exact scan into the same output buffers, which must still be valid) and returns the number
 * of re-run queries (0 = the results that were already consumed are final), or -1.  d_overflow (nullable, device)
 * receives, in stream order, the number of queries whose result is not final yet, so a multi-GPU step can carry
 * it through its gather and learn on every rank whether any shard has to be settled.  One pending search per
 * (handle, stream); the next search on the same stream settles that stream's pending one first, other streams'
 * pending searches are not touched. */
int32_t lance_b200_search_device_async(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                       int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                       int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts,
                                       uint32_t *d_overflow, void *stream, char *err_buf, int err_buf_len);
int32_t lance_b200_search_finish(void *handle, void *stream, char *err_buf, int err_buf_len);

/* Final select of the multi-GPU merge: G gathered per-shard lists
 * d_labels/d_dists [G][nq][k], d_counts [G][nq] -> one [nq][k] list.  Device buffers, on `stream`. */
int32_t lance_b200_merge_topk_device(int32_t device, const int64_t *d_labels, const float *d_dists,
                                     const int32_t *d_counts, int32_t G, int32_t nq, int32_t k,
                                     int64_t *d_out_labels, float *d_out_dists, int32_t *d_out_counts, void *stream,
                                     char *err_buf, int err_buf_len);

/* Shard support: labels handed out by this handle start at `base` (call before the first add). */
int32_t lance_b200_set_label_base(void *handle, int64_t base, char *err_buf, int err_buf_len);
/* Pre-size the device store for `rows` rows (avoids growth copies on large ingests). */
int32_t lance_b200_reserve(void *handle, int64_t rows, char *err_buf, int err_buf_len);
/* add_batch with the vectors already on the device (d_vectors[num][dim] fp32); returns first label or -1. */
int64_t lance_b200_add_batch_device(void *handle, const float *d_vectors, int64_t num, int32_t dim, void *stream,
                                    char *err_buf, int err_buf_len);

/* IVF_PQ index exchange, in the layout of oracle/oracle.h (oracle_ivfpq_t): centroids[P][dim],
 * codebook[M][256][dim/M], part_offsets[P+1], row_ids[n_indexed] (labels), codes[n_indexed][M],
 * partition-major.  set: stages a host-trained index into HBM; get: exports the device-trained one
 * (pass NULL buffers to read the sizes first).  Both return 0 or -1. */
typedef struct {
	int32_t metric; /* 0 l2, 1 cosine, 2 dot */
	int32_t dim;
	int32_t P;
	int32_t M;
	int64_t n_indexed;
} lance_b200_ivfpq_info;
int32_t lance_b200_set_ivfpq(void *handle, int32_t P, int32_t M, int64_t n_indexed, const float *centroids,
                             const float *codebook, const int64_t *part_offsets, const int64_t *row_ids,
                             const uint8_t *codes, char *err_buf, int err_buf_len);
int32_t lance_b200_get_ivfpq(void *handle, lance_b200_ivfpq_info *info, float *centroids, float *codebook,
                             int64_t *part_offsets, int64_t *row_ids, uint8_t *codes, char *err_buf,
                             int err_buf_len);
/* Drops the IVF_PQ index (searches fall back to the exact flat scan, still on the GPU). */
int32_t lance_b200_drop_index(void *handle, char *err_buf, int err_buf_len);

/* One IVF_PQ query with the intermediate results BASELINE.json's parity criterion names:
 * out_parts[nprobes] probed partitions (ascending centroid distance), out_cand_labels /
 * out_cand_adc [k*max(rf,1)] the pre-refine ADC candidates, then the final list. */
int32_t lance_b200_ivfpq_search_debug(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                      int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                                      int32_t *out_parts, int64_t *out_cand_labels, float *out_cand_adc,
                                      int32_t *out_ncand, int64_t *out_labels, float *out_distances, char *err_buf,
                                      int err_buf_len);

/* Persistence of the device store in the library's own format under <db_path>/<table>.b200
 * (SURVEY.md 8f-1); lance_open_detached() reads it back. */
int32_t lance_b200_save(void *handle, char *err_buf, int err_buf_len);

/* Counters since handle creation (bench.py's gpu_launches, tests' "the CUDA path ran"). */
typedef struct {
	int64_t kernel_launches;   /* kernels of this library launched for this handle */
	int64_t searches;          /* search calls (a batch counts once) */
	int64_t flat_simt_batches; /* batches served by the fp32 streaming scan */
	int64_t flat_tc_batches;   /* batches served by the tcgen05 candidate GEMM + fp32 re-rank */
	int64_t tc_fallback_queries; /* queries the re-rank bound could not certify (re-run exactly) */
	int64_t ivfpq_batches;
	int64_t rows;              /* rows in the store (live + tombstoned) */
	int64_t live_rows;
	int32_t device;
	int32_t has_index;
} lance_b200_stats_t;
int32_t lance_b200_stats(void *handle, lance_b200_stats_t *out);

/* CUDA-event timing of the dominant kernels, recorded on the launching stream while enabled.
 * read() synchronises the device, returns the sums since the last read and clears them. */
typedef struct {
	double flat_scan_ms;   /* fp32 streaming scan kernel (flat_scan.cu) */
	int64_t flat_scan_launches;
	double flat_gemm_ms;   /* tcgen05 candidate GEMM kernel (flat_gemm.cu) */
	int64_t flat_gemm_launches;
	double adc_scan_ms;    /* PQ ADC scan kernel (ivfpq.cu) */
	int64_t adc_scan_launches;
	double rerank_ms;      /* exact fp32 re-rank of candidates */
	int64_t rerank_launches;
} lance_b200_profile_t;
int32_t lance_b200_profile_enable(void *handle, int32_t on);
int32_t lance_b200_profile_read(void *handle, lance_b200_profile_t *out);

/* Measurement hook (bench.py): `iters` calls of lance_b200_search_device, each bracketed by its own pair of CUDA events
 * on `stream` (out_ms[i] = that call's device time in milliseconds, recorded and read in C so that a 30 us step is not
 * measured through an interpreter).  d_flush (nullable, flush_bytes >= 2 x the L2 size): before every call, outside
 * the timed interval, half of it is written and the other half read, so the table is not served from L2 and no dirty
 * line is left to write back.  Returns 0 or -1. */
int32_t lance_b200_time_search_device(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                      int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                      int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts, void *stream,
                                      int32_t iters, void *d_flush, int64_t flush_bytes, float *out_ms, char *err_buf,
                                      int err_buf_len);

/* ---------------------------------------------------------------- extra columns + predicate pushdown
 * The newer revision of the extension passes the pushed-down WHERE as a Lance SQL string
 * (src/lance_index.cpp:452-453 calls LanceDetachedSearch(.., predicate, ..); generator src/lance_optimizer.cpp:204-344)
 * and ships the index's extra columns through the Arrow C data interface (src/lance_index.cpp:321-359 ->
 * rust_lib/src/ffi.rs:63-90, 148-182; both are implemented above).  Here the columns live in HBM and the string is
 * compiled and evaluated on the device (SQL three-valued logic, prefilter semantics of
 * test/sql/lance_optimizer_filter.test:23-99); the bitmap is cached per (predicate, column epoch).
 * predicate NULL or "" = no filter.  Same returns as lance_detached_search / _search_batch. */
int32_t lance_detached_search_predicate(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                        int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                        float *out_distances, char *err_buf, int err_buf_len);
int32_t lance_detached_search_batch_predicate(void *handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                              int32_t nprobes, int32_t refine_factor, const char *predicate,
                                              int64_t *out_labels, float *out_distances, int32_t *out_counts,
                                              char *err_buf, int err_buf_len);
/* src/lance_index.cpp:481-486 (LanceDetachedCreateScalarIndex, newer revision): nothing to build on the device —
 * validates that the column exists, returns 0 or -1. */
int32_t lance_detached_create_scalar_index(void *handle, const char *column, const char *index_type, char *err_buf,
                                           int err_buf_len);
/* Typed host-buffer ingest of one extra column for the rows of labels [first_label, first_label + count) (they must
 * exist).  type: 0 int64 values, 1 double values, 2 boolean (uint8 values), 3 UTF-8 (values = bytes,
 * offsets[count + 1]).  valid: one byte per row or NULL (all valid).  Rows never written are NULL.  Returns 0 or -1. */
int32_t lance_b200_set_column(void *handle, const char *name, int32_t type, int64_t first_label, int64_t count,
                              const void *values, const int64_t *offsets, const uint8_t *valid, char *err_buf,
                              int err_buf_len);
/* Evaluates (or fetches from the cache) the allow bitmap of a predicate.  out_bitmap (nullable): host copy of the
 * first out_words words; out_device_bitmap (nullable): the device pointer, valid while the predicate stays cached
 * (for lance_b200_search_device / _time_search_device); out_was_cached (nullable).  Returns the number of rows the
 * bitmap covers, or -1. */
int64_t lance_b200_predicate_bitmap(void *handle, const char *predicate, uint32_t *out_bitmap, int64_t out_words,
                                    const uint32_t **out_device_bitmap, int32_t *out_was_cached, char *err_buf,
                                    int err_buf_len);

/* Appends `num` rows that carry the caller's labels first_label .. first_label + num - 1 (first_label must not be
 * below any label handed out so far).  A store whose labels stop being consecutive keeps them explicitly (8 B per row
 * on the device and on the host); results, deletes, bitmaps and columns keep speaking labels.  This is how one global
 * label sequence (rust_lib/src/lance_manager.rs:214, 232: an atomic counter) is spread over several device stores by
 * lance_b200_create_sharded.  Returns 0 or -1. */
int32_t lance_b200_add_batch_labeled(void *handle, const float *vectors, int64_t num, int32_t dim, int64_t first_label,
                                     char *err_buf, int err_buf_len);

/* Device rows d_vectors[num][dim] with one label per row (labels[num] in HOST memory, strictly ascending, none below
 * a label already handed out): how a partition-sharded IVF_PQ shard is loaded (duckdb-lancedb_b200/sharding.py).  On a
 * store with explicit labels lance_b200_set_ivfpq / _get_ivfpq exchange ROW numbers, not labels.  Returns 0 or -1. */
int32_t lance_b200_add_batch_device_labels(void *handle, const float *d_vectors, int64_t num, int32_t dim,
                                           const int64_t *labels, void *stream, char *err_buf, int err_buf_len);

/* One handle over several GPUs of the box, in ONE process (the consumer is an embedded database — no
 * launcher, no ranks).  devices[n_devices]: CUDA ordinals (a device may be named twice: two shards on it).  The
 * returned handle is used with the SAME calls as a plain one — lance_detached_add / _add_batch / _add_batch_arrow,
 * _search / _search_filtered / _search_batch / _search_predicate / _search_batch_predicate, _count, _delete /
 * _delete_batch, _get_vector, _compact, _dimension, _has_extra_columns, lance_b200_set_column, lance_free_detached:
 * whole appended batches go to the least-filled shard under one global label counter, a search runs on every device
 * concurrently and the per-shard lists are merged under the library's (distance, label) order.
 * lance_detached_create_index trains ONE IVF_PQ INDEX PER SHARD over that shard's rows with the caller's parameters (a
 * search then probes nprobes lists on every shard); lance_detached_create_hnsw_index / lance_b200_drop_index drop them;
 * lance_b200_stats adds the shards' counters up; lance_b200_save works on a handle made by lance_b200_create_sharded_at.
 * The index exchange calls and the device-pointer entry points answer "not supported on a sharded handle".  NULL + message on failure. */
void *lance_b200_create_sharded(const int32_t *devices, int32_t n_devices, int32_t dimension, const char *metric,
                                char *err_buf, int err_buf_len);
/* The same handle with a place on disk: shard i persists as table "<table_name>.shard<i>" under db_path (one store.bin
 * each: rows, tombstones, labels, extra columns, its IVF_PQ index) and lance_b200_save adds "<table_name>.b200shards"
 * (shard count, label counter, batch -> shard map).  lance_b200_open_sharded reopens it; n_devices must be the saved
 * shard count (shard i goes to devices[i]).  These stand where LanceIndex::LoadFromStorage calls lance_open_detached
 * (src/lance_index.cpp:584) when the index was created over several GPUs. */
void *lance_b200_create_sharded_at(const char *db_path, const char *table_name, const int32_t *devices, int32_t n_devices,
                                   int32_t dimension, const char *metric, char *err_buf, int err_buf_len);
void *lance_b200_open_sharded(const char *db_path, const char *table_name, const char *metric, const int32_t *devices,
                              int32_t n_devices, char *err_buf, int err_buf_len);
/* Rows held by each shard (out_rows[capacity]); returns the number of shards, -1 if `handle` is not sharded. */
int32_t lance_b200_sharded_rows(void *handle, int64_t *out_rows, int32_t capacity);

/* ---------------------------------------------------------------- allocation-free flat search
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Exact flat top-k of ONE
 * query over a caller-owned device table d_vectors[rows][dim] (fp32 row-major); metric 0 l2 (squared), 1 cosine, 2 dot
 * (reported as 1 - q.x), the conventions of lance_detached_search.  d_allow: nullable prefilter bitmap over rows.
 * Writes d_out_labels[k] (label_base + row, -1 beyond the count), d_out_distances[k] (+inf beyond it) and
 * d_out_count[0]; enqueued on `stream`, no host synchronisation.  Performs NO allocation: its partial top-k lists live in
 * a file-scope device array of 2 424 832 bytes (296 blocks x 8 warps x 128 keys x 8 B) that every GPU loading the
 * library holds until exit.  Limits: 1 <= k <= 128, rows < 2^32 — larger values are rejected with a message.  All calls
 * on one device share that array: calls that overlap on the GPU (different streams) race — serialise them; an
 * overlap that the call can detect yields count -1.  Returns 0 or -1. */
int32_t lance_b200_flat_search_noalloc(int32_t device, const float *d_vectors, int64_t rows, int32_t dim, int32_t metric,
                                       const float *d_query, int32_t k, int64_t label_base, const uint32_t *d_allow,
                                       int64_t bitmap_bits, int64_t *d_out_labels, float *d_out_distances,
                                       int32_t *d_out_count, void *stream, char *err_buf, int err_buf_len);

/* Tuning / test hooks: force a flat path.  0 auto, 1 always the fp32 scan, 2 always tcgen05. */
int32_t lance_b200_set_flat_path(void *handle, int32_t mode);

/* Library-wide.  Device of handles created afterwards; default: $LANCE_B200_DEVICE, else $LOCAL_RANK
 * (one process per GPU under torchrun), else 0. */
int32_t lance_b200_set_default_device(int32_t device);
int32_t lance_b200_device_count(void);
const char *lance_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* LANCE_B200_H */
