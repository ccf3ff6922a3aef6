/*
 * oracle.h — CPU restatement of the kNN hot path of decisiongraph/duckdb-lancedb.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under duckdb-lancedb_b200/ may include, link or
 * call this.  Allowed users: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline
 * leg and `bench.py --impl reference`.
 *
 * PARITY STATUS
 *   flat search, squared-L2 : pinned — reproduces every id/distance the reference's SQL
 *                             tests hold (test/sql/lance_basic.test:28-42,
 *                             lance_optimizer.test:33-72, lance_optimizer_filter.test:23-99,
 *                             lance_delete.test:22-55) and DuckDB's array_* known answers
 *                             (duckdb/test/sql/function/array/array_distance.test:17-24,
 *                             array_cosine_distance.test:24-31, array_inner_product.test:34-41);
 *                             the distance arithmetic is additionally checked against the
 *                             reference's own DuckDB kernels compiled into oracle/_ref.
 *   cosine / dot flat, IVF_PQ: PARITY UNPINNED — the arithmetic lives in the un-vendored
 *                             crates lancedb 0.15.0 / lance-index 0.22.0 / lance-linalg 0.22.0
 *                             (rust_lib/Cargo.lock:2733-3128); no Rust toolchain, no golden
 *                             vectors for these in the reference's tests.  The published
 *                             algorithm is restated here and anchored on the call sites
 *                             rust_lib/src/lance_manager.rs:394-451 (search) and :483-515
 *                             (IVF_PQ build parameters / metric mapping).
 */
#ifndef LANCE_ORACLE_H
#define LANCE_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* metric ids follow rust_lib/src/lance_manager.rs:493-497 ("cosine", "dot"|"ip", else L2) */
enum { ORACLE_L2 = 0, ORACLE_COSINE = 1, ORACLE_DOT = 2 };

int oracle_metric_from_string(const char *metric);

/* Distance of one pair.
 *   L2     = sum (q-x)^2            (squared; pinned by test/sql/lance_basic.test:40-42)
 *   cosine = 1 - q.x/(|q||x|)       (clamped like array_kernels.hpp:62-72; 1.0 if a norm is 0)
 *   dot    = 1 - q.x                (lance-linalg dot_distance convention, no source on box)
 * _f32 accumulates sequentially in fp32 exactly like DuckDB's fold
 * (duckdb/extension/core_functions/include/core_functions/array_kernels.hpp:11-105);
 * _f64 accumulates in double and is the ranking truth. */
float oracle_distance_f32(int metric, const float *q, const float *x, int d);
double oracle_distance_f64(int metric, const float *q, const float *x, int d);

/* Exact flat top-k of one query (lance_manager.rs:394-451 with no ANN index).
 * Ranking: ascending fp64 distance, ties broken by ascending label (= row index).
 * allow  : optional prefilter bitmap, bit (i&31) of word i>>5 set = row i may be returned
 *          (prefilter semantics pinned by lance_optimizer_filter.test:23-99).
 * deleted: optional tombstone bitmap, bit set = row was deleted (lance_manager.rs:461-471).
 * Returns number of results (<= k; fewer when fewer rows survive, lance_delete.test:22-34). */
int oracle_flat_search(const float *X, int64_t n, int d, int metric, const float *q, int k,
                       const uint32_t *allow, const uint32_t *deleted, int64_t label_base,
                       int64_t *out_labels, float *out_dist32, double *out_dist64);

/* Multi-threaded fp32 batch version used as the timed CPU baseline (OpenMP over row ranges,
 * X read once per batch, per-thread heaps).  Same ranking rule on fp32 distances. */
int oracle_flat_search_batch_f32(const float *X, int64_t n, int d, int metric, const float *Q,
                                 int nq, int k, const uint32_t *allow, const uint32_t *deleted,
                                 int64_t *out_labels, float *out_dist, int32_t *out_counts,
                                 int nthreads);
int oracle_num_threads(void);

/* Merge G per-shard result lists (each sorted, counts[g] valid entries of k) into one. */
int oracle_merge_topk(const int64_t *labels, const float *dists, const int32_t *counts, int G,
                      int k, int64_t *out_labels, float *out_dists);

/* ---------------- IVF_PQ (8-bit codes, 256 centroids per sub-quantizer) ---------------- */
typedef struct {
	int32_t metric;  /* metric the index was built with */
	int32_t dim;     /* D */
	int32_t P;       /* partitions */
	int32_t M;       /* sub-quantizers; dsub = D / M */
	int64_t n_indexed; /* rows covered by the index: labels [0, n_indexed) */
	const float *centroids;     /* [P][D]   (cosine: trained on normalised rows) */
	const float *codebook;      /* [M][256][dsub] (L2/cosine: trained on residuals) */
	const int64_t *part_offsets; /* [P+1] into row_ids / codes */
	const int64_t *row_ids;     /* [n_indexed] labels, partition-major */
	const uint8_t *codes;       /* [n_indexed][M], partition-major */
} oracle_ivfpq_t;

/* a12: the nprobes nearest partitions, ascending (distance, partition id). */
int oracle_ivf_probe(const oracle_ivfpq_t *ix, const float *q, int nprobes, int32_t *out_parts,
                     float *out_dists);

/* a13: lookup table for one (query, partition).  L2/cosine: residual q' - centroid_p where q' is
 * q (L2) or q/|q| (cosine); entry = |r_m - cb[m][c]|^2.  dot: entry = -(q_m . cb[m][c]). */
void oracle_pq_lut(const oracle_ivfpq_t *ix, const float *q, int part, float *lut /*[M*256]*/);

/* ADC distance of a stored code under a LUT, in the index metric's reported units:
 * L2: sum; cosine: sum/2; dot: 1 + sum. */
float oracle_adc_distance(const oracle_ivfpq_t *ix, const float *lut, const uint8_t *code);

/* a12-a15 end to end.  Rows with label >= n_indexed (appended after the index was built) are
 * scanned exactly and merged, as Lance does for unindexed fragments.
 * refine_factor >= 1: the best k*rf ADC candidates are re-ranked with exact distances (F6);
 * refine_factor == 0: ADC distances are returned as they are.
 * out_cand_* (nullable) receive the pre-refine candidate list (k*max(rf,1) entries max). */
int oracle_ivfpq_search(const oracle_ivfpq_t *ix, const float *X, int64_t n, const float *q, int k,
                        int nprobes, int refine_factor, const uint32_t *allow,
                        const uint32_t *deleted, int32_t *out_parts, int64_t *out_cand_labels,
                        float *out_cand_adc, int32_t *out_ncand, int64_t *out_labels,
                        float *out_dist);

/* Same search over a GIVEN list of partitions (no probe).  Used when the device's probe and the fp64
 * probe legally differ at a tie (two centroids within 1e-5 relative): the rest of the pipeline is then
 * still compared on identical lists. */
int oracle_ivfpq_search_parts(const oracle_ivfpq_t *ix, const float *X, int64_t n, const float *q, int k,
                              const int32_t *parts, int nparts, int refine_factor,
                              const uint32_t *allow, const uint32_t *deleted, int64_t *out_cand_labels,
                              float *out_cand_adc, int32_t *out_ncand, int64_t *out_labels,
                              float *out_dist);

/* ---------------- chunked truth accumulator (full-size parity tests) ----------------
 * oracle_flat_search for several queries over a table that arrives in row chunks: fp64 distances,
 * ranking by (distance, label), OpenMP over the rows of a chunk.  Chunk rows are table rows
 * [row0, row0 + n); bitmaps are indexed by table row; label = label_base + table row; rows at or
 * beyond allow_bits are filtered out when `allow` is given (the C ABI's bitmap_bits rule). */
typedef struct oracle_acc oracle_acc_t;
oracle_acc_t *oracle_acc_new(int metric, const float *Q, int nq, int d, int k);
void oracle_acc_add(oracle_acc_t *a, const float *X, int64_t n, int64_t row0, const uint32_t *allow,
                    int64_t allow_bits, const uint32_t *deleted, int64_t label_base, int nthreads);
/* out_labels / out_dist64 are [nq][k], out_counts [nq]; frees the accumulator */
void oracle_acc_finish(oracle_acc_t *a, int64_t *out_labels, double *out_dist64, int32_t *out_counts);

/* Cache-blocked SGEMM-style fp32 batch search (6 x 16 register tiles over query panels), the timed
 * CPU baseline of SURVEY.md 8(d).  Same outputs as oracle_flat_search_batch_f32; L2 distances come
 * from |q|^2 + |x|^2 - 2 q.x (so they carry the cancellation error of that form: ~1e-6 of |q|^2+|x|^2). */
int oracle_flat_search_batch_blocked(const float *X, int64_t n, int d, int metric, const float *Q, int nq,
                                     int k, const uint32_t *allow, const uint32_t *deleted,
                                     int64_t *out_labels, float *out_dist, int32_t *out_counts,
                                     int nthreads);

/* ---------------- training (so that small indexes can be built on the CPU) ---------------- */
/* Lloyd k-means, deterministic: initial centres = rows floor(i*n/k); `iters` full passes;
 * an empty cluster keeps its previous centre.  metric selects the assignment distance
 * (L2 for L2/cosine-after-normalisation, dot = max inner product). */
void oracle_kmeans(const float *X, int64_t n, int d, int k, int iters, int metric,
                   float *centroids /*[k][d] out*/, int32_t *assign /*[n] out, nullable*/);

/* Build a complete index over rows [0,n): centroids, codebook, codes, partition order.
 * Buffers are caller-allocated with the sizes in oracle_ivfpq_t. */
void oracle_ivfpq_train(const float *X, int64_t n, int d, int metric, int P, int M, int iters,
                        float *centroids, float *codebook, int64_t *part_offsets,
                        int64_t *row_ids, uint8_t *codes);

#ifdef __cplusplus
}
#endif
#endif
