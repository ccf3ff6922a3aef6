/*
 * lance_b200_host.h — C test-driver ABI over the C++ host mirror (duckdb-lancedb_b200/csrc/host/lance_host.hpp).
 *
 * The product boundary is include/lance_b200.h.  This header only lets tests (ctypes) drive the C++ classes that
 * mirror the extension's own host code on the path — LanceIndex::Search (src/lance_index.cpp:442-465), the
 * lance_search table function (src/lance_search.cpp:33-104), the optimizer's replacement scan and predicate
 * push-down (src/lance_optimizer.cpp:35-109, 204-344) — so that the reference's SQL tests can be replayed
 * against them.  Errors: -1 / NULL and a message in err_buf, as everywhere else.
 */
#ifndef LANCE_B200_HOST_H
#define LANCE_B200_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* extra columns (host side): a standalone set, or the one owned by an index (lbh_index_columns) */
void *lbh_columns_create(void);
void lbh_columns_free(void *cols);
/* append `count` values to column `name` (created on first use); valid[i]==0 means NULL; valid may be NULL */
int32_t lbh_columns_append_i64(void *cols, const char *name, const int64_t *values, const uint8_t *valid, int64_t count);
int32_t lbh_columns_append_f64(void *cols, const char *name, const double *values, const uint8_t *valid, int64_t count);
int32_t lbh_columns_append_bool(void *cols, const char *name, const uint8_t *values, const uint8_t *valid, int64_t count);
int32_t lbh_columns_append_str(void *cols, const char *name, const char *const *values, int64_t count); /* NULL ptr = NULL */
/* predicate string (grammar of src/lance_optimizer.cpp:204-344) -> allow bitmap of ceil(n/32) words; CPU only */
int32_t lbh_eval_predicate(void *cols, const char *predicate, int64_t n, uint32_t *out_bitmap, char *err_buf, int err_buf_len);

/* src/lance_optimizer.cpp:350-371: returns 1 and writes "l2" / "cosine" / "dot" for an array distance function, else 0 */
int32_t lbh_distance_function_metric(const char *function_name, char *out_metric, int out_len);
int32_t lbh_metric_matches_index(const char *function_metric, const char *index_metric);

/* LanceIndex (needs a CUDA device) */
void *lbh_index_create(const char *name, int32_t dimension, const char *metric, int32_t nprobes, int32_t refine_factor,
                       char *err_buf, int err_buf_len);
void lbh_index_free(void *index);
void *lbh_index_columns(void *index);
int32_t lbh_index_append(void *index, const float *vectors, int64_t count, const int64_t *row_ids, char *err_buf, int err_buf_len);
int32_t lbh_index_delete(void *index, const int64_t *row_ids, int64_t count, char *err_buf, int err_buf_len);
/* LanceIndex::Vacuum (src/lance_index.cpp:728-751) */
int32_t lbh_index_vacuum(void *index, char *err_buf, int err_buf_len);
/* LanceIndex::MergeIndexes (src/lance_index.cpp:624-668): live rows of `other` into `index`; returns the merged count */
int32_t lbh_index_merge(void *index, void *other, char *err_buf, int err_buf_len);
int64_t lbh_index_count(void *index, char *err_buf, int err_buf_len);
int32_t lbh_index_create_ann(void *index, int32_t num_partitions, int32_t num_sub_vectors, char *err_buf, int err_buf_len);
/* LanceIndex::Search: returns the result count (0 on dimension mismatch), row ids ascending by distance */
int32_t lbh_index_search(void *index, const float *query, int32_t dimension, int32_t k, const char *predicate,
                         int64_t *out_row_ids, float *out_distances, char *err_buf, int err_buf_len);
int32_t lbh_index_search_batch(void *index, const float *queries, int32_t nq, int32_t dimension, int32_t k,
                               const char *predicate, int64_t *out_row_ids, float *out_distances, int32_t *out_counts,
                               char *err_buf, int err_buf_len);
/* lance_search(table, index, query, k): bind + init + scan-until-empty; out buffers hold k entries;
 * *out_chunks = number of non-empty scan calls (<= 2048 rows each).  Returns the total row count. */
int32_t lbh_lance_search(void *index, const float *query, int32_t query_len, int32_t k, int64_t *out_row_ids,
                         float *out_distances, int32_t *out_chunks, char *err_buf, int err_buf_len);
/* lance_index_scan (the ORDER BY .. LIMIT rewrite): row ids in result order */
int32_t lbh_lance_index_scan(void *index, const float *query, int32_t query_len, int32_t limit, const char *predicate,
                             int64_t *out_row_ids, char *err_buf, int err_buf_len);

#ifdef __cplusplus
}
#endif
#endif
