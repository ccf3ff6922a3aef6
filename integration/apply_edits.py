#!/usr/bin/env python
"""apply_edits.py SRC_DIR — applies the reference-side edits of INTEGRATION.md to a COPY of the extension's `src/`
(never to /root/reference).  Used once to produce integration/patch/*.patch; kept so the patches can be regenerated.
Every edit is an exact-match replacement that fails loudly when the reference text is not what it was written against
(commit 27066f8c)."""
import sys

base = sys.argv[1].rstrip("/") + "/"


def edit(path, pairs):
    s = open(base + path).read()
    for old, new in pairs:
        assert s.count(old) >= 1, (path, old[:80])
        s = s.replace(old, new, 1)
    open(base + path, "w").write(s)


# ---------------------------------------------------------------- include/rust_ffi.hpp
edit("include/rust_ffi.hpp", [
    ("int64_t LanceDetachedCount(LanceHandle handle);",
     """// Search with the pushed-down WHERE (nullable Lance SQL string).  liblance_b200 evaluates it on the GPU over the
// index's extra columns (include/lance_b200.h: lance_detached_search_predicate).
int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const char *predicate, int64_t *out_labels, float *out_distances);

// Query batch: queries[nq][dim] -> out_labels / out_distances [nq][k], out_counts[nq].  One H2D, one pass, one D2H.
int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                 float *out_distances, int32_t *out_counts);

int64_t LanceDetachedCount(LanceHandle handle);"""),
    ("void LanceDetachedCompact(LanceHandle handle);",
     """void LanceDetachedCreateScalarIndex(LanceHandle handle, const std::string &column, const std::string &index_type);
void LanceDetachedCompact(LanceHandle handle);
// Writes the device store (vectors, tombstones, IVF_PQ index, extra columns) under the dataset path.
void LanceDetachedSave(LanceHandle handle);"""),
])

# ---------------------------------------------------------------- rust_ffi.cpp
edit("rust_ffi.cpp", [
    ("""int64_t lance_detached_count(void *handle, char *err_buf, int err_buf_len);
int32_t lance_detached_delete(""",
     """int32_t lance_detached_search_predicate(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                        int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                        float *out_distances, char *err_buf, int err_buf_len);
int32_t lance_detached_search_batch_predicate(void *handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                              int32_t nprobes, int32_t refine_factor, const char *predicate,
                                              int64_t *out_labels, float *out_distances, int32_t *out_counts,
                                              char *err_buf, int err_buf_len);
int32_t lance_detached_create_scalar_index(void *handle, const char *column, const char *index_type, char *err_buf,
                                           int err_buf_len);
int32_t lance_b200_save(void *handle, char *err_buf, int err_buf_len);
int64_t lance_detached_count(void *handle, char *err_buf, int err_buf_len);
int32_t lance_detached_delete("""),
    ("int64_t LanceDetachedCount(LanceHandle handle) {",
     """int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const char *predicate, int64_t *out_labels, float *out_distances) {
	char err_buf[ERR_BUF_LEN] = {0};
	int32_t n = lance_detached_search_predicate(handle, query, dim, k, nprobes, refine_factor, predicate, out_labels,
	                                            out_distances, err_buf, ERR_BUF_LEN);
	if (n < 0) {
		throw IOException("Lance search: " + std::string(err_buf));
	}
	return n;
}

int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                 float *out_distances, int32_t *out_counts) {
	char err_buf[ERR_BUF_LEN] = {0};
	int32_t n = lance_detached_search_batch_predicate(handle, queries, nq, dim, k, nprobes, refine_factor, predicate,
	                                                  out_labels, out_distances, out_counts, err_buf, ERR_BUF_LEN);
	if (n < 0) {
		throw IOException("Lance search: " + std::string(err_buf));
	}
	return n;
}

void LanceDetachedCreateScalarIndex(LanceHandle handle, const std::string &column, const std::string &index_type) {
	char err_buf[ERR_BUF_LEN] = {0};
	if (lance_detached_create_scalar_index(handle, column.c_str(), index_type.c_str(), err_buf, ERR_BUF_LEN) != 0) {
		throw IOException("Lance create_scalar_index: " + std::string(err_buf));
	}
}

void LanceDetachedSave(LanceHandle handle) {
	char err_buf[ERR_BUF_LEN] = {0};
	if (lance_b200_save(handle, err_buf, ERR_BUF_LEN) != 0) {
		throw IOException("Lance save: " + std::string(err_buf));
	}
}

int64_t LanceDetachedCount(LanceHandle handle) {"""),
])

# ---------------------------------------------------------------- include/lance_index.hpp
edit("include/lance_index.hpp", [
    ("	vector<pair<row_t, float>> Search(const float *query, int32_t dimension, int32_t k);\n",
     """	// `predicate`: pushed-down WHERE as a Lance SQL string (lance_optimizer.cpp), evaluated as a prefilter
	vector<pair<row_t, float>> Search(const float *query, int32_t dimension, int32_t k, const string &predicate = "");
	// One call for a whole query batch: queries[nq][dimension]; result[q] as Search()
	vector<vector<pair<row_t, float>>> SearchBatch(const float *queries, int32_t nq, int32_t dimension, int32_t k,
	                                               const string &predicate = "");
"""),
    ("	void CreateHnswIndex(int32_t m, int32_t ef_construction);\n",
     """	void CreateHnswIndex(int32_t m, int32_t ef_construction);
	void CreateScalarIndex(const string &column, const string &index_type);
"""),
])

# ---------------------------------------------------------------- lance_index.cpp
edit("lance_index.cpp", [
    ("void LanceIndex::CreateAnnIndex(int32_t num_partitions, int32_t num_sub_vectors) {",
     """vector<vector<pair<row_t, float>>> LanceIndex::SearchBatch(const float *queries, int32_t nq, int32_t dimension,
                                                           int32_t k, const string &predicate) {
	vector<vector<pair<row_t, float>>> results(nq > 0 ? nq : 0);
	if (!rust_handle_ || dimension != dimension_ || nq <= 0 || k <= 0) {
		return results;
	}

	// One FFI call for the whole batch: a single H2D of the queries, one tensor-core pass over the table
	// (or one batched IVF_PQ probe + scan), a single D2H of the packed results.
	vector<int64_t> labels(static_cast<size_t>(nq) * k);
	vector<float> distances(static_cast<size_t>(nq) * k);
	vector<int32_t> counts(nq);
	LanceDetachedSearchBatch(rust_handle_, queries, nq, dimension, k, nprobes_, refine_factor_,
	                         predicate.empty() ? nullptr : predicate.c_str(), labels.data(), distances.data(),
	                         counts.data());

	for (int32_t q = 0; q < nq; q++) {
		results[q].reserve(counts[q]);
		for (int32_t i = 0; i < counts[q]; i++) {
			auto label = labels[static_cast<size_t>(q) * k + i];
			if (label >= 0 && label < static_cast<int64_t>(label_to_rowid_.size())) {
				results[q].emplace_back(label_to_rowid_[label], distances[static_cast<size_t>(q) * k + i]);
			}
		}
	}
	return results;
}

void LanceIndex::CreateAnnIndex(int32_t num_partitions, int32_t num_sub_vectors) {"""),
    ("""	writer.Write(reinterpret_cast<const uint8_t *>(lance_path.data()), path_len);

	is_dirty_ = false;""",
     """	writer.Write(reinterpret_cast<const uint8_t *>(lance_path.data()), path_len);

	// Lance commits every add/delete to its dataset; the device store lives in HBM and is written here, next to the
	// metadata, so that LoadFromStorage (LanceOpenDetached) finds vectors, tombstones, IVF_PQ index and extra columns.
	LanceDetachedSave(rust_handle_);

	is_dirty_ = false;"""),
])
