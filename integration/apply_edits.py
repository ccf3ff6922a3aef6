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

# ---------------------------------------------------------------- lance_search.cpp
# lance_search keeps its four positional arguments and its (row_id, distance) schema; optional named parameter
# `prefilter` (a predicate over the index's extra columns, same grammar as the optimizer's pushdown).  New:
# lance_search_batch(table, index, LIST(LIST(FLOAT)), k [, prefilter := ..]) -> (query_idx, row_id, distance):
# the whole batch is one FFI call (LanceIndex::SearchBatch -> lance_detached_search_batch_predicate).
edit("lance_search.cpp", [
    ("""	vector<float> query;
	int32_t k;
};""",
     """	vector<float> query;
	int32_t k;
	string prefilter;
	//! lance_search_batch: row-major queries [nq][dim]
	int32_t nq = 0;
	int32_t dim = 0;
};"""),
    ("""	vector<row_t> row_ids;
	vector<float> distances;
	idx_t position = 0;""",
     """	vector<row_t> row_ids;
	vector<float> distances;
	vector<int32_t> query_idx; // lance_search_batch only
	idx_t position = 0;"""),
    ("""	bind_data->k = input.inputs[3].GetValue<int32_t>();

	return_types.push_back(LogicalType::BIGINT);""",
     """	bind_data->k = input.inputs[3].GetValue<int32_t>();
	auto prefilter_entry = input.named_parameters.find("prefilter");
	if (prefilter_entry != input.named_parameters.end() && !prefilter_entry->second.IsNull()) {
		bind_data->prefilter = prefilter_entry->second.GetValue<string>();
	}

	return_types.push_back(LogicalType::BIGINT);"""),
    ("""	    lance_idx.Search(bind.query.data(), static_cast<int32_t>(bind.query.size()), bind.k);""",
     """	    lance_idx.Search(bind.query.data(), static_cast<int32_t>(bind.query.size()), bind.k, bind.prefilter);"""),
])
edit("lance_search.cpp", [
    ("""void RegisterLanceSearchFunction(ExtensionLoader &loader) {""",
     """// ========================================
// lance_search_batch(table, index, [[..], [..], ..], k)
// Returns (query_idx INTEGER, row_id BIGINT, distance FLOAT), query by query, ascending distance within a query
// ========================================

static unique_ptr<FunctionData> LanceSearchBatchBind(ClientContext &context, TableFunctionBindInput &input,
                                                     vector<LogicalType> &return_types, vector<string> &names) {
	auto bind_data = make_uniq<LanceSearchBindData>();
	bind_data->table_name = input.inputs[0].GetValue<string>();
	bind_data->index_name = input.inputs[1].GetValue<string>();

	auto &queries = ListValue::GetChildren(input.inputs[2]);
	bind_data->nq = static_cast<int32_t>(queries.size());
	for (auto &query : queries) {
		if (query.IsNull()) {
			throw InvalidInputException("lance_search_batch: NULL query vector");
		}
		auto &children = ListValue::GetChildren(query);
		if (bind_data->dim == 0) {
			bind_data->dim = static_cast<int32_t>(children.size());
		} else if (static_cast<int32_t>(children.size()) != bind_data->dim) {
			throw InvalidInputException("lance_search_batch: all query vectors must have the same length");
		}
		for (auto &child : children) {
			bind_data->query.push_back(child.GetValue<float>());
		}
	}
	bind_data->k = input.inputs[3].GetValue<int32_t>();
	auto prefilter_entry = input.named_parameters.find("prefilter");
	if (prefilter_entry != input.named_parameters.end() && !prefilter_entry->second.IsNull()) {
		bind_data->prefilter = prefilter_entry->second.GetValue<string>();
	}

	return_types.push_back(LogicalType::INTEGER);
	return_types.push_back(LogicalType::BIGINT);
	return_types.push_back(LogicalType::FLOAT);
	names.push_back("query_idx");
	names.push_back("row_id");
	names.push_back("distance");
	return std::move(bind_data);
}

static unique_ptr<GlobalTableFunctionState> LanceSearchBatchInit(ClientContext &context,
                                                                 TableFunctionInitInput &input) {
	auto state = make_uniq<LanceSearchState>();
	auto &bind = input.bind_data->Cast<LanceSearchBindData>();

	auto &catalog = Catalog::GetCatalog(context, "");
	auto &table_entry = catalog.GetEntry<TableCatalogEntry>(context, DEFAULT_SCHEMA, bind.table_name);
	auto &duck_table = table_entry.Cast<DuckTableEntry>();
	auto &storage = duck_table.GetStorage();
	auto &table_info = *storage.GetDataTableInfo();
	auto &indexes = table_info.GetIndexes();

	indexes.Bind(context, table_info, LanceIndex::TYPE_NAME);

	auto index_ptr = indexes.Find(bind.index_name);
	if (!index_ptr) {
		throw InvalidInputException("Index '%s' not found on table '%s'", bind.index_name, bind.table_name);
	}

	auto &lance_idx = index_ptr->Cast<LanceIndex>();
	auto results = lance_idx.SearchBatch(bind.query.data(), bind.nq, bind.dim, bind.k, bind.prefilter);

	for (idx_t q = 0; q < results.size(); q++) {
		for (auto &result : results[q]) {
			state->query_idx.push_back(static_cast<int32_t>(q));
			state->row_ids.push_back(result.first);
			state->distances.push_back(result.second);
		}
	}

	return std::move(state);
}

static void LanceSearchBatchScan(ClientContext &context, TableFunctionInput &data, DataChunk &output) {
	auto &state = data.global_state->Cast<LanceSearchState>();

	if (state.position >= state.row_ids.size()) {
		output.SetCardinality(0);
		return;
	}

	idx_t chunk_size = MinValue<idx_t>(STANDARD_VECTOR_SIZE, state.row_ids.size() - state.position);

	auto query_data = FlatVector::GetData<int32_t>(output.data[0]);
	auto rowid_data = FlatVector::GetData<int64_t>(output.data[1]);
	auto dist_data = FlatVector::GetData<float>(output.data[2]);

	for (idx_t i = 0; i < chunk_size; i++) {
		query_data[i] = state.query_idx[state.position + i];
		rowid_data[i] = state.row_ids[state.position + i];
		dist_data[i] = state.distances[state.position + i];
	}

	state.position += chunk_size;
	output.SetCardinality(chunk_size);
}

void RegisterLanceSearchFunction(ExtensionLoader &loader) {"""),
    ("""	func.cardinality = LanceSearchCardinality;
	loader.RegisterFunction(func);""",
     """	func.cardinality = LanceSearchCardinality;
	func.named_parameters["prefilter"] = LogicalType::VARCHAR;
	loader.RegisterFunction(func);

	TableFunction batch_func("lance_search_batch",
	                         {LogicalType::VARCHAR, LogicalType::VARCHAR,
	                          LogicalType::LIST(LogicalType::LIST(LogicalType::FLOAT)), LogicalType::INTEGER},
	                         LanceSearchBatchScan, LanceSearchBatchBind, LanceSearchBatchInit);
	batch_func.named_parameters["prefilter"] = LogicalType::VARCHAR;
	loader.RegisterFunction(batch_func);"""),
])
