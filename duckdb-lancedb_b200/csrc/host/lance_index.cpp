// lance_index.cpp — throwing FFI wrappers, LanceIndex and the lance_search / lance_index_scan functions of the
// host mirror (see lance_host.hpp for the reference lines each part follows).
#include "lance_host.hpp"

#include <algorithm>
#include <cstdlib>

#include "../../../include/lance_b200.h"

namespace lance_host {

static constexpr int ERR_BUF = 2048; // src/rust_ffi.cpp:46

// ------------------------------------------------------------------ wrappers (src/rust_ffi.cpp:44-206)
LanceHandle LanceCreateDetached(const std::string &db_path, int32_t dimension, const std::string &metric,
                                const std::string &table_name) {
	char err[ERR_BUF] = {0};
	LanceHandle h = lance_create_detached(db_path.c_str(), dimension, metric.c_str(), table_name.c_str(), err, ERR_BUF);
	if (!h) throw IOException(std::string("Lance create: ") + err);
	return h;
}
LanceHandle LanceOpenDetached(const std::string &db_path, const std::string &table_name, const std::string &metric) {
	char err[ERR_BUF] = {0};
	LanceHandle h = lance_open_detached(db_path.c_str(), table_name.c_str(), metric.c_str(), err, ERR_BUF);
	if (!h) throw IOException(std::string("Lance open: ") + err);
	return h;
}
void LanceFreeDetached(LanceHandle handle) {
	if (handle) lance_free_detached(handle);
}
int32_t LanceDetachedDimension(LanceHandle handle) { return lance_detached_dimension(handle); }
int64_t LanceDetachedAdd(LanceHandle handle, const float *vector, int32_t dimension) {
	char err[ERR_BUF] = {0};
	int64_t label = lance_detached_add(handle, vector, dimension, err, ERR_BUF);
	if (label < 0) throw IOException(std::string("Lance add: ") + err);
	return label;
}
int32_t LanceDetachedAddBatch(LanceHandle handle, const float *vectors, int32_t num, int32_t dim, int64_t *out_labels) {
	char err[ERR_BUF] = {0};
	int32_t n = lance_detached_add_batch(handle, vectors, num, dim, out_labels, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance add_batch: ") + err);
	return n;
}
int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                            int64_t *out_labels, float *out_distances) {
	char err[ERR_BUF] = {0};
	int32_t n = allow_bitmap ? lance_detached_search_filtered(handle, query, dim, k, nprobes, refine_factor, allow_bitmap,
	                                                          bitmap_bits, out_labels, out_distances, err, ERR_BUF)
	                         : lance_detached_search(handle, query, dim, k, nprobes, refine_factor, out_labels,
	                                                 out_distances, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance search: ") + err);
	return n;
}
// the newer revision's signature (called at src/lance_index.cpp:452-453): predicate string, nullable
int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const char *predicate, int64_t *out_labels, float *out_distances) {
	char err[ERR_BUF] = {0};
	int32_t n = lance_detached_search_predicate(handle, query, dim, k, nprobes, refine_factor, predicate, out_labels,
	                                            out_distances, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance search: ") + err);
	return n;
}
int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                 float *out_distances, int32_t *out_counts) {
	char err[ERR_BUF] = {0};
	int32_t n = lance_detached_search_batch_predicate(handle, queries, nq, dim, k, nprobes, refine_factor, predicate,
	                                                  out_labels, out_distances, out_counts, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance search: ") + err);
	return n;
}
int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const uint32_t *allow_bitmap,
                                 int64_t bitmap_bits, int64_t *out_labels, float *out_distances, int32_t *out_counts) {
	char err[ERR_BUF] = {0};
	int32_t n = lance_detached_search_batch(handle, queries, nq, dim, k, nprobes, refine_factor, allow_bitmap, bitmap_bits,
	                                        out_labels, out_distances, out_counts, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance search: ") + err);
	return n;
}
int64_t LanceDetachedCount(LanceHandle handle) {
	char err[ERR_BUF] = {0};
	int64_t n = lance_detached_count(handle, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance count: ") + err);
	return n;
}
void LanceDetachedDelete(LanceHandle handle, int64_t label) {
	char err[ERR_BUF] = {0};
	if (lance_detached_delete(handle, label, err, ERR_BUF) != 0) throw IOException(std::string("Lance delete: ") + err);
}
void LanceDetachedDeleteBatch(LanceHandle handle, const int64_t *labels, int32_t count) {
	char err[ERR_BUF] = {0};
	if (lance_detached_delete_batch(handle, labels, count, err, ERR_BUF) != 0)
		throw IOException(std::string("Lance delete_batch: ") + err);
}
int32_t LanceDetachedMerge(LanceHandle target, LanceHandle source, const int64_t *live_source_labels, int32_t live_count,
                           int64_t *out_old_labels, int64_t *out_new_labels) {
	char err[ERR_BUF] = {0};
	const int32_t n = lance_detached_merge(target, source, live_source_labels, live_count, out_old_labels, out_new_labels,
	                                       err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance merge: ") + err);
	return n;
}
void LanceDetachedCreateIndex(LanceHandle handle, int32_t num_partitions, int32_t num_sub_vectors) {
	char err[ERR_BUF] = {0};
	if (lance_detached_create_index(handle, num_partitions, num_sub_vectors, err, ERR_BUF) != 0)
		throw IOException(std::string("Lance create_index: ") + err);
}
void LanceDetachedCreateHnswIndex(LanceHandle handle, int32_t m, int32_t ef_construction) {
	char err[ERR_BUF] = {0};
	if (lance_detached_create_hnsw_index(handle, m, ef_construction, err, ERR_BUF) != 0)
		throw IOException(std::string("Lance create_hnsw_index: ") + err);
}
void LanceDetachedCompact(LanceHandle handle) {
	char err[ERR_BUF] = {0};
	if (lance_detached_compact(handle, err, ERR_BUF) != 0) throw IOException(std::string("Lance compact: ") + err);
}
int32_t LanceDetachedGetVector(LanceHandle handle, int64_t label, float *out_vec, int32_t capacity) {
	char err[ERR_BUF] = {0};
	int32_t n = lance_detached_get_vector(handle, label, out_vec, capacity, err, ERR_BUF);
	if (n < 0) throw IOException(std::string("Lance get_vector: ") + err);
	return n;
}

// ------------------------------------------------------------------ optimizer helpers (src/lance_optimizer.cpp:350-371)
bool IsArrayDistanceFunction(const std::string &name) {
	return name == "array_distance" || name == "array_cosine_distance" || name == "array_inner_product";
}
std::string DistanceFunctionToMetric(const std::string &name) {
	if (name == "array_cosine_distance") return "cosine";
	if (name == "array_inner_product") return "dot";
	return "l2";
}
bool MetricMatchesIndex(const std::string &fm, const std::string &im) {
	auto norm = [](const std::string &m) { return m == "ip" ? std::string("dot") : m; };
	return norm(fm) == norm(im);
}

// ------------------------------------------------------------------ LanceIndex
LanceIndex::LanceIndex(const std::string &name, int32_t dimension, const std::map<std::string, std::string> &options,
                       const std::string &db_path, const std::string &table_name)
    : name_(name), dimension_(dimension) {
	for (auto &kv : options) { // src/lance_index.cpp:157-165
		if (kv.first == "metric") metric_ = kv.second;
		else if (kv.first == "nprobes") nprobes_ = std::atoi(kv.second.c_str());
		else if (kv.first == "refine_factor") refine_factor_ = std::atoi(kv.second.c_str());
	}
	rust_handle_ = LanceCreateDetached(db_path, dimension_, metric_, table_name);
}
LanceIndex::~LanceIndex() { LanceFreeDetached(rust_handle_); }

int64_t LanceIndex::GetVectorCount() const { return rust_handle_ ? LanceDetachedCount(rust_handle_) : 0; }

void LanceIndex::Append(const float *vectors, int64_t count, const row_t *row_ids) {
	// one FFI call per <= 2048-row chunk, as the sink does (src/lance_index.cpp:941-954)
	std::vector<int64_t> labels;
	for (int64_t off = 0; off < count; off += STANDARD_VECTOR_SIZE) {
		const int32_t n = (int32_t)std::min<int64_t>(STANDARD_VECTOR_SIZE, count - off);
		labels.resize(n);
		LanceDetachedAddBatch(rust_handle_, vectors + (size_t)off * dimension_, n, dimension_, labels.data());
		for (int32_t i = 0; i < n; i++) {
			const int64_t label = labels[i];
			if ((int64_t)label_to_rowid_.size() <= label) label_to_rowid_.resize((size_t)label + 1, -1);
			label_to_rowid_[label] = row_ids[off + i];
			rowid_to_label_[row_ids[off + i]] = label;
		}
	}
}

void LanceIndex::Delete(const row_t *row_ids, int64_t count) {
	std::vector<int64_t> labels;
	for (int64_t i = 0; i < count; i++) {
		auto it = rowid_to_label_.find(row_ids[i]);
		if (it == rowid_to_label_.end()) continue;
		labels.push_back(it->second);
		// the slot is dead in the forward mapping too (:410-413)
		if (it->second >= 0 && (size_t)it->second < label_to_rowid_.size()) label_to_rowid_[(size_t)it->second] = (row_t)-1;
		rowid_to_label_.erase(it);
	}
	if (!labels.empty()) {
		LanceDetachedDeleteBatch(rust_handle_, labels.data(), (int32_t)labels.size());
		has_pending_deletes_ = true; // :419-421
	}
}

// src/lance_index.cpp:728-751: compact the backing store, then rebuild label_to_rowid_ from the authoritative
// rowid_to_label_ map (drops the dead slots at its tail; labels themselves never change)
void LanceIndex::Vacuum() {
	if (!has_pending_deletes_ || !rust_handle_) return;
	LanceDetachedCompact(rust_handle_);
	has_pending_deletes_ = false;
	std::vector<row_t> new_map;
	for (auto &kv : rowid_to_label_) {
		if (kv.second < 0) continue;
		if ((size_t)kv.second >= new_map.size()) new_map.resize((size_t)kv.second + 1, (row_t)-1);
		new_map[(size_t)kv.second] = kv.first;
	}
	label_to_rowid_ = std::move(new_map);
}

static void AppendColumnValue(Column &dst, const Column *src, size_t at) {
	// NULL when the source index does not have the column or the value is NULL
	const bool ok = src && at < src->size() && src->valid[at];
	dst.valid.push_back(ok ? 1 : 0);
	switch (dst.type) {
	case ColumnType::VARCHAR: dst.s.push_back(ok && at < src->s.size() ? src->s[at] : std::string()); break;
	case ColumnType::DOUBLE: dst.d.push_back(ok && at < src->d.size() ? src->d[at] : 0.0); break;
	default: dst.i.push_back(ok && at < src->i.size() ? src->i[at] : 0); break;
	}
}

int32_t LanceIndex::MergeIndexes(LanceIndex &other) {
	if (!rust_handle_ || !other.rust_handle_) return 0; // :626-629
	std::vector<int64_t> live_labels;
	std::vector<row_t> live_rowids;
	for (size_t i = 0; i < other.label_to_rowid_.size(); i++) {
		if (other.label_to_rowid_[i] == (row_t)-1) continue;
		live_labels.push_back((int64_t)i);
		live_rowids.push_back(other.label_to_rowid_[i]);
	}
	if (live_labels.empty()) return 0;
	const int32_t count = (int32_t)live_labels.size();
	std::vector<int64_t> out_old(count), out_new(count);
	const int32_t n = LanceDetachedMerge(rust_handle_, other.rust_handle_, live_labels.data(), count, out_old.data(),
	                                     out_new.data());
	std::unordered_map<int64_t, row_t> old_label_to_rowid;
	for (size_t i = 0; i < live_labels.size(); i++) old_label_to_rowid[live_labels[i]] = live_rowids[i];
	for (int32_t i = 0; i < n; i++) {
		auto it = old_label_to_rowid.find(out_old[i]);
		if (it == old_label_to_rowid.end()) continue;
		const int64_t new_label = out_new[i];
		if ((int64_t)label_to_rowid_.size() <= new_label) label_to_rowid_.resize((size_t)new_label + 1, (row_t)-1);
		label_to_rowid_[(size_t)new_label] = it->second;
		rowid_to_label_[it->second] = new_label;
		// extra columns are indexed by label: pad to the new label, then take the source row's values
		for (auto &kv : columns_) {
			Column &c = kv.second;
			while ((int64_t)c.size() < new_label) AppendColumnValue(c, nullptr, 0);
			auto oc = other.columns_.find(kv.first);
			const Column *src = (oc != other.columns_.end() && oc->second.type == c.type) ? &oc->second : nullptr;
			if ((int64_t)c.size() == new_label) AppendColumnValue(c, src, (size_t)out_old[i]);
		}
	}
	return n;
}

// The index's extra columns travel to the device store (the reference ships them to Lance with every Append,
// src/lance_index.cpp:321-359); here the mirror pushes what was appended to `columns_` since the last search.
void LanceIndex::SyncColumnsToDevice() {
	std::lock_guard<std::mutex> g(predicate_mu_);
	const int64_t n_labels = (int64_t)label_to_rowid_.size();
	for (auto &kv : columns_) {
		const Column &c = kv.second;
		int64_t &done = synced_rows_[kv.first];
		const int64_t upto = std::min<int64_t>((int64_t)c.size(), n_labels);
		if (upto <= done) continue;
		const int64_t cnt = upto - done;
		char err[ERR_BUF] = {0};
		int32_t rc = 0;
		switch (c.type) {
		case ColumnType::BIGINT:
			rc = lance_b200_set_column(rust_handle_, kv.first.c_str(), 0, done, cnt, c.i.data() + done, nullptr, c.valid.data() + done, err, ERR_BUF);
			break;
		case ColumnType::DOUBLE:
			rc = lance_b200_set_column(rust_handle_, kv.first.c_str(), 1, done, cnt, c.d.data() + done, nullptr, c.valid.data() + done, err, ERR_BUF);
			break;
		case ColumnType::BOOLEAN: {
			std::vector<uint8_t> b((size_t)cnt);
			for (int64_t i = 0; i < cnt; i++) b[(size_t)i] = c.i[(size_t)(done + i)] ? 1 : 0;
			rc = lance_b200_set_column(rust_handle_, kv.first.c_str(), 2, done, cnt, b.data(), nullptr, c.valid.data() + done, err, ERR_BUF);
			break;
		}
		case ColumnType::VARCHAR: {
			std::string bytes;
			std::vector<int64_t> off((size_t)cnt + 1, 0);
			for (int64_t i = 0; i < cnt; i++) {
				bytes += c.s[(size_t)(done + i)];
				off[(size_t)i + 1] = (int64_t)bytes.size();
			}
			bytes.push_back('\0');
			rc = lance_b200_set_column(rust_handle_, kv.first.c_str(), 3, done, cnt, bytes.data(), off.data(), c.valid.data() + done, err, ERR_BUF);
			break;
		}
		}
		if (rc < 0) throw IOException(std::string("Lance set_column: ") + err);
		done = upto;
	}
}

std::vector<std::pair<row_t, float>> LanceIndex::Search(const float *query, int32_t dimension, int32_t k,
                                                        const std::string &predicate) {
	if (!rust_handle_ || dimension != dimension_) return {}; // :444-446
	if (k <= 0) return {};
	std::vector<int64_t> labels(k);
	std::vector<float> distances(k);
	if (!predicate.empty()) SyncColumnsToDevice();
	// :452-453 verbatim: the predicate string goes through the FFI; the device evaluates it
	const int32_t n = LanceDetachedSearch(rust_handle_, query, dimension, k, nprobes_, refine_factor_,
	                                      predicate.empty() ? nullptr : predicate.c_str(), labels.data(), distances.data());
	std::vector<std::pair<row_t, float>> results;
	results.reserve(n);
	for (int32_t i = 0; i < n; i++) {
		const int64_t label = labels[i];
		if (label >= 0 && label < (int64_t)label_to_rowid_.size()) results.emplace_back(label_to_rowid_[label], distances[i]);
	}
	return results;
}

std::vector<std::vector<std::pair<row_t, float>>> LanceIndex::SearchBatch(const float *queries, int32_t nq,
                                                                          int32_t dimension, int32_t k,
                                                                          const std::string &predicate) {
	std::vector<std::vector<std::pair<row_t, float>>> out((size_t)std::max(nq, 0));
	if (!rust_handle_ || dimension != dimension_ || k <= 0 || nq <= 0) return out;
	std::vector<int64_t> labels((size_t)nq * k);
	std::vector<float> distances((size_t)nq * k);
	std::vector<int32_t> counts(nq);
	if (!predicate.empty()) SyncColumnsToDevice();
	LanceDetachedSearchBatch(rust_handle_, queries, nq, dimension, k, nprobes_, refine_factor_,
	                         predicate.empty() ? nullptr : predicate.c_str(), labels.data(), distances.data(), counts.data());
	for (int32_t q = 0; q < nq; q++) {
		out[q].reserve(counts[q]);
		for (int32_t i = 0; i < counts[q]; i++) {
			const int64_t label = labels[(size_t)q * k + i];
			if (label >= 0 && label < (int64_t)label_to_rowid_.size())
				out[q].emplace_back(label_to_rowid_[label], distances[(size_t)q * k + i]);
		}
	}
	return out;
}

void LanceIndex::CreateAnnIndex(int32_t num_partitions, int32_t num_sub_vectors) {
	if (!rust_handle_) throw IOException("Lance index not initialized");
	LanceDetachedCreateIndex(rust_handle_, num_partitions, num_sub_vectors);
}
void LanceIndex::CreateHnswIndex(int32_t m, int32_t ef_construction) {
	if (!rust_handle_) throw IOException("Lance index not initialized");
	LanceDetachedCreateHnswIndex(rust_handle_, m, ef_construction);
}

// ------------------------------------------------------------------ lance_search (src/lance_search.cpp)
std::unique_ptr<LanceSearchBindData> LanceSearchBind(LanceIndex *index, const std::vector<float> &query, int32_t k) {
	if (!index) throw BinderException("lance_search: index not found");
	auto bind = std::unique_ptr<LanceSearchBindData>(new LanceSearchBindData());
	bind->index = index;
	bind->query = query;
	bind->k = k;
	return bind;
}
std::unique_ptr<LanceSearchState> LanceSearchInit(const LanceSearchBindData &bind) {
	auto state = std::unique_ptr<LanceSearchState>(new LanceSearchState());
	auto results = bind.index->Search(bind.query.data(), (int32_t)bind.query.size(), bind.k);
	state->row_ids.reserve(results.size());
	state->distances.reserve(results.size());
	for (auto &r : results) {
		state->row_ids.push_back(r.first);
		state->distances.push_back(r.second);
	}
	return state;
}
int64_t LanceSearchScan(LanceSearchState &state, int64_t *out_row_ids, float *out_distances) {
	const size_t remaining = state.row_ids.size() - state.offset;
	const size_t n = std::min<size_t>(remaining, (size_t)STANDARD_VECTOR_SIZE);
	for (size_t i = 0; i < n; i++) {
		out_row_ids[i] = state.row_ids[state.offset + i];
		out_distances[i] = state.distances[state.offset + i];
	}
	state.offset += n;
	return (int64_t)n;
}

std::vector<row_t> LanceIndexScanInit(const LanceIndexScanBindData &bind) {
	auto results = bind.index->Search(bind.query_vector.data(), (int32_t)bind.query_vector.size(), bind.limit, bind.predicate);
	std::vector<row_t> ids;
	ids.reserve(results.size());
	for (auto &r : results) ids.push_back(r.first);
	return ids;
}

} // namespace lance_host
