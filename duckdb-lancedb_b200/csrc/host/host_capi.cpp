// host_capi.cpp — extern "C" test-driver shim over the host mirror (include/lance_b200_host.h).
#include <cstring>

#include "../../../include/lance_b200_host.h"
#include "lance_host.hpp"

using namespace lance_host;

static void werr(char *buf, int len, const std::string &m) {
	if (!buf || len <= 0) return;
	size_t n = m.size() < (size_t)(len - 1) ? m.size() : (size_t)(len - 1);
	memcpy(buf, m.data(), n);
	buf[n] = '\0';
}
#define HTRY(fail, ...)                        \
	try {                                      \
		__VA_ARGS__                            \
	} catch (const std::exception &e) {        \
		werr(err_buf, err_buf_len, e.what());  \
		return fail;                           \
	} catch (...) {                            \
		werr(err_buf, err_buf_len, "unknown"); \
		return fail;                           \
	}

static Column &col_of(void *cols, const char *name, ColumnType t) {
	ColumnSet &cs = *static_cast<ColumnSet *>(cols);
	auto it = cs.find(name);
	if (it == cs.end()) {
		Column c;
		c.type = t;
		it = cs.emplace(name, c).first;
	}
	return it->second;
}

extern "C" {

void *lbh_columns_create(void) { return new ColumnSet(); }
void lbh_columns_free(void *cols) { delete static_cast<ColumnSet *>(cols); }

int32_t lbh_columns_append_i64(void *cols, const char *name, const int64_t *values, const uint8_t *valid, int64_t count) {
	Column &c = col_of(cols, name, ColumnType::BIGINT);
	for (int64_t i = 0; i < count; i++) {
		c.i.push_back(values[i]);
		c.valid.push_back(valid ? valid[i] : 1);
	}
	return 0;
}
int32_t lbh_columns_append_f64(void *cols, const char *name, const double *values, const uint8_t *valid, int64_t count) {
	Column &c = col_of(cols, name, ColumnType::DOUBLE);
	for (int64_t i = 0; i < count; i++) {
		c.d.push_back(values[i]);
		c.valid.push_back(valid ? valid[i] : 1);
	}
	return 0;
}
int32_t lbh_columns_append_bool(void *cols, const char *name, const uint8_t *values, const uint8_t *valid, int64_t count) {
	Column &c = col_of(cols, name, ColumnType::BOOLEAN);
	for (int64_t i = 0; i < count; i++) {
		c.i.push_back(values[i] ? 1 : 0);
		c.valid.push_back(valid ? valid[i] : 1);
	}
	return 0;
}
int32_t lbh_columns_append_str(void *cols, const char *name, const char *const *values, int64_t count) {
	Column &c = col_of(cols, name, ColumnType::VARCHAR);
	for (int64_t i = 0; i < count; i++) {
		c.s.push_back(values[i] ? values[i] : "");
		c.valid.push_back(values[i] ? 1 : 0);
	}
	return 0;
}
int32_t lbh_eval_predicate(void *cols, const char *predicate, int64_t n, uint32_t *out_bitmap, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		auto bm = EvaluatePredicate(predicate, *static_cast<ColumnSet *>(cols), n);
		memcpy(out_bitmap, bm.data(), (size_t)((n + 31) / 32) * sizeof(uint32_t));
		return 0;
	})
}

int32_t lbh_distance_function_metric(const char *function_name, char *out_metric, int out_len) {
	if (!IsArrayDistanceFunction(function_name)) return 0;
	werr(out_metric, out_len, DistanceFunctionToMetric(function_name));
	return 1;
}
int32_t lbh_metric_matches_index(const char *fm, const char *im) { return MetricMatchesIndex(fm, im) ? 1 : 0; }

void *lbh_index_create(const char *name, int32_t dimension, const char *metric, int32_t nprobes, int32_t refine_factor,
                       char *err_buf, int err_buf_len) {
	HTRY(nullptr, {
		std::map<std::string, std::string> opt;
		if (metric && *metric) opt["metric"] = metric;
		if (nprobes > 0) opt["nprobes"] = std::to_string(nprobes);
		if (refine_factor >= 0) opt["refine_factor"] = std::to_string(refine_factor);
		return new LanceIndex(name ? name : "idx", dimension, opt);
	})
}
void lbh_index_free(void *index) { delete static_cast<LanceIndex *>(index); }
void *lbh_index_columns(void *index) { return &static_cast<LanceIndex *>(index)->MutableColumns(); }

int32_t lbh_index_append(void *index, const float *vectors, int64_t count, const int64_t *row_ids, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		static_cast<LanceIndex *>(index)->Append(vectors, count, row_ids);
		return 0;
	})
}
int32_t lbh_index_delete(void *index, const int64_t *row_ids, int64_t count, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		static_cast<LanceIndex *>(index)->Delete(row_ids, count);
		return 0;
	})
}
int32_t lbh_index_vacuum(void *index, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		static_cast<LanceIndex *>(index)->Vacuum();
		return 0;
	})
}
int32_t lbh_index_merge(void *index, void *other, char *err_buf, int err_buf_len) {
	HTRY(-1, { return static_cast<LanceIndex *>(index)->MergeIndexes(*static_cast<LanceIndex *>(other)); })
}
int64_t lbh_index_count(void *index, char *err_buf, int err_buf_len) {
	HTRY(-1, { return static_cast<LanceIndex *>(index)->GetVectorCount(); })
}
int32_t lbh_index_create_ann(void *index, int32_t num_partitions, int32_t num_sub_vectors, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		static_cast<LanceIndex *>(index)->CreateAnnIndex(num_partitions, num_sub_vectors);
		return 0;
	})
}
int32_t lbh_index_search(void *index, const float *query, int32_t dimension, int32_t k, const char *predicate,
                         int64_t *out_row_ids, float *out_distances, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		auto res = static_cast<LanceIndex *>(index)->Search(query, dimension, k, predicate ? predicate : "");
		for (size_t i = 0; i < res.size(); i++) {
			out_row_ids[i] = res[i].first;
			out_distances[i] = res[i].second;
		}
		return (int32_t)res.size();
	})
}
int32_t lbh_index_search_batch(void *index, const float *queries, int32_t nq, int32_t dimension, int32_t k,
                               const char *predicate, int64_t *out_row_ids, float *out_distances, int32_t *out_counts,
                               char *err_buf, int err_buf_len) {
	HTRY(-1, {
		auto res = static_cast<LanceIndex *>(index)->SearchBatch(queries, nq, dimension, k, predicate ? predicate : "");
		for (int32_t q = 0; q < nq; q++) {
			out_counts[q] = (int32_t)res[q].size();
			for (size_t i = 0; i < res[q].size(); i++) {
				out_row_ids[(size_t)q * k + i] = res[q][i].first;
				out_distances[(size_t)q * k + i] = res[q][i].second;
			}
		}
		return nq;
	})
}
int32_t lbh_lance_search(void *index, const float *query, int32_t query_len, int32_t k, int64_t *out_row_ids,
                         float *out_distances, int32_t *out_chunks, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		std::vector<float> q(query, query + query_len);
		auto bind = LanceSearchBind(static_cast<LanceIndex *>(index), q, k);
		auto state = LanceSearchInit(*bind);
		std::vector<int64_t> ids((size_t)STANDARD_VECTOR_SIZE);
		std::vector<float> ds((size_t)STANDARD_VECTOR_SIZE);
		int32_t total = 0, chunks = 0;
		while (true) {
			int64_t n = LanceSearchScan(*state, ids.data(), ds.data());
			if (n == 0) break;
			memcpy(out_row_ids + total, ids.data(), (size_t)n * sizeof(int64_t));
			memcpy(out_distances + total, ds.data(), (size_t)n * sizeof(float));
			total += (int32_t)n;
			chunks++;
		}
		if (out_chunks) *out_chunks = chunks;
		return total;
	})
}
int32_t lbh_lance_index_scan(void *index, const float *query, int32_t query_len, int32_t limit, const char *predicate,
                             int64_t *out_row_ids, char *err_buf, int err_buf_len) {
	HTRY(-1, {
		LanceIndexScanBindData bind;
		bind.index = static_cast<LanceIndex *>(index);
		bind.query_vector.assign(query, query + query_len);
		bind.limit = limit;
		bind.predicate = predicate ? predicate : "";
		auto ids = LanceIndexScanInit(bind);
		for (size_t i = 0; i < ids.size(); i++) out_row_ids[i] = ids[i];
		return (int32_t)ids.size();
	})
}
}
