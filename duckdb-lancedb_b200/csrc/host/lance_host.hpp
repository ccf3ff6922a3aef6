// lance_host.hpp — host side above the C ABI: a DuckDB-free mirror of the extension's own classes on the
// kNN path, with the reference's names, argument meaning and error behaviour, so that the edits listed in
// INTEGRATION.md are mechanical.  It calls ONLY the public C ABI (include/lance_b200.h).
//
//   LanceDetached*  throwing wrappers            <- src/rust_ffi.cpp:44-206, src/include/rust_ffi.hpp:9-58
//   LanceIndex      Search / Append / Delete ... <- src/lance_index.cpp:147-190, 273-425, 442-472; include/lance_index.hpp:53,79-99
//   LanceSearch*    bind / init / scan            <- src/lance_search.cpp:17-104
//   LanceIndexScan* the optimizer's replacement   <- src/lance_optimizer.cpp:35-109
//   DistanceFunctionToMetric, predicate grammar   <- src/lance_optimizer.cpp:116-371
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

namespace lance_host {

typedef int64_t row_t; // duckdb::row_t
typedef void *LanceHandle;
static constexpr int64_t STANDARD_VECTOR_SIZE = 2048; // duckdb's chunk size (src/lance_search.cpp:88)

// duckdb::IOException("Lance <op>: <message>") of src/rust_ffi.cpp
struct IOException : std::runtime_error {
	explicit IOException(const std::string &m) : std::runtime_error(m) {}
};
struct BinderException : std::runtime_error {
	explicit BinderException(const std::string &m) : std::runtime_error(m) {}
};

// ------------------------------------------------------------------ throwing wrappers (rust_ffi.hpp)
LanceHandle LanceCreateDetached(const std::string &db_path, int32_t dimension, const std::string &metric,
                                const std::string &table_name);
LanceHandle LanceOpenDetached(const std::string &db_path, const std::string &table_name, const std::string &metric);
void LanceFreeDetached(LanceHandle handle);
int32_t LanceDetachedDimension(LanceHandle handle);
int64_t LanceDetachedAdd(LanceHandle handle, const float *vector, int32_t dimension);
int32_t LanceDetachedAddBatch(LanceHandle handle, const float *vectors, int32_t num, int32_t dim, int64_t *out_labels);
// rust_ffi.cpp:130-139; the newer revision's `predicate` (src/lance_index.cpp:452-453) arrives as a row bitmap
int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                            int64_t *out_labels, float *out_distances);
// the newer revision's form (src/lance_index.cpp:452-453): the pushed-down WHERE as a string, evaluated on the device
int32_t LanceDetachedSearch(LanceHandle handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                            int32_t refine_factor, const char *predicate, int64_t *out_labels, float *out_distances);
int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                 float *out_distances, int32_t *out_counts);
int32_t LanceDetachedSearchBatch(LanceHandle handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                 int32_t nprobes, int32_t refine_factor, const uint32_t *allow_bitmap,
                                 int64_t bitmap_bits, int64_t *out_labels, float *out_distances, int32_t *out_counts);
int64_t LanceDetachedCount(LanceHandle handle);
void LanceDetachedDelete(LanceHandle handle, int64_t label);
void LanceDetachedDeleteBatch(LanceHandle handle, const int64_t *labels, int32_t count);
// src/rust_ffi.cpp:119-128
int32_t LanceDetachedMerge(LanceHandle target, LanceHandle source, const int64_t *live_source_labels, int32_t live_count,
                           int64_t *out_old_labels, int64_t *out_new_labels);
void LanceDetachedCreateIndex(LanceHandle handle, int32_t num_partitions, int32_t num_sub_vectors);
void LanceDetachedCreateHnswIndex(LanceHandle handle, int32_t m, int32_t ef_construction);
void LanceDetachedCompact(LanceHandle handle);
int32_t LanceDetachedGetVector(LanceHandle handle, int64_t label, float *out_vec, int32_t capacity);

// ------------------------------------------------------------------ extra columns + predicate
// The index's extra columns (src/lance_index.cpp:233-245 allows VARCHAR, INTEGER, BIGINT, DOUBLE, BOOLEAN),
// staged on the host per label and pushed to the device store before a filtered search; the reference ships them to
// Lance and lets it evaluate the predicate string — the device store does the same (csrc/columns.h).  The host
// evaluator below is the CPU checker of the device one.
enum class ColumnType { VARCHAR, BIGINT, DOUBLE, BOOLEAN };
struct Column {
	ColumnType type = ColumnType::BIGINT;
	std::vector<std::string> s;
	std::vector<int64_t> i;
	std::vector<double> d; // also holds BOOLEAN as 0/1 in `i`
	std::vector<uint8_t> valid;
	size_t size() const { return valid.size(); }
};
typedef std::map<std::string, Column> ColumnSet;

// Evaluates a predicate in the grammar src/lance_optimizer.cpp:204-344 emits (comparisons col <op> const or
// const <op> col, AND / OR with parentheses, NOT (...), IS [NOT] NULL, [NOT] IN (...), BETWEEN rewritten as two
// comparisons) over `cols` for labels [0, n).  SQL three-valued logic: only TRUE keeps a row.
// Returns the allow bitmap (bit label&31 of word label>>5).  Throws BinderException on a malformed predicate.
std::vector<uint32_t> EvaluatePredicate(const std::string &predicate, const ColumnSet &cols, int64_t n);

// src/lance_optimizer.cpp:350-371
bool IsArrayDistanceFunction(const std::string &name);
std::string DistanceFunctionToMetric(const std::string &name);
bool MetricMatchesIndex(const std::string &function_metric, const std::string &index_metric); // :512-541 ("dot" == "ip")

// ------------------------------------------------------------------ LanceIndex (src/lance_index.cpp)
class LanceIndex {
public:
	// options: metric | nprobes | refine_factor (src/lance_index.cpp:157-165), defaults l2 / 20 / 1
	LanceIndex(const std::string &name, int32_t dimension, const std::map<std::string, std::string> &options,
	           const std::string &db_path = "", const std::string &table_name = "vectors");
	~LanceIndex();
	LanceIndex(const LanceIndex &) = delete;
	LanceIndex &operator=(const LanceIndex &) = delete;

	// Append (src/lance_index.cpp:273-387): `count` vectors [count][dimension] with their DuckDB row ids;
	// extra column values are appended by the caller through MutableColumns() before or after.
	void Append(const float *vectors, int64_t count, const row_t *row_ids);
	// Delete (src/lance_index.cpp:389-425): unknown row ids are ignored
	void Delete(const row_t *row_ids, int64_t count);
	// Vacuum (src/lance_index.cpp:728-751): LanceDetachedCompact when deletes are pending, then the forward map is
	// rebuilt from the reverse one
	void Vacuum();
	bool HasPendingDeletes() const { return has_pending_deletes_; }
	// ANN search (src/lance_index.cpp:442-465): dimension mismatch -> empty result (lance_basic.test:55-59);
	// labels outside label_to_rowid_ are dropped; results ascending by distance.
	std::vector<std::pair<row_t, float>> Search(const float *query, int32_t dimension, int32_t k,
	                                            const std::string &predicate = "");
	// query batch (north_star); out[q] as Search()
	std::vector<std::vector<std::pair<row_t, float>>> SearchBatch(const float *queries, int32_t nq, int32_t dimension,
	                                                             int32_t k, const std::string &predicate = "");
	// MergeIndexes (src/lance_index.cpp:624-668): the live rows of `other` move into this index on the device
	// (lance_detached_merge), label <-> row id maps follow the returned old -> new mapping, and the rows' extra
	// column values (kept on the host here) are carried over under the new labels.
	// Returns the number of merged rows.
	int32_t MergeIndexes(LanceIndex &other);
	void CreateAnnIndex(int32_t num_partitions, int32_t num_sub_vectors); // :467-472
	void CreateHnswIndex(int32_t m, int32_t ef_construction);              // :474-479

	int32_t GetDimension() const { return dimension_; }
	const std::string &GetMetric() const { return metric_; }
	const std::string &GetName() const { return name_; }
	int64_t GetVectorCount() const;
	ColumnSet &MutableColumns() { return columns_; }

private:
	std::string name_;
	LanceHandle rust_handle_ = nullptr; // the reference's field name: the C-ABI handle
	int32_t dimension_ = 0;
	std::string metric_ = "l2";
	int32_t nprobes_ = 20;
	int32_t refine_factor_ = 1;
	std::vector<row_t> label_to_rowid_;
	std::unordered_map<row_t, int64_t> rowid_to_label_;
	ColumnSet columns_;
	bool has_pending_deletes_ = false;
	// rows of every extra column already pushed to the device store (lance_b200_set_column); the predicate string
	// itself goes through the FFI as in the reference and is evaluated (and cached) on the device
	std::map<std::string, int64_t> synced_rows_;
	std::mutex predicate_mu_;
	void SyncColumnsToDevice();
};

// ------------------------------------------------------------------ lance_search table function (src/lance_search.cpp)
struct LanceSearchBindData {
	LanceIndex *index = nullptr; // the catalog lookup of :58-67 is the caller's (no catalog here)
	std::vector<float> query;
	int32_t k = 0;
};
struct LanceSearchState {
	std::vector<row_t> row_ids;
	std::vector<float> distances;
	size_t offset = 0;
};
// :33-52 — k must be positive, the query non-empty
std::unique_ptr<LanceSearchBindData> LanceSearchBind(LanceIndex *index, const std::vector<float> &query, int32_t k);
// :54-82 — the whole search runs here
std::unique_ptr<LanceSearchState> LanceSearchInit(const LanceSearchBindData &bind);
// :84-104 — emits <= 2048 rows per call into (row_id BIGINT, distance FLOAT); returns the chunk cardinality
int64_t LanceSearchScan(LanceSearchState &state, int64_t *out_row_ids, float *out_distances);

// ------------------------------------------------------------------ lance_index_scan (src/lance_optimizer.cpp:35-109)
struct LanceIndexScanBindData {
	LanceIndex *index = nullptr;
	std::vector<float> query_vector;
	int32_t limit = 0;
	std::string predicate;
};
// Init = Search (:60-82); the row fetch of :84-109 (DataTable::Fetch) stays with DuckDB: this returns the row ids
// in result order (ascending distance), which is the order the replaced ORDER BY must produce.
std::vector<row_t> LanceIndexScanInit(const LanceIndexScanBindData &bind);

} // namespace lance_host
