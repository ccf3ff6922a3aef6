// columns.h — the index's extra columns on the device and the pushed-down WHERE evaluated there.
//
// Reference side: a LANCE index may carry extra columns (VARCHAR / INTEGER / BIGINT / DOUBLE / BOOLEAN,
// src/lance_index.cpp:233-245), shipped to Lance through the Arrow C data interface
// (src/lance_index.cpp:321-359 -> rust_lib/src/ffi.rs:148-182 -> lance_manager.rs:245-305), and the optimizer
// pushes the WHERE of the rewritten query down as a Lance SQL string (src/lance_optimizer.cpp:204-344) that Lance
// evaluates as a PREFILTER (lance_manager.rs:411-419, test/sql/lance_optimizer_filter.test:23-99).
// Here: one typed array per column in HBM, row r of a column = row r of the vector table; the predicate string is
// compiled on the host into a short postfix program and one kernel evaluates it for every row under SQL
// three-valued logic, writing the allow bitmap the scans already take.  Bitmaps are cached on the device per
// (predicate, column epoch), so the steady-state cost of a filtered query is zero extra bytes over PCIe.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.cuh"

namespace lb {

enum ColType : int32_t { COL_I64 = 0, COL_F64 = 1, COL_BOOL = 2, COL_STR = 3 };

// One instruction of a compiled predicate (postfix order).  Leaves push a (true, false) bit pair — neither bit set
// = NULL; PRED_NOT / PRED_AND / PRED_OR combine the top of the stack (Kleene logic).
enum PredOpcode : int32_t {
	PRED_CMP_I64 = 0, // column (int64 storage) <cmp> ival
	PRED_CMP_F64 = 1, // column (int64 or double storage, compared as double) <cmp> dval
	PRED_IN_I64 = 2,  // column (int64 storage) IN aux_i64[aux_off .. aux_off + n_list); has_null: list held a NULL
	PRED_IN_F64 = 3,  // column (as double) IN aux_f64[...]
	PRED_STR_LUT = 4, // string column: aux_u8[aux_off + dictionary code] = 1 true / 0 false / 2 null
	PRED_IS_NULL = 5,
	PRED_NOT_NULL = 6,
	PRED_CONST = 7, // ival: 1 true / 0 false / 2 null for every row (type mismatch, comparison with NULL)
	PRED_NOT = 8,
	PRED_AND = 9,
	PRED_OR = 10,
};
// a comparison as the set of orderings that satisfy it: bit 0 = less, bit 1 = equal, bit 2 = greater (branch-free on the device)
enum PredCmp : int32_t { CMP_LT = 1, CMP_EQ = 2, CMP_LE = 3, CMP_GT = 4, CMP_NE = 5, CMP_GE = 6 };

struct PredOp {
	int32_t op = PRED_CONST;
	int32_t cmp = CMP_EQ;
	int32_t col_is_f64 = 0; // PRED_CMP_F64 / PRED_IN_F64: storage of the column
	int32_t n_list = 0;
	int32_t has_null = 0;
	int32_t negate = 0; // PRED_IN_*: NOT IN
	int64_t aux_off = 0;
	int64_t ival = 0;
	double dval = 0;
	const void *data = nullptr;      // device: the column's values (int64 / double / int32 codes)
	const uint32_t *valid = nullptr; // device: validity bitmap of the column
};

static constexpr int PRED_MAX_OPS = 48;   // one kernel parameter block (48 x 64 B = 3 KB of the 4 KB limit)
static constexpr int PRED_MAX_DEPTH = 30; // stack bits per thread

struct PredProgram {
	PredOp ops[PRED_MAX_OPS];
	int n_ops = 0;
	const int64_t *aux_i64 = nullptr;
	const double *aux_f64 = nullptr;
	const uint8_t *aux_u8 = nullptr;
};

// out[w] bit b = predicate is TRUE for row 32 w + b (rows >= n: 0).  `words` = words written.
void launch_eval_predicate(const PredProgram &prog, long long n, uint32_t *out, long long words, cudaStream_t s,
                           int64_t *launches);
// dst[dst_first + i] = src[rows[i]] for 8-byte / 4-byte elements and validity bits (index merge carries the columns)
void launch_gather_column(const void *src, const uint32_t *src_valid, const int64_t *d_rows, long long m, int elem_bytes,
                          void *dst, uint32_t *dst_valid, long long dst_first, cudaStream_t s, int64_t *launches);

// A device-resident allow bitmap, kept alive by whoever searches with it
struct DevBitmap {
	void *p = nullptr;
	int64_t bits = 0;
	int64_t words = 0;
	~DevBitmap();
	const uint32_t *data() const { return static_cast<const uint32_t *>(p); }
};

class ColumnStore {
public:
	struct Column {
		std::string name;
		ColType type = COL_I64;
		void *data = nullptr;      // device [cap]: int64 (I64, BOOL) / double (F64) / int32 dictionary codes (STR)
		uint32_t *valid = nullptr; // device [cap / 32]; rows never written are NULL
		int64_t cap = 0;
		std::vector<std::string> dict; // STR: code -> text
		std::unordered_map<std::string, int32_t> dict_index;
		int elem_bytes() const { return type == COL_STR ? 4 : 8; }
	};

	~ColumnStore();
	bool empty() const { return cols_.empty(); }
	size_t size() const { return cols_.size(); }
	const Column &column(size_t i) const { return *cols_[i]; }
	int find(const std::string &name) const;
	// creates the column when it does not exist (type must match when it does)
	Column &define(const std::string &name, ColType type);
	// rows [first_row, first_row + count): values = int64[] / double[] / uint8[] (BOOL) / UTF-8 bytes with
	// offsets[count + 1] (STR); valid = one byte per row or NULL (all valid).  Synchronous on `s`.
	void write(const std::string &name, ColType type, int64_t first_row, int64_t count, const void *values,
	           const int64_t *offsets, const uint8_t *valid, int64_t cap_rows, cudaStream_t s);
	// Arrow C data interface flavour: validity as an Arrow bitmap, element offset `off` into every buffer
	void write_arrow(const std::string &name, ColType type, int64_t first_row, int64_t count, const void *values,
	                 int value_width, const void *offsets, bool large_offsets, const uint8_t *valid_bits, int64_t off,
	                 int64_t cap_rows, cudaStream_t s);
	void grow(int64_t cap_rows, cudaStream_t s); // the vector table was reallocated / a column was just defined
	void define_like(const ColumnStore &other); // same columns (names, types, dictionaries), no rows
	void swap(ColumnStore &other);              // contents and cache; the epoch moves forward on both
	void touch() { epoch_++; }                   // rows appended / merged / compacted
	uint64_t epoch() const { return epoch_; }

	// The allow bitmap of `predicate` over rows [0, n_rows), evaluated on `s` unless cached.  Throws Error on a
	// malformed predicate / unknown column (message as the host mirror's BinderException).
	std::shared_ptr<DevBitmap> bitmap(const std::string &predicate, int64_t n_rows, cudaStream_t s, int64_t *launches,
	                                  bool *was_cached = nullptr);

	// persistence (store.bin, behind the IVF_PQ section) and merge
	void save(FILE *f, int64_t n_rows, cudaStream_t s) const;
	void load(FILE *f, const std::string &path, int64_t n_rows, int64_t cap_rows, cudaStream_t s);
	// appends rows `d_rows` (device, m entries) of `src` at [dst_first, dst_first + m); columns are matched by name
	void gather_from(const ColumnStore &src, const int64_t *d_rows, int64_t m, int64_t dst_first, int64_t cap_rows,
	                 cudaStream_t s, int64_t *launches);

private:
	std::vector<std::unique_ptr<Column>> cols_;
	uint64_t epoch_ = 1;
	struct Cached {
		uint64_t epoch = 0;
		int64_t n_rows = 0;
		std::shared_ptr<DevBitmap> bm;
		uint64_t last_use = 0;
	};
	std::mutex cache_mu_;
	std::map<std::string, Cached> cache_;
	uint64_t use_clock_ = 0;
	void ensure_cap(Column &c, int64_t cap_rows, cudaStream_t s);
};

} // namespace lb
