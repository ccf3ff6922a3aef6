// store_columns.cu — the Store's side of the extra columns: typed host-buffer ingest, Arrow C data interface ingest
// (what src/lance_index.cpp:321-359 hands to lance_detached_add_batch_arrow), predicate search.
#include <cstring>

#include "store.h"

namespace lb {

// The Arrow C data interface (https://arrow.apache.org/docs/format/CDataInterface.html) — the two ABI-stable structs
// DuckDB's ArrowConverter fills (src/lance_index.cpp:343-349).
struct ArrowSchema {
	const char *format;
	const char *name;
	const char *metadata;
	int64_t flags;
	int64_t n_children;
	struct ArrowSchema **children;
	struct ArrowSchema *dictionary;
	void (*release)(struct ArrowSchema *);
	void *private_data;
};
struct ArrowArray {
	int64_t length;
	int64_t null_count;
	int64_t offset;
	int64_t n_buffers;
	int64_t n_children;
	const void **buffers;
	struct ArrowArray **children;
	struct ArrowArray *dictionary;
	void (*release)(struct ArrowArray *);
	void *private_data;
};

namespace {

struct ArrowColType {
	bool ok = false;
	ColType type = COL_I64;
	int width = 8;      // bytes per value, negative = unsigned
	bool large = false; // 64-bit string offsets
};

ArrowColType arrow_extra_type(const char *fmt) {
	ArrowColType t;
	if (!fmt) return t;
	const std::string f(fmt);
	t.ok = true;
	if (f == "c") t.type = COL_I64, t.width = 1;
	else if (f == "C") t.type = COL_I64, t.width = -1;
	else if (f == "s") t.type = COL_I64, t.width = 2;
	else if (f == "S") t.type = COL_I64, t.width = -2;
	else if (f == "i") t.type = COL_I64, t.width = 4;
	else if (f == "I") t.type = COL_I64, t.width = -4;
	else if (f == "l" || f == "L") t.type = COL_I64, t.width = 8;
	else if (f == "f") t.type = COL_F64, t.width = 4;
	else if (f == "g") t.type = COL_F64, t.width = 8;
	else if (f == "b") t.type = COL_BOOL, t.width = 0;
	else if (f == "u") t.type = COL_STR, t.large = false;
	else if (f == "U") t.type = COL_STR, t.large = true;
	else t.ok = false;
	return t;
}

// "+w:<D>" with a float child = the vector column
int arrow_vector_dim(const ArrowSchema *s) {
	if (!s || !s->format || strncmp(s->format, "+w:", 3) != 0) return 0;
	if (s->n_children != 1 || !s->children || !s->children[0] || !s->children[0]->format) return 0;
	if (strcmp(s->children[0]->format, "f") != 0) return 0;
	return atoi(s->format + 3);
}

int find_vector_child(const ArrowSchema *s) {
	if (!s || !s->format || strcmp(s->format, "+s") != 0) throw Error("arrow schema is not a struct (\"+s\")");
	for (int64_t i = 0; i < s->n_children; i++)
		if (arrow_vector_dim(s->children[i]) > 0) return (int)i;
	throw Error("no FixedSizeList column found in schema"); // lance_manager.rs:86-88
}

std::string bad_type(const ArrowSchema *c) {
	return std::string("unsupported arrow type \"") + (c->format ? c->format : "") + "\" for extra column \"" +
	       (c->name ? c->name : "") + "\"";
}

} // namespace

int arrow_schema_vector_dim(void *arrow_schema) {
	const ArrowSchema *s = static_cast<const ArrowSchema *>(arrow_schema);
	return arrow_vector_dim(s->children[find_vector_child(s)]);
}

bool Store::has_extra_columns() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	return !columns_.empty();
}

int Store::column_count() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	return (int)columns_.size();
}

bool Store::column_info(int i, std::string *name, int32_t *type) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	if (i < 0 || (size_t)i >= columns_.size()) return false;
	if (name) *name = columns_.column((size_t)i).name;
	if (type) *type = (int32_t)columns_.column((size_t)i).type;
	return true;
}

void Store::set_column(const std::string &name, ColType type, int64_t first_label, int64_t count, const void *values,
                       const int64_t *offsets, const uint8_t *valid) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	const int64_t first = count > 0 ? row_of_label(first_label) : 0;
	// the labels of the range must sit in consecutive rows (true for rows as they were appended)
	if (count < 0 || first < 0 || first + count > n_rows_ || (count > 0 && label_of_row(first + count - 1) != first_label + count - 1))
		throw Error("column values for labels [" + std::to_string(first_label) + ", " + std::to_string(first_label + count) +
		            ") fall outside the table");
	if (count == 0) {
		columns_.define(name, type);
		columns_.grow(cap_rows_, copy_stream_);
		return;
	}
	columns_.write(name, type, first, count, values, offsets, valid, cap_rows_, copy_stream_);
}

void Store::define_columns_from_arrow(void *arrow_schema) {
	const ArrowSchema *s = static_cast<const ArrowSchema *>(arrow_schema);
	const int vi = find_vector_child(s);
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	for (int64_t i = 0; i < s->n_children; i++) {
		if (i == vi) continue;
		const ArrowSchema *c = s->children[i];
		const ArrowColType t = arrow_extra_type(c->format);
		if (!t.ok) throw Error(bad_type(c));
		columns_.define(c->name ? c->name : "", t.type);
	}
	columns_.grow(cap_rows_, copy_stream_);
}

int64_t Store::add_arrow(void *arrow_schema, void *arrow_array, int64_t *out_count, int64_t first_label) {
	const ArrowSchema *s = static_cast<const ArrowSchema *>(arrow_schema);
	ArrowArray *a = static_cast<ArrowArray *>(arrow_array);
	struct Releaser { // the callee consumes the array (ffi.rs:148-182: "takes ownership of arrow_array")
		ArrowArray *a;
		~Releaser() {
			if (a && a->release) a->release(a);
		}
	} rel{a};
	const int vi = find_vector_child(s);
	if (a->n_children != s->n_children || !a->children) throw Error("arrow array does not match its schema");
	const int64_t n = a->length;
	const ArrowArray *va = a->children[vi];
	const int d = arrow_vector_dim(s->children[vi]);
	if (d != dim_) throw Error("dimension mismatch: expected " + std::to_string(dim_) + ", got " + std::to_string(d));
	if (!va || va->n_children != 1 || !va->children || !va->children[0] || va->children[0]->n_buffers < 2)
		throw Error("malformed FixedSizeList array");
	if (va->length < n) throw Error("vector column shorter than the batch");
	if (va->null_count > 0) throw Error("NULL vectors cannot be indexed");
	const ArrowArray *fa = va->children[0];
	const float *vec = static_cast<const float *>(fa->buffers[1]) + (size_t)fa->offset + (size_t)(a->offset + va->offset) * d;
	// validate the extra columns before anything is appended
	std::vector<ArrowColType> types((size_t)s->n_children);
	for (int64_t i = 0; i < s->n_children; i++) {
		if (i == vi) continue;
		types[(size_t)i] = arrow_extra_type(s->children[i]->format);
		const ArrowArray *c = a->children[i];
		if (!types[(size_t)i].ok) throw Error(bad_type(s->children[i]));
		if (!c || c->length < n || c->n_buffers < (types[(size_t)i].type == COL_STR ? 3 : 2)) throw Error("malformed arrow column");
	}
	int64_t first;
	if (first_label >= 0) {
		add_host_labeled(vec, n, first_label);
		first = first_label;
	} else {
		first = add_host(vec, n);
	}
	if (out_count) *out_count = n;
	if (n == 0) return first;
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	for (int64_t i = 0; i < s->n_children; i++) {
		if (i == vi) continue;
		const ArrowArray *c = a->children[i];
		const ArrowColType &t = types[(size_t)i];
		const int64_t off = a->offset + c->offset;
		const uint8_t *vbits = (c->null_count != 0 && c->buffers[0]) ? static_cast<const uint8_t *>(c->buffers[0]) : nullptr;
		const char *nm = s->children[i]->name ? s->children[i]->name : "";
		if (t.type == COL_STR)
			columns_.write_arrow(nm, t.type, row_of_label(first), n, c->buffers[2], 1, c->buffers[1], t.large, vbits, off, cap_rows_, copy_stream_);
		else
			columns_.write_arrow(nm, t.type, row_of_label(first), n, c->buffers[1], t.width, nullptr, false, vbits, off, cap_rows_, copy_stream_);
	}
	return first;
}

std::shared_ptr<DevBitmap> Store::predicate_bitmap(const std::string &predicate, bool *was_cached) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	int64_t l = 0;
	// evaluated on a pooled io stream: searches on other streams only ever see the finished bitmap
	auto io = acquire_io();
	std::shared_ptr<DevBitmap> bm;
	try {
		bm = columns_.bitmap(predicate, n_rows_, io->stream, &l, was_cached);
	} catch (...) {
		release_io(std::move(io));
		throw;
	}
	release_io(std::move(io));
	add_launches(l);
	return bm;
}

void Store::search_host_predicate(const float *queries, int nq, int k, int nprobes, int refine, const char *predicate,
                                  int64_t *out_labels, float *out_dist, int32_t *out_counts) {
	if (!predicate || !*predicate)
		return search_host(queries, nq, k, nprobes, refine, nullptr, 0, out_labels, out_dist, out_counts);
	std::shared_ptr<DevBitmap> bm = predicate_bitmap(predicate);
	search_host(queries, nq, k, nprobes, refine, nullptr, 0, out_labels, out_dist, out_counts, bm.get());
}

} // namespace lb
