// ffi.cu — the C boundary (include/lance_b200.h).  Same symbols, argument meaning and error
// convention as the Rust crate's exports (rust_lib/src/ffi.rs:37-541) that src/rust_ffi.cpp:9-41 binds.
// Nothing throws across this file: every entry catches and writes err_buf.
#include <cstdio>
#include <cstring>
#include <memory>
#include <new>

#include "sharded.h"
#include "store.h"

using lb::Store;

static void write_err(char *buf, int len, const std::string &msg) {
	// rust_lib/src/ffi.rs:15-24
	if (!buf || len <= 0) return;
	size_t n = msg.size() < (size_t)(len - 1) ? msg.size() : (size_t)(len - 1);
	memcpy(buf, msg.data(), n);
	buf[n] = '\0';
}
static std::string cstr(const char *s) { return s ? std::string(s) : std::string(); }

#define LB_TRY(prefix, fail_value, body)                         \
	try {                                                        \
		body                                                     \
	} catch (const std::exception &e) {                          \
		write_err(err_buf, err_buf_len, std::string(prefix) + e.what()); \
		return fail_value;                                       \
	} catch (...) {                                              \
		write_err(err_buf, err_buf_len, std::string(prefix) + "unknown error"); \
		return fail_value;                                       \
	}

#define LB_HANDLE(fail_value)                               \
	if (!handle) {                                          \
		write_err(err_buf, err_buf_len, "null handle");     \
		return fail_value;                                  \
	}                                                       \
	if (lb::as_sharded(handle)) {                           \
		write_err(err_buf, err_buf_len, "not supported on a sharded handle"); \
		return fail_value;                                  \
	}                                                       \
	Store *h = static_cast<Store *>(handle);

extern "C" {

// LANCE_B200_DEVICES="0,1,2,3" (or "all"): lance_create_detached[_from_arrow] spread the table over those devices — the
// sharded handle of sharded.h behind the reference's own entry points, so the unmodified extension reaches every GPU.
// Unset or one device: a plain one-device store.
static std::vector<int> env_devices() {
	std::vector<int> dv;
	const char *e = getenv("LANCE_B200_DEVICES");
	if (!e || !*e) return dv;
	if (std::string(e) == "all") {
		int count = 0;
		if (cudaGetDeviceCount(&count) != cudaSuccess) count = 0;
		for (int i = 0; i < count; i++) dv.push_back(i);
		return dv;
	}
	const char *p = e;
	while (*p) {
		char *end = nullptr;
		const long v = strtol(p, &end, 10);
		if (end == p) throw lb::Error(std::string("LANCE_B200_DEVICES: cannot parse \"") + e + "\"");
		dv.push_back((int)v);
		p = end;
		while (*p == ',' || *p == ' ') p++;
	}
	return dv;
}

void *lance_create_detached(const char *db_path, int32_t dimension, const char *metric, const char *table_name,
                            char *err_buf, int err_buf_len) {
	LB_TRY("create failed: ", nullptr, {
		const std::vector<int> dv = env_devices();
		if (dv.size() > 1) {
			std::unique_ptr<lb::ShardedStore> sh(new lb::ShardedStore(dv, dimension, cstr(metric), cstr(db_path), cstr(table_name)));
			lb::register_sharded(sh.get());
			return static_cast<void *>(sh.release());
		}
		lb::ShardedStore::forget(cstr(db_path), cstr(table_name)); // a new plain table replaces a sharded one of that name
		return static_cast<void *>(new Store(cstr(db_path), cstr(table_name), dimension, cstr(metric),
		                                     dv.size() == 1 ? dv[0] : lb::default_device()));
	})
}

void *lance_create_detached_from_arrow(const char *db_path, void *arrow_schema, const char *metric,
                                       const char *table_name, char *err_buf, int err_buf_len) {
	// rust_lib/src/ffi.rs:63-90, lance_manager.rs:67-135: the vector dimension comes from the FixedSizeList column,
	// the other fields become extra columns (on the device here: columns.h); the schema is borrowed, not consumed
	if (!arrow_schema) {
		write_err(err_buf, err_buf_len, "null arrow schema");
		return nullptr;
	}
	LB_TRY("create_from_arrow failed: ", nullptr, {
		const int dim = lb::arrow_schema_vector_dim(arrow_schema);
		const std::vector<int> dv = env_devices();
		if (dv.size() > 1) {
			std::unique_ptr<lb::ShardedStore> sh(new lb::ShardedStore(dv, dim, cstr(metric), cstr(db_path), cstr(table_name)));
			sh->define_columns_from_arrow(arrow_schema);
			lb::register_sharded(sh.get());
			return static_cast<void *>(sh.release());
		}
		lb::ShardedStore::forget(cstr(db_path), cstr(table_name));
		std::unique_ptr<Store> st(new Store(cstr(db_path), cstr(table_name), dim, cstr(metric),
		                                    dv.size() == 1 ? dv[0] : lb::default_device()));
		st->define_columns_from_arrow(arrow_schema);
		return static_cast<void *>(st.release());
	})
}

void *lance_open_detached(const char *db_path, const char *table_name, const char *metric, char *err_buf,
                          int err_buf_len) {
	LB_TRY("open failed: ", nullptr, {
		// a table that was saved sharded reopens sharded: on LANCE_B200_DEVICES when it names as many devices as there are
		// shards, else shard i on device i modulo the visible ones
		const int G = lb::ShardedStore::saved_shards(cstr(db_path), cstr(table_name));
		std::vector<int> dv = env_devices();
		if (G > 0) {
			if ((int)dv.size() != G) {
				int count = 1;
				if (cudaGetDeviceCount(&count) != cudaSuccess || count < 1) count = 1;
				dv.clear();
				for (int i = 0; i < G; i++) dv.push_back(i % count);
			}
			std::unique_ptr<lb::ShardedStore> sh(lb::ShardedStore::open(cstr(db_path), cstr(table_name), cstr(metric), dv));
			lb::register_sharded(sh.get());
			return static_cast<void *>(sh.release());
		}
		return static_cast<void *>(Store::open(cstr(db_path), cstr(table_name), cstr(metric),
		                                       dv.size() == 1 ? dv[0] : lb::default_device()));
	})
}

void lance_free_detached(void *handle) {
	if (!handle) return;
	try {
		if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
			lb::unregister_sharded(sh);
			delete sh;
			return;
		}
		delete static_cast<Store *>(handle);
	} catch (...) {
	}
}

int32_t lance_detached_has_extra_columns(void *handle) {
	// rust_lib/src/ffi.rs:115-123
	if (!handle) return 0;
	try {
		if (lb::ShardedStore *sh = lb::as_sharded(handle)) return sh->has_extra_columns() ? 1 : 0;
		return static_cast<Store *>(handle)->has_extra_columns() ? 1 : 0;
	} catch (...) {
		return 0;
	}
}

int32_t lance_detached_dimension(void *handle) {
	if (!handle) return 0;
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) return sh->dim();
	return static_cast<Store *>(handle)->dim();
}

int64_t lance_detached_add(void *handle, const float *vector, int32_t dimension, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("add failed: ", -1, {
			if (dimension != sh->dim())
				throw lb::Error("expected dimension " + std::to_string(sh->dim()) + ", got " + std::to_string(dimension));
			return sh->add_host(vector, 1);
		})
	}

	LB_HANDLE(-1)
	LB_TRY("add failed: ", -1, {
		if (dimension != h->dim())
			throw lb::Error("expected dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dimension));
		return h->add_host(vector, 1);
	})
}

int32_t lance_detached_add_batch(void *handle, const float *vectors, int32_t num, int32_t dim, int64_t *out_labels,
                                 char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("add_batch failed: ", -1, {
			if (dim != sh->dim())
				throw lb::Error("expected dimension " + std::to_string(sh->dim()) + ", got " + std::to_string(dim));
			if (num < 0) throw lb::Error("negative batch size");
			const int64_t first = sh->add_host(vectors, num);
			if (out_labels)
				for (int32_t i = 0; i < num; i++) out_labels[i] = first + i;
			return num;
		})
	}

	LB_HANDLE(-1)
	LB_TRY("add_batch failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("expected dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		if (num < 0) throw lb::Error("negative batch size");
		int64_t first = h->add_host(vectors, num);
		if (out_labels)
			for (int32_t i = 0; i < num; i++) out_labels[i] = first + i;
		return num;
	})
}

int32_t lance_detached_add_batch_arrow(void *handle, void *arrow_schema, void *arrow_array, int64_t *out_labels,
                                       char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("add_batch_arrow failed: ", -1, {
			if (!arrow_schema || !arrow_array) throw lb::Error("null arrow schema/array");
			int64_t n = 0;
			const int64_t first = sh->add_arrow(arrow_schema, arrow_array, &n);
			if (out_labels)
				for (int64_t i = 0; i < n; i++) out_labels[i] = first + i;
			return (int32_t)n;
		})
	}

	// rust_lib/src/ffi.rs:148-182: consumes the array, borrows the schema, labels are consecutive
	LB_HANDLE(-1)
	if (!arrow_schema || !arrow_array) {
		write_err(err_buf, err_buf_len, "null arrow schema/array");
		return -1;
	}
	LB_TRY("add_batch_arrow failed: ", -1, {
		int64_t n = 0;
		const int64_t first = h->add_arrow(arrow_schema, arrow_array, &n);
		if (out_labels)
			for (int64_t i = 0; i < n; i++) out_labels[i] = first + i;
		return (int32_t)n;
	})
}

int32_t lance_detached_search_batch_predicate(void *handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                              int32_t nprobes, int32_t refine_factor, const char *predicate,
                                              int64_t *out_labels, float *out_distances, int32_t *out_counts,
                                              char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("search failed: ", -1, {
			if (dim != sh->dim())
				throw lb::Error("expected query dimension " + std::to_string(sh->dim()) + ", got " + std::to_string(dim));
			if (nq < 0 || k < 0) throw lb::Error("negative batch size or k");
			if (nq == 0) return 0;
			if (!queries || !out_labels || !out_distances || !out_counts) throw lb::Error("null buffer");
			sh->search_host(queries, nq, k, nprobes, refine_factor, nullptr, 0, predicate, out_labels, out_distances, out_counts);
			return nq;
		})
	}

	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("expected query dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		if (nq < 0 || k < 0) throw lb::Error("negative batch size or k");
		if (nq == 0) return 0;
		if (!queries || !out_labels || !out_distances || !out_counts) throw lb::Error("null buffer");
		h->search_host_predicate(queries, nq, k, nprobes, refine_factor, predicate, out_labels, out_distances, out_counts);
		return nq;
	})
}

int32_t lance_detached_search_predicate(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                        int32_t refine_factor, const char *predicate, int64_t *out_labels,
                                        float *out_distances, char *err_buf, int err_buf_len) {
	int32_t count = 0;
	int32_t rc = lance_detached_search_batch_predicate(handle, query, 1, dim, k, nprobes, refine_factor, predicate,
	                                                   out_labels, out_distances, &count, err_buf, err_buf_len);
	return rc < 0 ? -1 : count;
}

int32_t lance_detached_create_scalar_index(void *handle, const char *column, const char *index_type, char *err_buf,
                                           int err_buf_len) {
	// src/lance_index.cpp:481-486 (newer revision): a scalar index only speeds up Lance's prefilter; here the
	// predicate kernel scans the column at HBM rate and its bitmap is cached, so there is nothing to build —
	// the call validates its arguments and succeeds
	LB_HANDLE(-1)
	LB_TRY("create_scalar_index failed: ", -1, {
		(void)index_type;
		const std::string name = cstr(column);
		for (int i = 0; i < h->column_count(); i++) {
			std::string n;
			if (h->column_info(i, &n, nullptr) && n == name) return 0;
		}
		throw lb::Error("no extra column named \"" + name + "\"");
	})
}

void *lance_b200_create_sharded(const int32_t *devices, int32_t n_devices, int32_t dimension, const char *metric,
                                char *err_buf, int err_buf_len) {
	LB_TRY("create_sharded failed: ", nullptr, {
		if (!devices || n_devices <= 0) throw lb::Error("no devices given");
		std::vector<int> dv(devices, devices + n_devices);
		std::unique_ptr<lb::ShardedStore> sh(new lb::ShardedStore(dv, dimension, cstr(metric)));
		lb::register_sharded(sh.get());
		return sh.release();
	})
}

void *lance_b200_create_sharded_at(const char *db_path, const char *table_name, const int32_t *devices, int32_t n_devices,
                                   int32_t dimension, const char *metric, char *err_buf, int err_buf_len) {
	LB_TRY("create_sharded failed: ", nullptr, {
		if (!devices || n_devices <= 0) throw lb::Error("no devices given");
		std::vector<int> dv(devices, devices + n_devices);
		std::unique_ptr<lb::ShardedStore> sh(new lb::ShardedStore(dv, dimension, cstr(metric), cstr(db_path), cstr(table_name)));
		lb::register_sharded(sh.get());
		return sh.release();
	})
}

void *lance_b200_open_sharded(const char *db_path, const char *table_name, const char *metric, const int32_t *devices,
                              int32_t n_devices, char *err_buf, int err_buf_len) {
	LB_TRY("open_sharded failed: ", nullptr, {
		if (!devices || n_devices <= 0) throw lb::Error("no devices given");
		std::vector<int> dv(devices, devices + n_devices);
		std::unique_ptr<lb::ShardedStore> sh(lb::ShardedStore::open(cstr(db_path), cstr(table_name), cstr(metric), dv));
		lb::register_sharded(sh.get());
		return sh.release();
	})
}

int32_t lance_b200_sharded_rows(void *handle, int64_t *out_rows, int32_t capacity) {
	lb::ShardedStore *sh = lb::as_sharded(handle);
	if (!sh) return -1;
	try {
		for (int i = 0; i < sh->num_shards() && i < capacity; i++) out_rows[i] = sh->shard(i).rows();
		return sh->num_shards();
	} catch (...) {
		return -1;
	}
}

int32_t lance_b200_add_batch_labeled(void *handle, const float *vectors, int64_t num, int32_t dim, int64_t first_label,
                                     char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("add_batch failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("dimension mismatch: expected " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		if (num < 0) throw lb::Error("negative count");
		if (num > 0 && !vectors) throw lb::Error("null buffer");
		h->add_host_labeled(vectors, num, first_label);
		return 0;
	})
}

int32_t lance_b200_add_batch_device_labels(void *handle, const float *d_vectors, int64_t num, int32_t dim,
                                           const int64_t *labels, void *stream, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("add_batch failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("dimension mismatch: expected " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		if (num < 0) throw lb::Error("negative count");
		if (num > 0 && (!d_vectors || !labels)) throw lb::Error("null buffer");
		h->add_device_with_labels(d_vectors, num, labels, static_cast<cudaStream_t>(stream));
		return 0;
	})
}

int32_t lance_b200_set_column(void *handle, const char *name, int32_t type, int64_t first_label, int64_t count,
                              const void *values, const int64_t *offsets, const uint8_t *valid, char *err_buf,
                              int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("set_column failed: ", -1, {
			if (type < 0 || type > lb::COL_STR) throw lb::Error("unknown column type");
			sh->set_column(cstr(name), (lb::ColType)type, first_label, count, values, offsets, valid);
			return 0;
		})
	}

	LB_HANDLE(-1)
	LB_TRY("set_column failed: ", -1, {
		if (type < 0 || type > lb::COL_STR) throw lb::Error("unknown column type");
		h->set_column(cstr(name), (lb::ColType)type, first_label, count, values, offsets, valid);
		return 0;
	})
}

int64_t lance_b200_predicate_bitmap(void *handle, const char *predicate, uint32_t *out_bitmap, int64_t out_words,
                                    const uint32_t **out_device_bitmap, int32_t *out_was_cached, char *err_buf,
                                    int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("predicate failed: ", -1, {
		bool cached = false;
		std::shared_ptr<lb::DevBitmap> bm = h->predicate_bitmap(cstr(predicate), &cached);
		if (out_was_cached) *out_was_cached = cached ? 1 : 0;
		if (out_device_bitmap) *out_device_bitmap = bm->data(); // stays valid while the predicate is cached
		if (out_bitmap && out_words > 0) {
			const int64_t w = bm->words < out_words ? bm->words : out_words;
			memset(out_bitmap, 0, (size_t)out_words * sizeof(uint32_t));
			if (w > 0) LB_CUDA(cudaMemcpy(out_bitmap, bm->data(), (size_t)w * sizeof(uint32_t), cudaMemcpyDeviceToHost));
		}
		return bm->bits;
	})
}

int32_t lance_detached_merge(void *target_handle, void *source_handle, const int64_t *live_source_labels,
                             int32_t live_count, int64_t *out_old_labels, int64_t *out_new_labels, char *err_buf,
                             int err_buf_len) {
	// rust_lib/src/ffi.rs:187-228
	if (!target_handle || !source_handle) {
		write_err(err_buf, err_buf_len, "null handle");
		return -1;
	}
	if (lb::as_sharded(target_handle) || lb::as_sharded(source_handle)) {
		// a sharded side: the listed live rows of the source travel through the host (get_vector / add), in the order of
		// the list as it stands in the source; same contract, no device-to-device gather
		LB_TRY("merge failed: ", -1, {
			if (target_handle == source_handle) throw lb::Error("cannot merge an index into itself");
			lb::ShardedStore *ts = lb::as_sharded(target_handle);
			lb::ShardedStore *ss = lb::as_sharded(source_handle);
			const int dim_t = ts ? ts->dim() : static_cast<Store *>(target_handle)->dim();
			const int dim_s = ss ? ss->dim() : static_cast<Store *>(source_handle)->dim();
			if (dim_t != dim_s)
				throw lb::Error("dimension mismatch: expected " + std::to_string(dim_t) + ", got " + std::to_string(dim_s));
			if (live_count < 0) throw lb::Error("negative count");
			std::vector<int64_t> sorted(live_source_labels, live_source_labels + live_count);
			std::sort(sorted.begin(), sorted.end());
			sorted.erase(std::unique(sorted.begin(), sorted.end()), sorted.end());
			std::vector<float> rows;
			std::vector<int64_t> olds;
			std::vector<float> one((size_t)dim_s);
			for (int64_t lab : sorted) {
				try {
					if (ss) ss->get_vector(lab, one.data());
					else static_cast<Store *>(source_handle)->get_vector(lab, one.data());
				} catch (const lb::Error &) {
					continue; // unknown or deleted in the source: not merged
				}
				rows.insert(rows.end(), one.begin(), one.end());
				olds.push_back(lab);
			}
			const int64_t m = (int64_t)olds.size();
			if (m == 0) return 0;
			int64_t first;
			if (ts) first = ts->add_host(rows.data(), m);
			else first = static_cast<Store *>(target_handle)->add_host(rows.data(), m);
			for (int64_t i = 0; i < m; i++) {
				if (out_old_labels) out_old_labels[i] = olds[(size_t)i];
				if (out_new_labels) out_new_labels[i] = first + i;
			}
			return (int)m;
		})
	}
	void *handle = target_handle;
	LB_HANDLE(-1)
	LB_TRY("merge failed: ", -1, {
		Store *src = static_cast<Store *>(source_handle);
		return h->merge_from(*src, live_source_labels, live_count, out_old_labels, out_new_labels);
	})
}

int32_t lance_detached_search_batch(void *handle, const float *queries, int32_t nq, int32_t dim, int32_t k,
                                    int32_t nprobes, int32_t refine_factor, const uint32_t *allow_bitmap,
                                    int64_t bitmap_bits, int64_t *out_labels, float *out_distances,
                                    int32_t *out_counts, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("search failed: ", -1, {
			if (dim != sh->dim())
				throw lb::Error("expected query dimension " + std::to_string(sh->dim()) + ", got " + std::to_string(dim));
			if (nq < 0 || k < 0) throw lb::Error("negative batch size or k");
			if (nq == 0) return 0;
			if (!queries || !out_labels || !out_distances || !out_counts) throw lb::Error("null buffer");
			sh->search_host(queries, nq, k, nprobes, refine_factor, allow_bitmap, bitmap_bits, nullptr, out_labels, out_distances,
			                out_counts);
			return nq;
		})
	}

	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, {
		// rust_lib/src/lance_manager.rs:401-407
		if (dim != h->dim())
			throw lb::Error("expected query dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		if (nq < 0 || k < 0) throw lb::Error("negative batch size or k");
		if (nq == 0) return 0;
		if (!queries || !out_labels || !out_distances || !out_counts) throw lb::Error("null buffer");
		h->search_host(queries, nq, k, nprobes, refine_factor, allow_bitmap, bitmap_bits, out_labels, out_distances,
		               out_counts);
		return nq;
	})
}

int32_t lance_detached_search_filtered(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                       int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                                       int64_t *out_labels, float *out_distances, char *err_buf, int err_buf_len) {
	int32_t count = 0;
	int32_t rc = lance_detached_search_batch(handle, query, 1, dim, k, nprobes, refine_factor, allow_bitmap,
	                                         bitmap_bits, out_labels, out_distances, &count, err_buf, err_buf_len);
	return rc < 0 ? -1 : count;
}

int32_t lance_detached_search(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                              int32_t refine_factor, int64_t *out_labels, float *out_distances, char *err_buf,
                              int err_buf_len) {
	return lance_detached_search_filtered(handle, query, dim, k, nprobes, refine_factor, nullptr, 0, out_labels,
	                                      out_distances, err_buf, err_buf_len);
}

int64_t lance_detached_count(void *handle, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("count failed: ", -1, {
			return sh->live_rows();
		})
	}

	LB_HANDLE(-1)
	LB_TRY("count failed: ", -1, { return h->live_rows(); })
}

int32_t lance_detached_delete_batch(void *handle, const int64_t *labels, int32_t count, char *err_buf,
                                    int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("delete failed: ", -1, {
			if (count > 0 && !labels) throw lb::Error("null buffer");
			sh->delete_labels(labels, count);
			return 0;
		})
	}

	LB_HANDLE(-1)
	LB_TRY("delete_batch failed: ", -1, {
		h->delete_labels(labels, count);
		return 0;
	})
}

int32_t lance_detached_delete(void *handle, int64_t label, char *err_buf, int err_buf_len) {
	if (lb::as_sharded(handle)) return lance_detached_delete_batch(handle, &label, 1, err_buf, err_buf_len);
	LB_HANDLE(-1)
	LB_TRY("delete failed: ", -1, {
		h->delete_labels(&label, 1);
		return 0;
	})
}

int32_t lance_detached_create_index(void *handle, int32_t num_partitions, int32_t num_sub_vectors, char *err_buf,
                                    int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("create_index failed: ", -1, {
			sh->create_index(num_partitions, num_sub_vectors); // one index per shard (sharded.h)
			return 0;
		})
	}
	LB_HANDLE(-1)
	LB_TRY("create_index failed: ", -1, {
		h->create_index(num_partitions, num_sub_vectors);
		return 0;
	})
}

int32_t lance_detached_create_hnsw_index(void *handle, int32_t m, int32_t ef_construction, char *err_buf, int err_buf_len) {
	// rust_lib/src/ffi.rs:425-446 -> lance_manager.rs:521-551: IVF_HNSW_SQ, `.replace(true)` (it takes the place of whatever
	// vector index the table had).  No graph is built here: pointer chasing is what this machine is worst at, and the
	// HBM-resident table answers an HNSW-sized request exactly at memory speed (4.1 / 4.2), i.e. with recall 1.0 >= any
	// graph's.  The call is accepted (m, ef_construction <= 0 = defaults, as in the reference), an existing IVF_PQ index
	// is dropped as a replaced index would be, and searches are exact from then on.
	(void)m, (void)ef_construction;
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("create_hnsw_index failed: ", -1, {
			sh->drop_index();
			return 0;
		})
	}
	LB_HANDLE(-1)
	LB_TRY("create_hnsw_index failed: ", -1, {
		h->drop_index();
		return 0;
	})
}

int32_t lance_detached_compact(void *handle, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("compact failed: ", -1, {
			sh->compact();
			return 0;
		})
	}

	// rust_lib/src/ffi.rs:448-465 -> lance_manager.rs:553-566 (compact_files): tombstoned rows are dropped for good, the
	// HBM they held is released, labels are kept
	LB_HANDLE(-1)
	LB_TRY("compact failed: ", -1, {
		h->compact();
		return 0;
	})
}

int32_t lance_detached_get_vector(void *handle, int64_t label, float *out_vec, int32_t capacity, char *err_buf,
                                  int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("get_vector failed: ", -1, {
			if (capacity < sh->dim()) throw lb::Error("output buffer too small");
			sh->get_vector(label, out_vec);
			return sh->dim();
		})
	}
	LB_HANDLE(-1)
	if (capacity < h->dim()) {
		write_err(err_buf, err_buf_len, "output buffer too small");
		return -1;
	}
	LB_TRY("get_vector failed: ", -1, {
		h->get_vector(label, out_vec);
		return h->dim();
	})
}

int32_t lance_detached_get_all_vectors(void *handle, int64_t *out_labels, float *out_vectors, int64_t *out_count,
                                       char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("get_all_vectors failed: ", -1, {
			const int64_t c = sh->get_all(out_labels, out_vectors);
			if (out_count) *out_count = c;
			return 0;
		})
	}
	LB_HANDLE(-1)
	LB_TRY("get_all_vectors failed: ", -1, {
		int64_t c = h->get_all(out_labels, out_vectors);
		if (out_count) *out_count = c;
		return 0;
	})
}

// ------------------------------------------------------------------ Part 2

int32_t lance_b200_search_device(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                 int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                 int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts, void *stream,
                                 char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, {
		h->search_device(d_queries, nq, k, nprobes, refine_factor, d_allow, bitmap_bits, d_out_labels, d_out_distances,
		                 d_out_counts, static_cast<cudaStream_t>(stream));
		return 0;
	})
}

int32_t lance_b200_search_device_async(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                       int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                       int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts,
                                       uint32_t *d_overflow, void *stream, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, {
		h->search_device(d_queries, nq, k, nprobes, refine_factor, d_allow, bitmap_bits, d_out_labels, d_out_distances,
		                 d_out_counts, static_cast<cudaStream_t>(stream), true, d_overflow);
		return 0;
	})
}

int32_t lance_b200_search_finish(void *handle, void *stream, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, { return h->search_finish(static_cast<cudaStream_t>(stream)); })
}

int32_t lance_b200_merge_topk_strided_device(int32_t device, const int64_t *d_labels, const float *d_dists,
                                             const int32_t *d_counts, int64_t labels_stride, int64_t dists_stride,
                                             int64_t counts_stride, int32_t G, int32_t nq, int32_t k,
                                             int64_t *d_out_labels, float *d_out_dists, int32_t *d_out_counts,
                                             void *stream, char *err_buf, int err_buf_len) {
	LB_TRY("merge failed: ", -1, {
		if (G <= 0 || nq < 0 || k < 0) throw lb::Error("bad shape");
		if (nq == 0 || k == 0) return 0;
		LB_CUDA(cudaSetDevice(device));
		lb::launch_merge_lists(d_labels, d_dists, d_counts, labels_stride, dists_stride, counts_stride, G, nq, k, d_out_labels,
		                       d_out_dists, d_out_counts, static_cast<cudaStream_t>(stream), nullptr);
		return 0;
	})
}

int32_t lance_b200_merge_topk_device(int32_t device, const int64_t *d_labels, const float *d_dists,
                                     const int32_t *d_counts, int32_t G, int32_t nq, int32_t k,
                                     int64_t *d_out_labels, float *d_out_dists, int32_t *d_out_counts, void *stream,
                                     char *err_buf, int err_buf_len) {
	return lance_b200_merge_topk_strided_device(device, d_labels, d_dists, d_counts, (int64_t)nq * k, (int64_t)nq * k, nq, G, nq,
	                                            k, d_out_labels, d_out_dists, d_out_counts, stream, err_buf, err_buf_len);
}

int32_t lance_b200_set_label_base(void *handle, int64_t base, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("set_label_base failed: ", -1, {
		h->set_label_base(base);
		return 0;
	})
}

int32_t lance_b200_reserve(void *handle, int64_t rows, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("reserve failed: ", -1, {
		h->reserve(rows);
		return 0;
	})
}

int64_t lance_b200_add_batch_device(void *handle, const float *d_vectors, int64_t num, int32_t dim, void *stream,
                                    char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("add_batch failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("expected dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		return h->add_device(d_vectors, num, static_cast<cudaStream_t>(stream));
	})
}

int32_t lance_b200_set_ivfpq(void *handle, int32_t P, int32_t M, int64_t n_indexed, const float *centroids,
                             const float *codebook, const int64_t *part_offsets, const int64_t *row_ids,
                             const uint8_t *codes, char *err_buf, int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("set_ivfpq failed: ", -1, {
		h->set_index(P, M, n_indexed, centroids, codebook, part_offsets, row_ids, codes);
		return 0;
	})
}

int32_t lance_b200_get_ivfpq(void *handle, lance_b200_ivfpq_info *info, float *centroids, float *codebook,
                             int64_t *part_offsets, int64_t *row_ids, uint8_t *codes, char *err_buf,
                             int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("get_ivfpq failed: ", -1, {
		h->get_index(info, centroids, codebook, part_offsets, row_ids, codes);
		return 0;
	})
}

int32_t lance_b200_drop_index(void *handle, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("drop_index failed: ", -1, {
			sh->drop_index();
			return 0;
		})
	}
	LB_HANDLE(-1)
	LB_TRY("drop_index failed: ", -1, {
		h->drop_index();
		return 0;
	})
}

int32_t lance_b200_ivfpq_search_debug(void *handle, const float *query, int32_t dim, int32_t k, int32_t nprobes,
                                      int32_t refine_factor, const uint32_t *allow_bitmap, int64_t bitmap_bits,
                                      int32_t *out_parts, int64_t *out_cand_labels, float *out_cand_adc,
                                      int32_t *out_ncand, int64_t *out_labels, float *out_distances, char *err_buf,
                                      int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("search failed: ", -1, {
		if (dim != h->dim())
			throw lb::Error("expected query dimension " + std::to_string(h->dim()) + ", got " + std::to_string(dim));
		int32_t count = 0;
		h->ivfpq_search_debug(query, k, nprobes, refine_factor, allow_bitmap, bitmap_bits, out_parts, out_cand_labels,
		                      out_cand_adc, out_ncand, out_labels, out_distances, &count);
		return count;
	})
}

int32_t lance_b200_save(void *handle, char *err_buf, int err_buf_len) {
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) {
		LB_TRY("save failed: ", -1, {
			sh->save();
			return 0;
		})
	}
	LB_HANDLE(-1)
	LB_TRY("save failed: ", -1, {
		h->save();
		return 0;
	})
}

int32_t lance_b200_stats(void *handle, lance_b200_stats_t *out) {
	if (!handle || !out) return -1;
	try {
		if (lb::ShardedStore *sh = lb::as_sharded(handle)) sh->stats(out);
		else static_cast<Store *>(handle)->stats(out);
		return 0;
	} catch (...) {
		return -1;
	}
}

int32_t lance_b200_profile_enable(void *handle, int32_t on) {
	if (!handle || lb::as_sharded(handle)) return -1; // per-kernel timers belong to one device store
	static_cast<Store *>(handle)->set_profile(on != 0);
	return 0;
}

int32_t lance_b200_profile_read(void *handle, lance_b200_profile_t *out) {
	if (!handle || !out || lb::as_sharded(handle)) return -1;
	try {
		static_cast<Store *>(handle)->read_profile(out);
		return 0;
	} catch (...) {
		return -1;
	}
}

int32_t lance_b200_time_search_device(void *handle, const float *d_queries, int32_t nq, int32_t k, int32_t nprobes,
                                      int32_t refine_factor, const uint32_t *d_allow, int64_t bitmap_bits,
                                      int64_t *d_out_labels, float *d_out_distances, int32_t *d_out_counts, void *stream,
                                      int32_t iters, void *d_flush, int64_t flush_bytes, float *out_ms, char *err_buf,
                                      int err_buf_len) {
	LB_HANDLE(-1)
	LB_TRY("timed search failed: ", -1, {
		if (iters <= 0 || !out_ms) throw lb::Error("bad arguments");
		cudaStream_t s = static_cast<cudaStream_t>(stream);
		LB_CUDA(cudaSetDevice(h->device()));
		cudaEvent_t a;
		cudaEvent_t b;
		LB_CUDA(cudaEventCreate(&a));
		LB_CUDA(cudaEventCreate(&b));
		const size_t half = d_flush ? ((size_t)flush_bytes / 2) & ~(size_t)255 : 0;
		for (int i = 0; i < iters; i++) {
			if (half) {
				// evict: write half the buffer (more than the L2 holds), then read the other half so that no dirty line is
				// left for the timed search to write back; all of it outside the timed interval
				LB_CUDA(cudaMemsetAsync(d_flush, i & 0xFF, half, s));
				lb::launch_read_pass(static_cast<char *>(d_flush) + half, half, reinterpret_cast<unsigned *>(d_flush), s);
				LB_CUDA(cudaStreamSynchronize(s));
			}
			LB_CUDA(cudaEventRecord(a, s));
			h->search_device(d_queries, nq, k, nprobes, refine_factor, d_allow, bitmap_bits, d_out_labels, d_out_distances,
			                 d_out_counts, s);
			LB_CUDA(cudaEventRecord(b, s));
			LB_CUDA(cudaEventSynchronize(b));
			LB_CUDA(cudaEventElapsedTime(&out_ms[i], a, b));
		}
		cudaEventDestroy(a);
		cudaEventDestroy(b);
		return 0;
	})
}

int32_t lance_b200_set_flat_path(void *handle, int32_t mode) {
	if (!handle || mode < 0 || mode > 2) return -1;
	if (lb::ShardedStore *sh = lb::as_sharded(handle)) sh->set_flat_path(mode);
	else static_cast<Store *>(handle)->set_flat_path(mode);
	return 0;
}

int32_t lance_b200_set_default_device(int32_t device) {
	lb::set_default_device(device);
	return 0;
}

int32_t lance_b200_device_count(void) {
	int c = 0;
	if (cudaGetDeviceCount(&c) != cudaSuccess) {
		(void)cudaGetLastError();
		return 0;
	}
	return c;
}

const char *lance_b200_version(void) { return "lance_b200 0.1 (sm_100a)"; }

} // extern "C"
