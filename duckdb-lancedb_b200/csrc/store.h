// store.h — the device-resident index behind one `lance_*` handle.
//
// Takes the place of rust_lib/src/lance_manager.rs's `LanceIndex` (fields :20-31, search :394-451,
// add_batch :227-242, delete_batch :461-471, count :473-478, create_ann_index :483-515) for the kNN
// path: vectors live in HBM as row-major fp32 [rows][dim] (the layout they arrive in from
// src/lance_index.cpp:361-367 / :941-946), labels are consecutive from an atomic counter so
// label == label_base + row, deletes are tombstone bits applied as a prefilter.
#pragma once
#include <atomic>
#include <cstdio>
#include <memory>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <utility>
#include <vector>

#include "columns.h"
#include "kernels.h"
#include "../../include/lance_b200.h"

namespace lb {

struct DevBuf {
	void *p = nullptr;
	size_t bytes = 0;
	DevBuf() = default;
	DevBuf(const DevBuf &) = delete;
	DevBuf &operator=(const DevBuf &) = delete;
	~DevBuf() { release(); }
	void release();
	void alloc(size_t n);          // exact size, contents undefined
	void ensure(size_t n);         // grow (contents lost) if smaller than n
	template <class T>
	T *as() const { return reinterpret_cast<T *>(p); }
};

// per-call scratch: its own stream and buffers, so concurrent searches on one handle do not share state
struct Workspace {
	cudaStream_t stream = nullptr; // owned (io workspaces only)
	std::mutex mu;                 // held while work that uses the buffers is being enqueued
	DevBuf rerun;      // gathered queries + results of the exact re-run of flagged tcgen05 queries (flat_gemm.cu)
	DevBuf allow_rows; // a caller's label-indexed prefilter translated to row space (explicit labels)
	DevBuf q, allow, out_labels, out_dist, out_counts, work, work2, work3, work4;
	DevBuf probe_mat; // batched IVF probe: (distance, partition) keys [queries][P]
	DevBuf scan_hdr;  // persistent header of the flat scan (kernels.h SCAN_HDR_BYTES), all-ones between passes
	bool scan_hdr_clean = false;
	DevBuf adc_tickets; // persistent arrival tickets of the ADC scan (one per query), zero between launches
	DevBuf ivf1_hdr;    // persistent ticket word of the single-query IVF_PQ kernel (ivfpq1.cuh), zero between launches
	DevBuf scan1_hdr;   // persistent header of the single-query scan (flat_scan1.cuh), handed back clean by every launch
	uint32_t done_seq = 0; // completion-word sequence of host calls on this workspace (io workspaces)
	// small host-buffer calls: pinned staging (query in, packed [labels | distances | counts] out) so that the
	// call is one async H2D + one D2H instead of four pageable copies that each block the host
	void *h_stage = nullptr;
	size_t h_stage_bytes = 0;
	void *stage(size_t bytes);
	explicit Workspace(bool own_stream);
	~Workspace();
};

// What a host-buffer call can tell the device search (all optional): the query as the host holds it (the single-query
// scan then carries it in its kernel parameters: no H2D copy) and a completion word in mapped pinned memory.  `armed`
// comes back true when the kernel that was launched will write done_value to done_flag.
struct HostHints {
	const float *h_queries = nullptr;
	float *d_q_buf = nullptr;    // device buffer the search reads the query from when it cannot take it inline
	bool q_uploaded = false;     // false: the caller skipped the H2D copy, the search uploads it if its path needs it
	volatile uint32_t *done_flag = nullptr;
	uint32_t done_value = 0;
	bool armed = false;
};

struct IvfPqIndex; // ivfpq.h
// an IVF_PQ index in its exchange layout on the host, restricted to the rows that survive a compaction
struct CompactedIndex {
	bool present = false;
	int P = 0, M = 0;
	int64_t n_indexed = 0;
	std::vector<float> centroids, codebook;
	std::vector<int64_t> part_offsets, row_ids;
	std::vector<uint8_t> codes;
};
class FlatGemm;    // flat_gemm.h

class Store {
public:
	Store(const std::string &db_path, const std::string &table, int dim, const std::string &metric, int device);
	~Store();

	int dim() const { return dim_; }
	int metric() const { return metric_; }
	int device() const { return device_; }
	const std::string &dir() const { return dir_; }

	void set_label_base(int64_t base);
	void reserve(int64_t rows);
	int64_t add_host(const float *vectors, int64_t num);                    // returns first label
	int64_t add_device(const float *d_vectors, int64_t num, cudaStream_t s); // returns first label
	// Rows with caller-chosen labels first_label .. first_label + num - 1 (first_label >= every label handed out so far):
	// what a sharded handle uses to place chunks of one global label sequence on several devices (store_labels.cu)
	void add_host_labeled(const float *vectors, int64_t num, int64_t first_label);
	// device rows with an explicit, strictly ascending label per row (labels in host memory): what a partition-sharded
	// IVF_PQ shard is loaded with (sharding.py)
	void add_device_with_labels(const float *d_vectors, int64_t num, const int64_t *labels, cudaStream_t s);
	// Drops the tombstoned rows for good (src/lance_index.cpp:728-751 Vacuum -> LanceDetachedCompact): vectors, extra
	// columns and the IVF_PQ index are rewritten without them, labels are kept (they become explicit).  Returns the
	// number of rows removed.
	int64_t compact();
	bool labels_explicit();
	void delete_labels(const int64_t *labels, int count);
	int64_t live_rows();
	int64_t rows();
	void get_vector(int64_t label, float *out);
	int64_t get_all(int64_t *out_labels, float *out_vectors); // returns live count
	// appends the listed live rows of `src` (storage order) with fresh labels; fills the old -> new mapping
	int merge_from(Store &src, const int64_t *live_labels, int count, int64_t *out_old, int64_t *out_new);

	// host buffers in, host buffers out (the reference-facing call)
	void search_host(const float *queries, int nq, int k, int nprobes, int refine, const uint32_t *allow,
	                 int64_t allow_bits, int64_t *out_labels, float *out_dist, int32_t *out_counts,
	                 const DevBitmap *resident = nullptr);
	// the newer revision's call (src/lance_index.cpp:452-453): the pushed-down WHERE arrives as a string and is
	// evaluated on the device over the extra columns (columns.h); NULL / empty predicate = no filter
	void search_host_predicate(const float *queries, int nq, int k, int nprobes, int refine, const char *predicate,
	                           int64_t *out_labels, float *out_dist, int32_t *out_counts);
	// the device-resident allow bitmap of a predicate (cached per predicate and column epoch)
	std::shared_ptr<DevBitmap> predicate_bitmap(const std::string &predicate, bool *was_cached = nullptr);

	// extra columns (columns.h): values for the rows of labels [first_label, first_label + count)
	void set_column(const std::string &name, ColType type, int64_t first_label, int64_t count, const void *values,
	                const int64_t *offsets, const uint8_t *valid);
	bool has_extra_columns();
	int column_count();
	bool column_info(int i, std::string *name, int32_t *type);
	// Arrow C data interface ingest (rust_lib/src/ffi.rs:148-182): a struct array whose first FixedSizeList<float>
	// child is the vector column and whose other children are extra columns; returns the first label
	// first_label >= 0: the rows take the caller's labels (sharded handles) instead of this store's counter
	int64_t add_arrow(void *arrow_schema, void *arrow_array, int64_t *out_count, int64_t first_label = -1);
	void force_explicit_labels(); // sharded handles: every shard speaks labels through its label array from the start
	void define_columns_from_arrow(void *arrow_schema);
	// everything on the device, enqueued on s
	void search_device(const float *d_queries, int nq, int k, int nprobes, int refine, const uint32_t *d_allow,
	                   int64_t allow_bits, int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts,
	                   cudaStream_t s, bool defer = false, uint32_t *d_overflow = nullptr, HostHints *hints = nullptr,
	                   bool allow_in_row_space = false); // true: d_allow is indexed by row (a predicate bitmap), not by label
	int search_finish(cudaStream_t s);

	// IVF_PQ (ivfpq.cu)
	void create_index(int num_partitions, int num_sub_vectors);
	void set_index(int P, int M, int64_t n_indexed, const float *centroids, const float *codebook,
	               const int64_t *part_offsets, const int64_t *row_ids, const uint8_t *codes);
	void get_index(lance_b200_ivfpq_info *info, float *centroids, float *codebook, int64_t *part_offsets,
	               int64_t *row_ids, uint8_t *codes);
	void drop_index();
	bool ivfpq_takes_inline_query(int k, int nprobes, int refine); // shared lock held, index_ set
	void ivfpq_search_debug(const float *query, int k, int nprobes, int refine, const uint32_t *allow,
	                        int64_t allow_bits, int32_t *out_parts, int64_t *out_cand_labels, float *out_cand_adc,
	                        int32_t *out_ncand, int64_t *out_labels, float *out_dist, int32_t *out_count);

	void save();
	void save_index(FILE *f);                          // ivfpq_host.cuh: the IVF_PQ index behind the rows of store.bin
	void load_index(FILE *f, const std::string &path); // ... and back
	static Store *open(const std::string &db_path, const std::string &table, const std::string &metric, int device);

	void stats(lance_b200_stats_t *out);
	void set_profile(bool on) { timer_.on = on; }
	void read_profile(lance_b200_profile_t *out);
	void set_flat_path(int mode) { flat_path_ = mode; }

private:
	friend struct IvfPqSearch;
	void ensure_capacity(int64_t rows); // exclusive lock held
	std::unique_ptr<Workspace> acquire_io();
	void release_io(std::unique_ptr<Workspace> ws);
	Workspace *scratch_for(cudaStream_t s); // work issued on one stream is ordered, so it can share scratch
	void flat_search_locked(const float *d_queries, int nq, int k, const uint32_t *d_allow, int64_t allow_bits,
	                        int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts, Workspace &ws,
	                        cudaStream_t s, bool defer = false, HostHints *hints = nullptr);
	void ivfpq_search_locked(const float *d_queries, int nq, int k, int nprobes, int refine, const uint32_t *d_allow,
	                         int64_t allow_bits, int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts,
	                         Workspace &ws, cudaStream_t s, int32_t *d_dbg_parts, u64 *d_dbg_cand, int32_t *d_dbg_ncand,
	                         HostHints *hints = nullptr);

	std::string db_path_, table_, dir_;
	int dim_;
	int metric_;
	int device_;
	int num_sms_ = 148;

	std::shared_mutex mu_; // shared: searches; exclusive: add / delete / index build
	DevBuf X_;             // fp32 [cap_rows_][dim_]
	DevBuf deleted_;       // uint32 [cap_rows_/32]
	int64_t cap_rows_ = 0;
	int64_t n_rows_ = 0;
	int64_t n_deleted_ = 0;
	int64_t label_base_ = 0;
	// Labels.  Implicit (default): label = label_base_ + row, nothing stored.  Explicit (after compact() or
	// add_host_labeled with a gap): label_base_ is 0, every kernel works in ROW space and the API boundary translates —
	// row_labels_ (device) / row_labels_host_ (ascending) give row -> label, a binary search gives label -> row.
	bool explicit_labels_ = false;
	int64_t next_label_ = 0; // the label the next appended row gets
	DevBuf row_labels_;      // int64 [cap_rows_], explicit only
	std::vector<int64_t> row_labels_host_;
	void make_labels_explicit();                 // exclusive lock held
	void append_labels(int64_t first, int64_t num); // exclusive lock held, capacity ensured, before n_rows_ grows
	int64_t row_of_label(int64_t label) const;   // -1 when the label is not in the table (tombstoned rows still are)
	int64_t label_of_row(int64_t row) const { return explicit_labels_ ? row_labels_host_[(size_t)row] : label_base_ + row; }
	void extract_index_for_compaction(const std::vector<int64_t> &new_row_of, CompactedIndex &out); // ivfpq_host.cuh
	void install_index_locked(int P, int M, int64_t n_indexed, const float *centroids, const float *codebook,
	                          const int64_t *part_offsets, const int64_t *row_ids, const uint8_t *codes);
	std::vector<uint32_t> deleted_host_; // mirror of the tombstones (count + save)
	cudaStream_t copy_stream_ = nullptr;

	std::unique_ptr<IvfPqIndex> index_;
	ColumnStore columns_; // mutated under the exclusive lock; bitmaps are evaluated under the shared one

	std::mutex ws_mu_;
	std::vector<std::unique_ptr<Workspace>> io_pool_;
	std::vector<std::pair<cudaStream_t, std::unique_ptr<Workspace>>> stream_ws_;

	std::mutex gemm_mu_;
	std::unique_ptr<FlatGemm> flat_gemm_;

	std::atomic<int> flat_path_{0};
	KernelTimer timer_;
	// counters
	int64_t launches_ = 0; // guarded by launches_mu_ when merged from a workspace
	std::mutex stat_mu_;
	int64_t searches_ = 0, simt_batches_ = 0, tc_batches_ = 0, tc_fallback_ = 0, ivfpq_batches_ = 0;
	void add_launches(int64_t n) {
		std::lock_guard<std::mutex> g(stat_mu_);
		launches_ += n;
	}
};

void launch_relabel(int64_t *labels, long long n, const int64_t *row_labels, long long n_rows, cudaStream_t s);   // store_labels.cu
void launch_allow_by_row(const uint32_t *allow_by_label, long long allow_bits, const int64_t *row_labels, long long n_rows,
                         uint32_t *out, cudaStream_t s);
int arrow_schema_vector_dim(void *arrow_schema); // store_columns.cu: dimension of the FixedSizeList<float> child; throws
int default_device();
void set_default_device(int d);

} // namespace lb
