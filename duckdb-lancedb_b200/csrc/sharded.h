// sharded.h — one `lance_*` handle over the GPUs of a box, inside ONE process (the consumer is an embedded
// database, so there is no launcher and no rank: LanceIndex::Search in a DuckDB process must reach every device).
//
// Rows are spread over per-device Stores in whole appended batches (the least-filled shard takes the next batch); the
// global label sequence (rust_lib/src/lance_manager.rs:214, 232: one atomic counter) is kept, every shard stores its
// labels explicitly (store.h "Labels").  A search replicates the queries to every shard — one persistent host thread
// per shard drives that device's stream, so the devices run concurrently — and the per-shard top-k lists are merged
// under the library's (distance, label) order.  The path has no collective: every list already sits in host memory
// when the call returns (the C ABI hands results back in host buffers), so the exchange step of the multi-process form
// (one NCCL all_gather, sharding.py) degenerates to a host-side k-way merge.
#pragma once
#include <atomic>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "store.h"

namespace lb {

class ShardedStore {
public:
	// db_path empty: in memory only.  Else shard i persists as table "<table>.shard<i>" under db_path (store.bin each) and
	// save() adds "<db_path>/<table>.b200shards": shard count, label counter and the batch -> shard map.
	ShardedStore(const std::vector<int> &devices, int dim, const std::string &metric, const std::string &db_path = "",
	             const std::string &table = "");
	~ShardedStore();
	void save();
	static int saved_shards(const std::string &db_path, const std::string &table); // 0: no sharded table saved there
	static void forget(const std::string &db_path, const std::string &table);      // a plain table takes that name: drop the manifest
	// reopens what save() wrote; devices.size() must be the saved shard count (shard i goes to devices[i])
	static ShardedStore *open(const std::string &db_path, const std::string &table, const std::string &metric,
	                          const std::vector<int> &devices);

	int dim() const { return dim_; }
	int num_shards() const { return (int)shards_.size(); }
	Store &shard(int i) { return *shards_[(size_t)i]; }

	int64_t add_host(const float *vectors, int64_t num); // returns the first label
	int64_t add_arrow(void *arrow_schema, void *arrow_array, int64_t *out_count);
	void define_columns_from_arrow(void *arrow_schema);
	void set_column(const std::string &name, ColType type, int64_t first_label, int64_t count, const void *values,
	                const int64_t *offsets, const uint8_t *valid);
	bool has_extra_columns();
	void delete_labels(const int64_t *labels, int count);
	int64_t live_rows();
	int64_t rows();
	void get_vector(int64_t label, float *out);
	int64_t get_all(int64_t *out_labels, float *out_vectors); // live rows of every shard, shard by shard; returns the count
	int64_t compact();
	// IVF_PQ on a sharded table: every shard trains and keeps an index of ITS rows with the caller's parameters (the
	// partition count is clamped to the shard's rows); a search probes nprobes lists on every shard and the lists are
	// merged like flat results.  (The multi-process form shards ONE index by partition instead: DESIGN.md 6.0.)
	void create_index(int num_partitions, int num_sub_vectors);
	void drop_index();
	bool has_index();
	void stats(lance_b200_stats_t *out);
	void set_flat_path(int mode);
	// predicate (nullable) XOR allow bitmap (nullable, indexed by label)
	void search_host(const float *queries, int nq, int k, int nprobes, int refine, const uint32_t *allow, int64_t allow_bits,
	                 const char *predicate, int64_t *out_labels, float *out_dist, int32_t *out_counts);

private:
	ShardedStore() = default; // open()
	void start_workers();
	std::string db_path_, table_;
	struct Worker { // one host thread per shard: it owns that device's calls of a fan-out
		std::thread th;
		std::mutex mu;
		std::condition_variable cv;
		std::function<void()> job;
		bool has_job = false, done = false, stop = false;
		std::atomic<bool> posted{false}, finished{false}; // lock-free mirrors of has_job / done for the polling phases
		std::string error;
	};
	void run_on_all(const std::function<void(int)> &fn); // fn(shard) on every worker, waits, rethrows the first error
	int least_filled();                                   // by rows held
	// the shards' staged lists of queries [q_lo, q_hi) -> the caller's buffers
	void merge_queries(int q_lo, int q_hi, int nq, int k, int64_t *out_labels, float *out_dist, int32_t *out_counts);

	int dim_ = 0;
	std::vector<std::unique_ptr<Store>> shards_;
	std::vector<std::unique_ptr<Worker>> workers_;
	std::mutex add_mu_;   // appends are serialised: they draw from one label counter
	std::mutex fan_mu_;   // one fan-out at a time uses the workers and the staging buffers
	int64_t next_label_ = 0;
	std::vector<int64_t> chunk_first_; // label -> shard: ascending first labels of the appended batches ...
	std::vector<int> chunk_shard_;     // ... and the shard each went to
	std::vector<int64_t> stage_labels_;
	std::vector<float> stage_dist_;
	std::vector<int32_t> stage_counts_;
	int shard_of_label(int64_t label);
};

ShardedStore *as_sharded(void *handle); // nullptr when `handle` is a plain Store
void register_sharded(ShardedStore *s);
void unregister_sharded(ShardedStore *s);

} // namespace lb
