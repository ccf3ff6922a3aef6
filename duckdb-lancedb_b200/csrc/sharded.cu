// sharded.cu — see sharded.h.
#include "sharded.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <shared_mutex>
#include <unordered_set>

namespace lb {

static std::shared_mutex g_reg_mu;
static std::unordered_set<void *> g_reg;

ShardedStore *as_sharded(void *handle) {
	std::shared_lock<std::shared_mutex> lk(g_reg_mu);
	return g_reg.count(handle) ? static_cast<ShardedStore *>(handle) : nullptr;
}
void register_sharded(ShardedStore *s) {
	std::unique_lock<std::shared_mutex> lk(g_reg_mu);
	g_reg.insert(s);
}
void unregister_sharded(ShardedStore *s) {
	std::unique_lock<std::shared_mutex> lk(g_reg_mu);
	g_reg.erase(s);
}

static void check_devices(const std::vector<int> &devices) {
	if (devices.empty()) throw Error("a sharded handle needs at least one device");
	int count = 0;
	LB_CUDA(cudaGetDeviceCount(&count));
	for (int d : devices)
		if (d < 0 || d >= count) throw Error("device " + std::to_string(d) + " does not exist (" + std::to_string(count) + " visible)");
}

static std::string shard_table(const std::string &table, size_t i) {
	return (table.empty() ? std::string("vectors") : table) + ".shard" + std::to_string(i);
}
static std::string manifest_path(const std::string &db_path, const std::string &table) {
	return db_path + "/" + (table.empty() ? std::string("vectors") : table) + ".b200shards";
}

ShardedStore::ShardedStore(const std::vector<int> &devices, int dim, const std::string &metric, const std::string &db_path,
                           const std::string &table)
    : dim_(dim), db_path_(db_path), table_(table) {
	check_devices(devices);
	for (size_t i = 0; i < devices.size(); i++) {
		shards_.emplace_back(new Store(db_path, shard_table(table, i), dim, metric, devices[i]));
		shards_.back()->force_explicit_labels();
	}
	start_workers();
}

void ShardedStore::save() {
	if (db_path_.empty()) throw Error("sharded handle has no db_path to save under");
	std::lock_guard<std::mutex> g(fan_mu_);
	std::lock_guard<std::mutex> a(add_mu_);
	run_on_all([&](int s) { shards_[(size_t)s]->save(); });
	const std::string path = manifest_path(db_path_, table_);
	FILE *f = fopen((path + ".tmp").c_str(), "w");
	if (!f) throw Error("cannot write " + path);
	bool ok = fprintf(f, "LB200SHARDS 1\n%zu %d %lld %zu\n", shards_.size(), dim_, (long long)next_label_, chunk_first_.size()) > 0;
	for (size_t i = 0; i < chunk_first_.size() && ok; i++)
		ok = fprintf(f, "%lld %d\n", (long long)chunk_first_[i], chunk_shard_[i]) > 0;
	ok = (fclose(f) == 0) && ok;
	if (!ok || rename((path + ".tmp").c_str(), path.c_str()) != 0) throw Error("cannot write " + path);
}

int ShardedStore::saved_shards(const std::string &db_path, const std::string &table) {
	FILE *f = fopen(manifest_path(db_path, table).c_str(), "r");
	if (!f) return 0;
	char magic[32] = {0};
	int version = 0;
	size_t G = 0;
	const bool ok = fscanf(f, "%31s %d %zu", magic, &version, &G) == 3 && std::string(magic) == "LB200SHARDS";
	fclose(f);
	return ok ? (int)G : 0;
}

void ShardedStore::forget(const std::string &db_path, const std::string &table) {
	if (!db_path.empty()) remove(manifest_path(db_path, table).c_str());
}

ShardedStore *ShardedStore::open(const std::string &db_path, const std::string &table, const std::string &metric,
                                 const std::vector<int> &devices) {
	check_devices(devices);
	const std::string path = manifest_path(db_path, table);
	FILE *f = fopen(path.c_str(), "r");
	if (!f) throw Error("no saved sharded store at " + path);
	char magic[32] = {0};
	int version = 0, dim = 0;
	size_t G = 0, nchunks = 0;
	long long next_label = 0;
	bool ok = fscanf(f, "%31s %d", magic, &version) == 2 && std::string(magic) == "LB200SHARDS" && version == 1 &&
	          fscanf(f, "%zu %d %lld %zu", &G, &dim, &next_label, &nchunks) == 4;
	std::unique_ptr<ShardedStore> sh(new ShardedStore());
	if (ok) {
		sh->chunk_first_.resize(nchunks);
		sh->chunk_shard_.resize(nchunks);
		for (size_t i = 0; i < nchunks && ok; i++) {
			long long first = 0;
			int shard = 0;
			ok = fscanf(f, "%lld %d", &first, &shard) == 2 && shard >= 0 && (size_t)shard < G &&
			     (i == 0 || first > sh->chunk_first_[i - 1]);
			sh->chunk_first_[i] = first, sh->chunk_shard_[i] = shard;
		}
	}
	fclose(f);
	if (!ok) throw Error("corrupt sharded manifest " + path);
	if (G != devices.size())
		throw Error("the table was saved with " + std::to_string(G) + " shards, " + std::to_string(devices.size()) + " devices were given");
	sh->dim_ = dim, sh->db_path_ = db_path, sh->table_ = table, sh->next_label_ = next_label;
	for (size_t i = 0; i < G; i++) {
		sh->shards_.emplace_back(Store::open(db_path, shard_table(table, i), metric, devices[i]));
		if (sh->shards_.back()->dim() != dim) throw Error("shard " + std::to_string(i) + " has another dimension than the manifest");
		sh->shards_.back()->force_explicit_labels();
	}
	sh->start_workers();
	return sh.release();
}

void ShardedStore::start_workers() {
	for (size_t i = 0; i < shards_.size(); i++) workers_.emplace_back(new Worker());
	for (size_t i = 0; i < workers_.size(); i++) {
		Worker *w = workers_[i].get();
		w->th = std::thread([w] {
			std::unique_lock<std::mutex> lk(w->mu);
			while (true) {
				// back-to-back searches: poll for the next job for a moment before sleeping (a futex wake-up costs
				// tens of microseconds on each side of a 1-2 ms search)
				lk.unlock();
				const auto t_spin = std::chrono::steady_clock::now();
				while (!w->posted.load(std::memory_order_acquire) &&
				       std::chrono::steady_clock::now() - t_spin < std::chrono::microseconds(200))
					std::this_thread::yield();
				lk.lock();
				w->cv.wait(lk, [w] { return w->has_job || w->stop; });
				if (w->stop) return;
				w->posted.store(false, std::memory_order_relaxed);
				std::function<void()> job = std::move(w->job);
				w->has_job = false;
				lk.unlock();
				std::string err;
				try {
					job();
				} catch (const std::exception &e) {
					err = e.what();
				} catch (...) {
					err = "unknown error";
				}
				lk.lock();
				w->error = err;
				w->done = true;
				w->finished.store(true, std::memory_order_release);
				w->cv.notify_all();
			}
		});
	}
}

ShardedStore::~ShardedStore() {
	for (auto &w : workers_) {
		{
			std::lock_guard<std::mutex> g(w->mu);
			w->stop = true;
		}
		w->cv.notify_all();
		if (w->th.joinable()) w->th.join();
	}
}

void ShardedStore::run_on_all(const std::function<void(int)> &fn) {
	for (size_t i = 0; i < workers_.size(); i++) {
		Worker *w = workers_[i].get();
		std::lock_guard<std::mutex> g(w->mu);
		w->job = [&fn, i] { fn((int)i); };
		w->has_job = true;
		w->done = false;
		w->finished.store(false, std::memory_order_relaxed);
		w->error.clear();
		w->posted.store(true, std::memory_order_release);
		w->cv.notify_all();
	}
	std::string first;
	for (auto &w : workers_) {
		// the caller polls like cudaStreamSynchronize does (the devices are busy for the whole wait)
		while (!w->finished.load(std::memory_order_acquire)) std::this_thread::yield();
		std::unique_lock<std::mutex> lk(w->mu);
		w->cv.wait(lk, [&w] { return w->done; });
		if (first.empty() && !w->error.empty()) first = w->error;
	}
	if (!first.empty()) throw Error(first);
}

int ShardedStore::least_filled() {
	int best = 0;
	int64_t best_rows = -1;
	for (size_t i = 0; i < shards_.size(); i++) {
		const int64_t r = shards_[i]->rows();
		if (best_rows < 0 || r < best_rows) best = (int)i, best_rows = r;
	}
	return best;
}

int ShardedStore::shard_of_label(int64_t label) {
	auto it = std::upper_bound(chunk_first_.begin(), chunk_first_.end(), label);
	if (it == chunk_first_.begin()) return -1;
	return chunk_shard_[(size_t)(it - chunk_first_.begin()) - 1];
}

int64_t ShardedStore::add_host(const float *vectors, int64_t num) {
	std::lock_guard<std::mutex> g(add_mu_);
	const int64_t first = next_label_;
	if (num <= 0) return first;
	const int s = least_filled();
	shards_[(size_t)s]->add_host_labeled(vectors, num, first);
	chunk_first_.push_back(first);
	chunk_shard_.push_back(s);
	next_label_ += num;
	return first;
}

int64_t ShardedStore::add_arrow(void *arrow_schema, void *arrow_array, int64_t *out_count) {
	std::lock_guard<std::mutex> g(add_mu_);
	const int s = least_filled();
	int64_t n = 0;
	const int64_t first = shards_[(size_t)s]->add_arrow(arrow_schema, arrow_array, &n, next_label_);
	if (out_count) *out_count = n;
	if (n > 0) {
		chunk_first_.push_back(first);
		chunk_shard_.push_back(s);
		next_label_ += n;
	}
	return first;
}

void ShardedStore::define_columns_from_arrow(void *arrow_schema) {
	for (auto &s : shards_) s->define_columns_from_arrow(arrow_schema);
}

void ShardedStore::set_column(const std::string &name, ColType type, int64_t first_label, int64_t count, const void *values,
                              const int64_t *offsets, const uint8_t *valid) {
	std::lock_guard<std::mutex> g(add_mu_);
	// the range is cut at the boundaries of the appended batches; every piece goes to the shard that holds it
	int64_t done = 0;
	if (count == 0) {
		for (auto &s : shards_) s->set_column(name, type, 0, 0, nullptr, nullptr, nullptr);
		return;
	}
	while (done < count) {
		const int64_t lab = first_label + done;
		auto it = std::upper_bound(chunk_first_.begin(), chunk_first_.end(), lab);
		if (it == chunk_first_.begin()) throw Error("column values for label " + std::to_string(lab) + " fall outside the table");
		const size_t ci = (size_t)(it - chunk_first_.begin()) - 1;
		const int64_t chunk_end = ci + 1 < chunk_first_.size() ? chunk_first_[ci + 1] : next_label_;
		const int64_t piece = std::min(count - done, chunk_end - lab);
		if (piece <= 0) throw Error("column values for label " + std::to_string(lab) + " fall outside the table");
		const void *vals = values;
		const int64_t *offs = offsets ? offsets + done : nullptr; // string offsets stay absolute into `values`
		if (type == COL_I64 || type == COL_F64) vals = static_cast<const char *>(values) + (size_t)done * 8;
		else if (type == COL_BOOL) vals = static_cast<const char *>(values) + (size_t)done;
		shards_[(size_t)chunk_shard_[ci]]->set_column(name, type, lab, piece, vals, offs, valid ? valid + done : nullptr);
		done += piece;
	}
}

bool ShardedStore::has_extra_columns() {
	for (auto &s : shards_)
		if (s->has_extra_columns()) return true;
	return false;
}

void ShardedStore::delete_labels(const int64_t *labels, int count) {
	std::lock_guard<std::mutex> g(add_mu_);
	std::vector<std::vector<int64_t>> per(shards_.size());
	for (int i = 0; i < count; i++) {
		const int s = shard_of_label(labels[i]);
		if (s >= 0) per[(size_t)s].push_back(labels[i]);
	}
	for (size_t s = 0; s < shards_.size(); s++)
		if (!per[s].empty()) shards_[s]->delete_labels(per[s].data(), (int)per[s].size());
}

int64_t ShardedStore::live_rows() {
	int64_t n = 0;
	for (auto &s : shards_) n += s->live_rows();
	return n;
}
int64_t ShardedStore::rows() {
	int64_t n = 0;
	for (auto &s : shards_) n += s->rows();
	return n;
}

void ShardedStore::get_vector(int64_t label, float *out) {
	int s;
	{
		std::lock_guard<std::mutex> g(add_mu_);
		s = shard_of_label(label);
	}
	if (s < 0) throw Error("label " + std::to_string(label) + " not found");
	shards_[(size_t)s]->get_vector(label, out);
}

int64_t ShardedStore::get_all(int64_t *out_labels, float *out_vectors) {
	int64_t total = 0;
	for (auto &s : shards_) {
		const int64_t c = s->get_all(out_labels ? out_labels + total : nullptr,
		                             out_vectors ? out_vectors + (size_t)total * (size_t)dim_ : nullptr);
		total += c;
	}
	return total;
}

int64_t ShardedStore::compact() {
	std::lock_guard<std::mutex> g(fan_mu_);
	std::vector<int64_t> removed(shards_.size(), 0);
	run_on_all([&](int s) { removed[(size_t)s] = shards_[(size_t)s]->compact(); });
	int64_t n = 0;
	for (auto r : removed) n += r;
	return n;
}

void ShardedStore::create_index(int num_partitions, int num_sub_vectors) {
	std::lock_guard<std::mutex> g(fan_mu_);
	if (rows() == 0) throw Error("cannot train an IVF_PQ index on an empty table");
	run_on_all([&](int s) {
		Store &st = *shards_[(size_t)s];
		if (st.rows() > 0) st.create_index(num_partitions, num_sub_vectors);
	});
}

void ShardedStore::drop_index() {
	std::lock_guard<std::mutex> g(fan_mu_);
	for (auto &s : shards_) s->drop_index();
}

bool ShardedStore::has_index() {
	lance_b200_stats_t st;
	stats(&st);
	return st.has_index != 0;
}

// the shards' counters added up; device = the first shard's, has_index = any shard holds one
void ShardedStore::stats(lance_b200_stats_t *out) {
	lance_b200_stats_t sum = {};
	for (size_t i = 0; i < shards_.size(); i++) {
		lance_b200_stats_t st;
		shards_[i]->stats(&st);
		sum.kernel_launches += st.kernel_launches, sum.searches += st.searches, sum.flat_simt_batches += st.flat_simt_batches;
		sum.flat_tc_batches += st.flat_tc_batches, sum.tc_fallback_queries += st.tc_fallback_queries;
		sum.ivfpq_batches += st.ivfpq_batches, sum.rows += st.rows, sum.live_rows += st.live_rows;
		if (i == 0) sum.device = st.device;
		sum.has_index |= st.has_index;
	}
	*out = sum;
}

void ShardedStore::set_flat_path(int mode) {
	for (auto &s : shards_) s->set_flat_path(mode);
}

void ShardedStore::search_host(const float *queries, int nq, int k, int nprobes, int refine, const uint32_t *allow,
                               int64_t allow_bits, const char *predicate, int64_t *out_labels, float *out_dist,
                               int32_t *out_counts) {
	if (nq <= 0) return;
	if (k <= 0) {
		for (int i = 0; i < nq; i++) out_counts[i] = 0;
		return;
	}
	std::lock_guard<std::mutex> g(fan_mu_);
	const size_t G = shards_.size(), per = (size_t)nq * k;
	stage_labels_.resize(G * per);
	stage_dist_.resize(G * per);
	stage_counts_.resize(G * (size_t)nq);
	// Every device searches its shard concurrently (one host thread per device issues and waits on that device's
	// stream).  The same threads then merge: a worker that has its lists waits until all have arrived and takes the
	// queries [s nq / G, (s + 1) nq / G) — one thread merging a 1024-query batch of 8 shards was a third of the call.
	std::atomic<int> arrived{0};
	std::atomic<bool> failed{false};
	static const bool trace = getenv("LANCE_B200_SHARD_TRACE") != nullptr; // per-shard phase times on stderr
	const auto t_call = std::chrono::steady_clock::now();
	auto us_since = [&] { return (long long)std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t_call).count(); };
	std::vector<long long> t_start(G), t_search(G), t_all(G), t_merge(G);
	run_on_all([&](int s) {
		t_start[(size_t)s] = us_since();
		try {
			Store &st = *shards_[(size_t)s];
			int64_t *L = stage_labels_.data() + (size_t)s * per;
			float *D = stage_dist_.data() + (size_t)s * per;
			int32_t *C = stage_counts_.data() + (size_t)s * nq;
			if (st.rows() == 0) std::fill(C, C + nq, 0);
			else if (predicate && *predicate) st.search_host_predicate(queries, nq, k, nprobes, refine, predicate, L, D, C);
			else st.search_host(queries, nq, k, nprobes, refine, allow, allow_bits, L, D, C);
		} catch (...) {
			failed.store(true);
			arrived.fetch_add(1, std::memory_order_release);
			throw;
		}
		t_search[(size_t)s] = us_since();
		arrived.fetch_add(1, std::memory_order_release);
		while (arrived.load(std::memory_order_acquire) < (int)G) std::this_thread::yield();
		t_all[(size_t)s] = us_since();
		if (failed.load()) return;
		const int q_lo = (int)((long long)nq * s / (long long)G), q_hi = (int)((long long)nq * (s + 1) / (long long)G);
		merge_queries(q_lo, q_hi, nq, k, out_labels, out_dist, out_counts);
		t_merge[(size_t)s] = us_since();
	});
	if (trace) {
		const long long t_end = us_since();
		for (size_t s = 0; s < G; s++)
			fprintf(stderr, "[lance_b200 shard %zu] start %lld us, search done %lld, all arrived %lld, merged %lld, call %lld\n", s,
			        t_start[s], t_search[s], t_all[s], t_merge[s], t_end);
	}
}

// per query a k-way merge of the shards' ascending lists under (distance, label) — the order of make_key (common.cuh);
// labels of different shards never coincide, so the order is total
void ShardedStore::merge_queries(int q_lo, int q_hi, int nq, int k, int64_t *out_labels, float *out_dist, int32_t *out_counts) {
	const size_t G = shards_.size(), per = (size_t)nq * k;
	struct Head {
		uint32_t key; // ordered distance
		int64_t label;
		const int64_t *L;
		const float *D;
		int left;
	};
	std::vector<Head> heads(G);
	for (int q = q_lo; q < q_hi; q++) {
		size_t live = 0;
		for (size_t s = 0; s < G; s++) {
			const int c = std::min(k, (int)stage_counts_[s * (size_t)nq + q]);
			if (c <= 0) continue;
			Head &h = heads[live++];
			h.L = stage_labels_.data() + s * per + (size_t)q * k;
			h.D = stage_dist_.data() + s * per + (size_t)q * k;
			h.left = c, h.key = float_to_ordered(h.D[0]), h.label = h.L[0];
		}
		int out = 0;
		int64_t *OL = out_labels + (size_t)q * k;
		float *OD = out_dist + (size_t)q * k;
		while (out < k && live > 0) {
			size_t best = 0;
			for (size_t i = 1; i < live; i++)
				if (heads[i].key < heads[best].key || (heads[i].key == heads[best].key && heads[i].label < heads[best].label)) best = i;
			Head &h = heads[best];
			OL[out] = h.label, OD[out] = h.D[0];
			out++;
			if (--h.left == 0) {
				heads[best] = heads[--live];
			} else {
				h.L++, h.D++;
				h.key = float_to_ordered(h.D[0]), h.label = h.L[0];
			}
		}
		out_counts[q] = out;
		for (int i = out; i < k; i++) OL[i] = -1, OD[i] = INFINITY;
	}
}

} // namespace lb
