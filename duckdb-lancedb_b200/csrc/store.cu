// store.cu — device store lifecycle and the flat-search orchestration (see store.h).
#include "store.h"

#include <sys/stat.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "ivfpq.h"
#include "flat_gemm.h"

namespace lb {

// ------------------------------------------------------------------ device selection
static std::atomic<int> g_default_device{-1};

int default_device() {
	int d = g_default_device.load();
	if (d >= 0) return d;
	const char *e = getenv("LANCE_B200_DEVICE");
	if (!e) e = getenv("LOCAL_RANK"); // one process per GPU under torchrun
	int count = 0;
	if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0)
		throw Error("no CUDA device available: liblance_b200 has no CPU fallback");
	d = e ? atoi(e) : 0;
	if (d < 0 || d >= count) d = 0;
	return d;
}
void set_default_device(int d) { g_default_device.store(d); }

// ------------------------------------------------------------------ buffers
void DevBuf::release() {
	if (p) cudaFree(p);
	p = nullptr;
	bytes = 0;
}
void DevBuf::alloc(size_t n) {
	release();
	if (n == 0) return;
	cudaError_t e = cudaMalloc(&p, n);
	if (e != cudaSuccess) {
		p = nullptr;
		(void)cudaGetLastError();
		throw Error("device allocation of " + std::to_string(n) + " bytes failed: " + cudaGetErrorString(e));
	}
	bytes = n;
}
void DevBuf::ensure(size_t n) {
	if (n > bytes) alloc(n + n / 4 + 256);
}

Workspace::Workspace(bool own_stream) {
	if (own_stream) LB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
}
Workspace::~Workspace() {
	if (stream) cudaStreamDestroy(stream);
	if (h_stage) cudaFreeHost(h_stage);
}
void *Workspace::stage(size_t bytes) {
	if (bytes > h_stage_bytes) {
		if (h_stage) cudaFreeHost(h_stage);
		h_stage = nullptr, h_stage_bytes = 0;
		LB_CUDA(cudaMallocHost(&h_stage, bytes + bytes / 2 + 4096));
		h_stage_bytes = bytes + bytes / 2 + 4096;
	}
	return h_stage;
}

// ------------------------------------------------------------------ Store
static int metric_from_string(const std::string &m) {
	// rust_lib/src/lance_manager.rs:493-497
	if (m == "cosine") return METRIC_COSINE;
	if (m == "dot" || m == "ip") return METRIC_DOT;
	return METRIC_L2;
}

Store::Store(const std::string &db_path, const std::string &table, int dim, const std::string &metric, int device)
    : db_path_(db_path), table_(table), dim_(dim), metric_(metric_from_string(metric)), device_(device) {
	if (dim <= 0) throw Error("dimension must be positive");
	LB_CUDA(cudaSetDevice(device_));
	cudaDeviceProp prop;
	LB_CUDA(cudaGetDeviceProperties(&prop, device_));
	num_sms_ = prop.multiProcessorCount;
	if (prop.major < 10)
		throw Error(std::string("liblance_b200 targets sm_100a (B200); device is ") + prop.name);
	LB_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
	if (!db_path_.empty()) dir_ = db_path_ + "/" + (table_.empty() ? std::string("vectors") : table_) + ".b200";
}

Store::~Store() {
	cudaSetDevice(device_);
	cudaDeviceSynchronize();
	stream_ws_.clear();
	io_pool_.clear();
	flat_gemm_.reset();
	index_.reset();
	if (copy_stream_) cudaStreamDestroy(copy_stream_);
}

std::unique_ptr<Workspace> Store::acquire_io() {
	{
		std::lock_guard<std::mutex> g(ws_mu_);
		if (!io_pool_.empty()) {
			auto ws = std::move(io_pool_.back());
			io_pool_.pop_back();
			return ws;
		}
	}
	return std::unique_ptr<Workspace>(new Workspace(true));
}
void Store::release_io(std::unique_ptr<Workspace> ws) {
	std::lock_guard<std::mutex> g(ws_mu_);
	io_pool_.push_back(std::move(ws));
}
Workspace *Store::scratch_for(cudaStream_t s) {
	std::lock_guard<std::mutex> g(ws_mu_);
	for (auto &e : stream_ws_)
		if (e.first == s) return e.second.get();
	stream_ws_.emplace_back(s, std::unique_ptr<Workspace>(new Workspace(false)));
	return stream_ws_.back().second.get();
}

void Store::set_label_base(int64_t base) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	if (n_rows_ != 0) throw Error("label base can only be set on an empty store");
	label_base_ = base;
	next_label_ = base;
}

void Store::ensure_capacity(int64_t rows) {
	if (rows <= cap_rows_) return;
	if (rows >= (1ll << 32)) throw Error("more than 2^32 rows per device store are not supported");
	int64_t ncap = std::max<int64_t>(rows, std::max<int64_t>(cap_rows_ * 2, 4096));
	ncap = (ncap + 255) & ~255ll;
	if (ncap >= (1ll << 32)) ncap = (1ll << 32) - 256;
	DevBuf nx, nd;
	nx.alloc((size_t)ncap * dim_ * sizeof(float));
	nd.alloc((size_t)(ncap / 32) * sizeof(uint32_t));
	LB_CUDA(cudaMemsetAsync(nd.p, 0, nd.bytes, copy_stream_));
	if (n_rows_ > 0) {
		LB_CUDA(cudaMemcpyAsync(nx.p, X_.p, (size_t)n_rows_ * dim_ * sizeof(float), cudaMemcpyDeviceToDevice, copy_stream_));
		LB_CUDA(cudaMemcpyAsync(nd.p, deleted_.p, (size_t)((n_rows_ + 31) / 32) * sizeof(uint32_t),
		                        cudaMemcpyDeviceToDevice, copy_stream_));
	}
	LB_CUDA(cudaStreamSynchronize(copy_stream_));
	std::swap(X_.p, nx.p);
	std::swap(X_.bytes, nx.bytes);
	std::swap(deleted_.p, nd.p);
	std::swap(deleted_.bytes, nd.bytes);
	cap_rows_ = ncap;
	columns_.grow(ncap, copy_stream_);
	if (explicit_labels_) {
		DevBuf nl;
		nl.alloc((size_t)ncap * sizeof(int64_t));
		if (n_rows_ > 0)
			LB_CUDA(cudaMemcpyAsync(nl.p, row_labels_.p, (size_t)n_rows_ * sizeof(int64_t), cudaMemcpyDeviceToDevice, copy_stream_));
		LB_CUDA(cudaStreamSynchronize(copy_stream_));
		std::swap(row_labels_.p, nl.p);
		std::swap(row_labels_.bytes, nl.bytes);
	}
	deleted_host_.resize((size_t)(ncap / 32), 0u);
	if (flat_gemm_) flat_gemm_->invalidate();
}

void Store::reserve(int64_t rows) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	ensure_capacity(rows);
}

int64_t Store::add_host(const float *vectors, int64_t num) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (num <= 0) return next_label_;
	ensure_capacity(n_rows_ + num);
	LB_CUDA(cudaMemcpyAsync(X_.as<float>() + (size_t)n_rows_ * dim_, vectors, (size_t)num * dim_ * sizeof(float),
	                        cudaMemcpyHostToDevice, copy_stream_));
	LB_CUDA(cudaStreamSynchronize(copy_stream_));
	const int64_t first = next_label_;
	append_labels(first, num);
	n_rows_ += num;
	return first;
}

int64_t Store::add_device(const float *d_vectors, int64_t num, cudaStream_t s) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (num <= 0) return next_label_;
	ensure_capacity(n_rows_ + num);
	LB_CUDA(cudaMemcpyAsync(X_.as<float>() + (size_t)n_rows_ * dim_, d_vectors, (size_t)num * dim_ * sizeof(float),
	                        cudaMemcpyDeviceToDevice, s));
	LB_CUDA(cudaStreamSynchronize(s));
	const int64_t first = next_label_;
	append_labels(first, num);
	n_rows_ += num;
	return first;
}

void Store::delete_labels(const int64_t *labels, int count) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (count <= 0) return;
	std::vector<int64_t> rows;
	rows.reserve(count);
	for (int i = 0; i < count; i++) {
		int64_t r = row_of_label(labels[i]);
		if (r < 0 || r >= n_rows_) continue; // deleting an unknown label matches nothing (lance_manager.rs:461-471)
		uint32_t &w = deleted_host_[(size_t)(r >> 5)];
		uint32_t bit = 1u << (r & 31);
		if (w & bit) continue;
		w |= bit;
		n_deleted_++;
		rows.push_back(r);
	}
	if (rows.empty()) return;
	DevBuf tmp;
	tmp.alloc(rows.size() * sizeof(int64_t));
	LB_CUDA(cudaMemcpyAsync(tmp.p, rows.data(), rows.size() * sizeof(int64_t), cudaMemcpyHostToDevice, copy_stream_));
	int64_t l = 0;
	launch_set_bits(deleted_.as<uint32_t>(), tmp.as<int64_t>(), (int)rows.size(), n_rows_, copy_stream_, &l);
	LB_CUDA(cudaStreamSynchronize(copy_stream_));
	add_launches(l);
}

int64_t Store::live_rows() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	return n_rows_ - n_deleted_;
}
int64_t Store::rows() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	return n_rows_;
}

void Store::get_vector(int64_t label, float *out) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	int64_t r = row_of_label(label);
	if (r < 0 || r >= n_rows_ || ((deleted_host_[(size_t)(r >> 5)] >> (r & 31)) & 1u))
		throw Error("label " + std::to_string(label) + " not found");
	LB_CUDA(cudaMemcpy(out, X_.as<float>() + (size_t)r * dim_, (size_t)dim_ * sizeof(float), cudaMemcpyDeviceToHost));
}

int64_t Store::get_all(int64_t *out_labels, float *out_vectors) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	int64_t live = n_rows_ - n_deleted_;
	if (!out_labels && !out_vectors) return live;
	// copy runs of live rows
	int64_t w = 0, r = 0;
	while (r < n_rows_) {
		while (r < n_rows_ && ((deleted_host_[(size_t)(r >> 5)] >> (r & 31)) & 1u)) r++;
		int64_t b = r;
		while (r < n_rows_ && !((deleted_host_[(size_t)(r >> 5)] >> (r & 31)) & 1u)) r++;
		if (r > b) {
			if (out_vectors)
				LB_CUDA(cudaMemcpy(out_vectors + (size_t)w * dim_, X_.as<float>() + (size_t)b * dim_,
				                   (size_t)(r - b) * dim_ * sizeof(float), cudaMemcpyDeviceToHost));
			if (out_labels)
				for (int64_t i = b; i < r; i++) out_labels[w + (i - b)] = label_of_row(i);
			w += r - b;
		}
	}
	return live;
}

// rust_lib/src/lance_manager.rs:307-393 (merge_from): the rows of `src` whose label is listed, exists and is not
// tombstoned are appended here in the source's storage order (what its `label IN (...)` query returns) under
// fresh sequential labels.  Rows move device to device; only the label lists touch the host.
int Store::merge_from(Store &src, const int64_t *live_labels, int count, int64_t *out_old, int64_t *out_new) {
	if (&src == this) throw Error("cannot merge an index into itself");
	if (src.dim_ != dim_)
		throw Error("dimension mismatch: target " + std::to_string(dim_) + ", source " + std::to_string(src.dim_));
	if (count <= 0 || !live_labels) return 0;
	if (!out_old || !out_new) throw Error("null buffer");
	std::vector<int64_t> rows;
	rows.reserve(count);
	DevBuf tmp; // [m][dim] on the source's device
	DevBuf d_rows;
	int64_t src_base = 0, l = 0;
	{
		std::shared_lock<std::shared_mutex> lk(src.mu_);
		LB_CUDA(cudaSetDevice(src.device_));
		src_base = src.label_base_;
		for (int i = 0; i < count; i++) {
			const int64_t r = src.row_of_label(live_labels[i]);
			if (r < 0 || r >= src.n_rows_ || ((src.deleted_host_[(size_t)(r >> 5)] >> (r & 31)) & 1u)) continue;
			rows.push_back(r);
		}
		std::sort(rows.begin(), rows.end());
		rows.erase(std::unique(rows.begin(), rows.end()), rows.end());
		if (rows.empty()) return 0;
		d_rows.alloc(rows.size() * sizeof(int64_t));
		tmp.alloc(rows.size() * (size_t)dim_ * sizeof(float));
		LB_CUDA(cudaMemcpyAsync(d_rows.p, rows.data(), rows.size() * sizeof(int64_t), cudaMemcpyHostToDevice, src.copy_stream_));
		launch_gather_rows(src.X_.as<float>(), d_rows.as<int64_t>(), (long long)rows.size(), dim_, tmp.as<float>(),
		                   src.copy_stream_, &l);
		LB_CUDA(cudaStreamSynchronize(src.copy_stream_));
		src.add_launches(l);
	}
	const int64_t m = (int64_t)rows.size();
	int64_t first;
	if (src.device_ == device_) {
		first = add_device(tmp.as<float>(), m, copy_stream_);
	} else {
		LB_CUDA(cudaSetDevice(device_));
		DevBuf here;
		here.alloc((size_t)m * dim_ * sizeof(float));
		LB_CUDA(cudaMemcpyPeer(here.p, device_, tmp.p, src.device_, (size_t)m * dim_ * sizeof(float)));
		first = add_device(here.as<float>(), m, copy_stream_);
	}
	if (!src.columns_.empty()) {
		// the rows' extra column values follow them (the reference moves whole Arrow batches, lance_manager.rs:307-345)
		if (src.device_ != device_) throw Error("merging extra columns across devices is not supported");
		std::unique_lock<std::shared_mutex> lk(mu_);
		std::shared_lock<std::shared_mutex> ls(src.mu_);
		LB_CUDA(cudaSetDevice(device_));
		int64_t lc = 0;
		columns_.gather_from(src.columns_, d_rows.as<int64_t>(), m, row_of_label(first), cap_rows_, copy_stream_, &lc);
		add_launches(lc);
	}
	{
		std::shared_lock<std::shared_mutex> ls(src.mu_);
		for (int64_t i = 0; i < m; i++) {
			out_old[i] = src.label_of_row(rows[(size_t)i]);
			out_new[i] = first + i;
		}
	}
	(void)src_base;
	return (int)m;
}

// ------------------------------------------------------------------ search
void Store::search_device(const float *d_queries, int nq, int k, int nprobes, int refine, const uint32_t *d_allow,
                          int64_t allow_bits, int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts,
                          cudaStream_t s, bool defer, uint32_t *d_overflow, HostHints *hints, bool allow_in_row_space) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (nq <= 0) return;
	if (k <= 0) {
		LB_CUDA(cudaMemsetAsync(d_out_counts, 0, (size_t)nq * sizeof(int32_t), s));
		return;
	}
	// scratch is keyed by stream: everything enqueued on one stream is ordered, so back-to-back searches
	// reuse it without a host synchronisation (growing a buffer frees the old one, which synchronises).
	Workspace *ws = scratch_for(s);
	std::lock_guard<std::mutex> wg(ws->mu);
	if (explicit_labels_) {
		// kernels work in ROW space: a caller's bitmap (by label) is translated, results are settled at once (a later
		// search_finish would rewrite re-run queries in row space) and mapped row -> label before the call returns
		defer = false;
		if (d_allow && !allow_in_row_space) {
			const size_t words = (size_t)((n_rows_ + 31) / 32);
			ws->allow_rows.ensure(std::max<size_t>(words, 1) * sizeof(uint32_t));
			launch_allow_by_row(d_allow, allow_bits, row_labels_.as<int64_t>(), n_rows_, ws->allow_rows.as<uint32_t>(), s);
			add_launches(1);
			d_allow = ws->allow_rows.as<uint32_t>();
			allow_bits = n_rows_;
		}
	}
	if (hints && !hints->q_uploaded) {
		// the host call left the query on the host: only the flat single-query scan takes it from its launch parameters
		// ... and the single-query IVF_PQ kernel, which decides for itself (ivfpq_search_locked): it uploads nothing when
		// it takes the query inline.  Conservative here: with an index, upload unless the shape is the fused one.
		const bool inline_ok = nq == 1 && dim_ <= 1024 &&
		                       (index_ ? ivfpq_takes_inline_query(k, nprobes, refine)
		                               : (n_rows_ > 0 && scan1_eligible(X_.as<float>(), dim_, k, 1)));
		if (!inline_ok) {
			LB_CUDA(cudaMemcpyAsync(hints->d_q_buf, hints->h_queries, (size_t)nq * dim_ * sizeof(float), cudaMemcpyHostToDevice, s));
			hints->q_uploaded = true;
		}
	}
	if (index_)
		ivfpq_search_locked(d_queries, nq, k, nprobes, refine, d_allow, allow_bits, d_out_labels, d_out_dist,
		                    d_out_counts, *ws, s, nullptr, nullptr, nullptr, hints);
	else
		flat_search_locked(d_queries, nq, k, d_allow, allow_bits, d_out_labels, d_out_dist, d_out_counts, *ws, s, defer, hints);
	if (explicit_labels_) {
		launch_relabel(d_out_labels, (long long)nq * k, row_labels_.as<int64_t>(), n_rows_, s);
		add_launches(1);
	}
	if (d_overflow) {
		// number of queries whose result is not final until search_finish (0 on every path but a pending tcgen05 batch)
		const uint32_t *t = (defer && flat_gemm_) ? flat_gemm_->pending_total(s) : nullptr;
		if (t)
			LB_CUDA(cudaMemcpyAsync(d_overflow, t, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
		else
			LB_CUDA(cudaMemsetAsync(d_overflow, 0, sizeof(uint32_t), s));
	}
	std::lock_guard<std::mutex> g(stat_mu_);
	searches_++;
}

// Completes an asynchronous device search (search_device with defer = true): waits for its stream and settles the
// candidate-list overflow check.  Returns how many queries had to be re-run by the exact scan.
int Store::search_finish(cudaStream_t s) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	int rerun = 0;
	if (flat_gemm_ && flat_gemm_->has_pending(s)) {
		Workspace *ws = scratch_for(s);
		std::lock_guard<std::mutex> wg(ws->mu);
		int64_t launches = 0, fallback = 0;
		rerun = flat_gemm_->finish(*ws, s, &launches, &fallback);
		std::lock_guard<std::mutex> g(stat_mu_);
		tc_fallback_ += fallback;
		launches_ += launches;
	} else {
		LB_CUDA(cudaStreamSynchronize(s));
	}
	return rerun;
}

void Store::search_host(const float *queries, int nq, int k, int nprobes, int refine, const uint32_t *allow,
                        int64_t allow_bits, int64_t *out_labels, float *out_dist, int32_t *out_counts,
                        const DevBitmap *resident) {
	if (nq <= 0) return;
	if (k <= 0) {
		for (int i = 0; i < nq; i++) out_counts[i] = 0;
		return;
	}
	LB_CUDA(cudaSetDevice(device_));
	auto io = acquire_io();
	try {
		cudaStream_t s = io->stream;
		const size_t qb = (size_t)nq * dim_ * sizeof(float);
		const size_t lb = (size_t)nq * k * sizeof(int64_t), db = (size_t)nq * k * sizeof(float), cb = (size_t)nq * sizeof(int32_t);
		// small calls (single queries, modest batches) go through pinned staging and one packed result buffer
		const bool staged = qb + lb + db + cb <= (1u << 20);
		// zero-copy (LANCE_B200_ZEROCOPY, default 1): the staging buffer is pinned and, under unified addressing,
		// mapped at the same address on the device.  >= 1: the kernels write labels / distances / counts straight
		// into it (posted writes, visible once the stream has drained) instead of a D2H copy (B200: C1 89 -> 81 us,
		// C3 111 -> 104 us per call); 2: tiny query blocks are also read in place instead of an H2D copy (measured
		// slower: every CTA fetches the query over PCIe, C1 98 us).
		static const int zc_mode = [] {
			const char *e = getenv("LANCE_B200_ZEROCOPY");
			return e ? atoi(e) : 1;
		}();
		const bool zc_out = staged && zc_mode >= 1;
		const bool zc_q = zc_out && zc_mode >= 2 && qb <= 8192 && nq < 8;
		const size_t qoff = (qb + 15) & ~(size_t)15; // results start 16-byte aligned behind the queries
		const size_t foff = (lb + db + cb + 15) & ~(size_t)15; // completion word behind the results
		char *hs = staged ? static_cast<char *>(io->stage(qoff + foff + 16)) : nullptr;
		// single query, results written in place: the kernel can also take the query from its launch parameters (no
		// H2D copy at all) and publish a completion word next to the results, which is polled instead of a stream
		// synchronise (LANCE_B200_POLL=0 switches the polling off)
		static const bool poll_mode = [] {
			const char *e = getenv("LANCE_B200_POLL");
			return e ? atoi(e) != 0 : true;
		}();
		HostHints hints;
		// (explicit labels: the result is re-labelled after the search kernel, so no completion word from inside it)
		const bool single = zc_out && nq == 1 && !explicit_labels_;
		if (single) {
			hints.h_queries = queries;
			if (poll_mode) {
				hints.done_flag = reinterpret_cast<volatile uint32_t *>(hs + qoff + foff);
				hints.done_value = ++io->done_seq;
				if (hints.done_value == 0) hints.done_value = ++io->done_seq;
			}
		}
		if (!zc_q) io->q.ensure(qb);
		if (!zc_out) io->out_labels.ensure(lb + db + cb);
		char *res = zc_out ? hs + qoff : static_cast<char *>(io->out_labels.p);
		int64_t *d_labels = reinterpret_cast<int64_t *>(res);
		float *d_dist = reinterpret_cast<float *>(res + lb);
		int32_t *d_counts = reinterpret_cast<int32_t *>(res + lb + db);
		const float *d_q = zc_q ? reinterpret_cast<const float *>(hs) : io->q.as<float>();
		if (single) {
			// the search decides under the store lock whether it can take the query inline; if not it uploads it itself
			hints.d_q_buf = io->q.as<float>();
			hints.q_uploaded = false;
			if (zc_q) { // (LANCE_B200_ZEROCOPY=2) the device reads the query in place from the mapped staging buffer
				std::memcpy(hs, queries, qb);
				hints.q_uploaded = true;
			}
		} else if (staged) {
			std::memcpy(hs, queries, qb);
			if (!zc_q) LB_CUDA(cudaMemcpyAsync(io->q.p, hs, qb, cudaMemcpyHostToDevice, s));
		} else {
			LB_CUDA(cudaMemcpyAsync(io->q.p, queries, qb, cudaMemcpyHostToDevice, s));
		}
		const uint32_t *d_allow = nullptr;
		if (resident) { // evaluated on the device (columns.h): nothing to upload
			d_allow = resident->data();
			allow_bits = resident->bits;
		} else if (allow) {
			size_t words = (size_t)((allow_bits + 31) / 32);
			io->allow.ensure(std::max<size_t>(words, 1) * sizeof(uint32_t));
			LB_CUDA(cudaMemcpyAsync(io->allow.p, allow, words * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
			d_allow = io->allow.as<uint32_t>();
		}
		search_device(d_q, nq, k, nprobes, refine, d_allow, allow_bits, d_labels, d_dist, d_counts, s, false, nullptr,
		              single ? &hints : nullptr, resident != nullptr);
		if (staged) {
			if (!zc_out) LB_CUDA(cudaMemcpyAsync(hs + qoff, d_labels, lb + db + cb, cudaMemcpyDeviceToHost, s));
			if (hints.armed) {
				// poll the completion word (the kernel writes it behind a system-scope fence, after the results); look at
				// the stream now and then so that a faulted launch surfaces as an error instead of a hang
				const volatile uint32_t *flag = hints.done_flag;
				for (uint32_t spins = 0; *flag != hints.done_value; spins++) {
					if ((spins & 0x3FFF) == 0x3FFF) {
						cudaError_t e = cudaStreamQuery(s);
						if (e == cudaSuccess) break; // drained: the word is there (or the fast path did not run)
						if (e != cudaErrorNotReady) LB_CUDA(e);
					}
#if defined(__x86_64__)
					__builtin_ia32_pause();
#endif
				}
			} else {
				LB_CUDA(cudaStreamSynchronize(s));
			}
			std::memcpy(out_labels, hs + qoff, lb);
			std::memcpy(out_dist, hs + qoff + lb, db);
			std::memcpy(out_counts, hs + qoff + lb + db, cb);
		} else {
			LB_CUDA(cudaMemcpyAsync(out_labels, d_labels, lb, cudaMemcpyDeviceToHost, s));
			LB_CUDA(cudaMemcpyAsync(out_dist, d_dist, db, cudaMemcpyDeviceToHost, s));
			LB_CUDA(cudaMemcpyAsync(out_counts, d_counts, cb, cudaMemcpyDeviceToHost, s));
			LB_CUDA(cudaStreamSynchronize(s));
		}
	} catch (...) {
		release_io(std::move(io));
		throw;
	}
	release_io(std::move(io));
}

void Store::flat_search_locked(const float *d_queries, int nq, int k, const uint32_t *d_allow, int64_t allow_bits,
                               int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts, Workspace &ws,
                               cudaStream_t s, bool defer, HostHints *hints) {
	int64_t launches = 0;
	const int mode = flat_path_.load();
	bool use_tc = false;
	if (mode == 2)
		use_tc = true;
	else if (mode == 0)
		use_tc = FlatGemm::worthwhile(nq, n_rows_, dim_, k);
	if (use_tc && !FlatGemm::supported(dim_, metric_, k)) {
		if (mode == 2) throw Error("the tcgen05 flat path needs 16 <= dim (<= 768; 704 for l2) and k <= 1024");
		use_tc = false;
	}
	if (use_tc) {
		{
			std::lock_guard<std::mutex> g(gemm_mu_);
			if (!flat_gemm_) flat_gemm_.reset(new FlatGemm(device_, num_sms_, dim_, metric_));
			flat_gemm_->sync_rows(X_.as<float>(), n_rows_, cap_rows_, s, &launches);
		}
		int64_t fallback = 0;
		flat_gemm_->search(X_.as<float>(), n_rows_, n_deleted_ > 0 ? deleted_.as<uint32_t>() : nullptr, d_queries, nq, k, d_allow, allow_bits,
		                   label_base_, d_out_labels, d_out_dist, d_out_counts, ws, s, &launches, &fallback, &timer_, defer);
		std::lock_guard<std::mutex> g(stat_mu_);
		tc_batches_++;
		tc_fallback_ += fallback;
		launches_ += launches;
		return;
	}
	FlatScanArgs a;
	a.X = X_.as<float>();
	a.n = n_rows_;
	a.d = dim_;
	a.metric = metric_;
	a.dQ = d_queries;
	a.nq = nq;
	a.k = k;
	a.allow = d_allow;
	a.allow_bits = allow_bits;
	a.deleted = n_deleted_ > 0 ? deleted_.as<uint32_t>() : nullptr;
	a.label_base = label_base_;
	a.out_labels = d_out_labels;
	a.out_dist = d_out_dist;
	a.out_counts = d_out_counts;
	a.timer = &timer_;
	if (n_rows_ > 0 && scan1_eligible(a.X, dim_, k, nq)) {
		// one query (what the SQL surface issues): the TMA-staged latency path, one launch, optional inline query and
		// completion word
		if (!ws.scan1_hdr.p) {
			ws.scan1_hdr.alloc(64);
			init_scan1_header(ws.scan1_hdr.as<u64>(), s);
		}
		ws.work.ensure(scan1_scratch_bytes(k, num_sms_));
		a.scan1_hdr = ws.scan1_hdr.as<u64>();
		if (hints) {
			a.hQ = hints->q_uploaded ? nullptr : hints->h_queries;
			a.done_flag = hints->done_flag, a.done_value = hints->done_value;
			hints->armed = hints->done_flag != nullptr;
		}
		run_flat_scan1(a, ws.work.p, num_sms_, s, &launches);
		std::lock_guard<std::mutex> g(stat_mu_);
		simt_batches_++;
		launches_ += launches;
		return;
	}
	FlatScanPlan p = plan_flat_scan(a.X, metric_, dim_, k, nq, n_rows_, num_sms_);
	ws.work.ensure(flat_scan_scratch_bytes(p, k));
	if (!ws.scan_hdr.p) {
		ws.scan_hdr.alloc(SCAN_HDR_BYTES);
		ws.scan_hdr_clean = false;
	}
	a.hdr = ws.scan_hdr.as<u64>();
	a.hdr_clean = &ws.scan_hdr_clean;
	run_flat_scan(a, p, ws.work.p, s, &launches);
	std::lock_guard<std::mutex> g(stat_mu_);
	simt_batches_++;
	launches_ += launches;
}

void Store::stats(lance_b200_stats_t *out) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	std::lock_guard<std::mutex> g(stat_mu_);
	memset(out, 0, sizeof(*out));
	out->kernel_launches = launches_;
	out->searches = searches_;
	out->flat_simt_batches = simt_batches_;
	out->flat_tc_batches = tc_batches_;
	out->tc_fallback_queries = tc_fallback_;
	out->ivfpq_batches = ivfpq_batches_;
	out->rows = n_rows_;
	out->live_rows = n_rows_ - n_deleted_;
	out->device = device_;
	out->has_index = index_ ? 1 : 0;
}

void Store::read_profile(lance_b200_profile_t *out) {
	LB_CUDA(cudaSetDevice(device_));
	LB_CUDA(cudaDeviceSynchronize());
	double ms[KC_COUNT];
	int64_t n[KC_COUNT];
	timer_.read(ms, n);
	out->flat_scan_ms = ms[KC_FLAT_SCAN], out->flat_scan_launches = n[KC_FLAT_SCAN];
	out->flat_gemm_ms = ms[KC_FLAT_GEMM], out->flat_gemm_launches = n[KC_FLAT_GEMM];
	out->adc_scan_ms = ms[KC_ADC_SCAN], out->adc_scan_launches = n[KC_ADC_SCAN];
	out->rerank_ms = ms[KC_RERANK], out->rerank_launches = n[KC_RERANK];
}

// ------------------------------------------------------------------ persistence (own format)
// <db_path>/<table>.b200/store.bin:
//   magic "LB200v2\0" | int64 hdr[8] = {dim, metric, n_rows, n_deleted, label_base, has_index, 0, 0}
//   | tombstone words [(n_rows + 31) / 32] | rows fp32 [n_rows][dim]
//   | (has_index) the IVF_PQ index in the exchange layout of lance_b200_get_ivfpq (save_index / load_index, ivfpq_host.cuh)
// The reference keeps its ANN index across a restart (src/lance_index.cpp:492-587 reopens the Lance dataset with its
// _indices; test/sql/lance_persist.test:20-34): so does this format — an indexed table reopens indexed.
static const char MAGIC[8] = {'L', 'B', '2', '0', '0', 'v', '2', '\0'};
static const char MAGIC_V1[8] = {'L', 'B', '2', '0', '0', 'v', '1', '\0'};

void Store::save() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (dir_.empty()) throw Error("store has no db_path to save under");
	mkdir(db_path_.c_str(), 0777);
	mkdir(dir_.c_str(), 0777);
	std::string path = dir_ + "/store.bin";
	FILE *f = fopen((path + ".tmp").c_str(), "wb");
	if (!f) throw Error("cannot write " + path);
	int64_t hdr[8] = {dim_, metric_, n_rows_, n_deleted_, label_base_, index_ ? 1 : 0, columns_.empty() ? 0 : 1, explicit_labels_ ? 1 : 0};
	bool ok = fwrite(MAGIC, 1, 8, f) == 8 && fwrite(hdr, sizeof(int64_t), 8, f) == 8;
	size_t words = (size_t)((n_rows_ + 31) / 32);
	ok = ok && fwrite(deleted_host_.data(), sizeof(uint32_t), words, f) == words;
	const int64_t chunk = std::max<int64_t>(1, (64ll << 20) / ((int64_t)dim_ * 4));
	std::vector<float> buf((size_t)chunk * dim_);
	for (int64_t r = 0; r < n_rows_ && ok; r += chunk) {
		int64_t c = std::min(chunk, n_rows_ - r);
		LB_CUDA(cudaMemcpy(buf.data(), X_.as<float>() + (size_t)r * dim_, (size_t)c * dim_ * sizeof(float),
		                   cudaMemcpyDeviceToHost));
		ok = fwrite(buf.data(), sizeof(float), (size_t)c * dim_, f) == (size_t)c * dim_;
	}
	if (ok && index_) {
		try {
			save_index(f);
		} catch (...) {
			fclose(f);
			throw;
		}
	}
	if (ok && explicit_labels_) {
		int64_t lh[2] = {next_label_, n_rows_};
		ok = fwrite(lh, sizeof(int64_t), 2, f) == 2 &&
		     fwrite(row_labels_host_.data(), sizeof(int64_t), (size_t)n_rows_, f) == (size_t)n_rows_;
	}
	if (ok && !columns_.empty()) {
		try {
			columns_.save(f, n_rows_, copy_stream_);
		} catch (...) {
			fclose(f);
			throw;
		}
	}
	ok = (fclose(f) == 0) && ok;
	if (!ok || rename((path + ".tmp").c_str(), path.c_str()) != 0) throw Error("short write to " + path);
}

Store *Store::open(const std::string &db_path, const std::string &table, const std::string &metric, int device) {
	std::string dir = db_path + "/" + (table.empty() ? std::string("vectors") : table) + ".b200";
	std::string path = dir + "/store.bin";
	FILE *f = fopen(path.c_str(), "rb");
	if (!f) throw Error("no saved store at " + path);
	char magic[8];
	int64_t hdr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
	bool v1 = false;
	if (fread(magic, 1, 8, f) != 8) {
		fclose(f);
		throw Error("not a lance_b200 store: " + path);
	}
	if (memcmp(magic, MAGIC_V1, 8) == 0) v1 = true;
	else if (memcmp(magic, MAGIC, 8) != 0) {
		fclose(f);
		throw Error("not a lance_b200 store: " + path);
	}
	const size_t nh = v1 ? 6 : 8;
	if (fread(hdr, sizeof(int64_t), nh, f) != nh) {
		fclose(f);
		throw Error("truncated store header: " + path);
	}
	std::unique_ptr<Store> st;
	try {
		const int64_t dim = hdr[0], n = hdr[2];
		if (dim <= 0 || dim > (1 << 20) || n < 0 || n >= (1ll << 32)) throw Error("corrupt store header: " + path);
		// the file must hold what the header promises (a truncated or foreign file is an error, not a short table)
		struct stat sb;
		if (fstat(fileno(f), &sb) != 0) throw Error("cannot stat " + path);
		const size_t words = (size_t)((n + 31) / 32);
		const uint64_t need = 8 + nh * 8 + (uint64_t)words * 4 + (uint64_t)n * (uint64_t)dim * 4;
		if ((uint64_t)sb.st_size < need) throw Error("truncated store: " + path);
		// the caller's metric wins, as in lance_manager.rs:117-161 (open takes the metric from the index options)
		std::string m = !metric.empty() ? metric : (hdr[1] == METRIC_COSINE ? "cosine" : hdr[1] == METRIC_DOT ? "dot" : "l2");
		st.reset(new Store(db_path, table, (int)dim, m, device));
		st->label_base_ = hdr[4];
		st->ensure_capacity(std::max<int64_t>(n, 1));
		if (fread(st->deleted_host_.data(), sizeof(uint32_t), words, f) != words) throw Error("truncated store: " + path);
		if (n & 31) st->deleted_host_[words - 1] &= (1u << (n & 31)) - 1u; // bits beyond the table mean nothing
		LB_CUDA(cudaMemcpy(st->deleted_.p, st->deleted_host_.data(), words * sizeof(uint32_t), cudaMemcpyHostToDevice));
		const int64_t chunk = std::max<int64_t>(1, (64ll << 20) / ((int64_t)st->dim_ * 4));
		std::vector<float> buf((size_t)chunk * st->dim_);
		for (int64_t r = 0; r < n; r += chunk) {
			int64_t c = std::min(chunk, n - r);
			if (fread(buf.data(), sizeof(float), (size_t)c * st->dim_, f) != (size_t)c * st->dim_)
				throw Error("truncated store: " + path);
			LB_CUDA(cudaMemcpy(st->X_.as<float>() + (size_t)r * st->dim_, buf.data(), (size_t)c * st->dim_ * sizeof(float),
			                   cudaMemcpyHostToDevice));
		}
		st->n_rows_ = n;
		// the tombstone count is what the bitmap says, not what the header claims
		int64_t nd = 0;
		for (size_t w = 0; w < words; w++) nd += __builtin_popcount(st->deleted_host_[w]);
		st->n_deleted_ = nd;
		st->next_label_ = st->label_base_ + n;
		if (!v1 && hdr[7] != 0) st->explicit_labels_ = true; // the index installs in row space: decide before it loads
		if (st->explicit_labels_) st->label_base_ = 0;
		if (!v1 && hdr[5] != 0) st->load_index(f, path);
		if (st->explicit_labels_) {
			int64_t lh[2];
			if (fread(lh, sizeof(int64_t), 2, f) != 2 || lh[1] != n) throw Error("corrupt label section: " + path);
			st->row_labels_host_.resize((size_t)n);
			if (fread(st->row_labels_host_.data(), sizeof(int64_t), (size_t)n, f) != (size_t)n) throw Error("truncated store: " + path);
			for (int64_t i = 1; i < n; i++)
				if (st->row_labels_host_[(size_t)i] <= st->row_labels_host_[(size_t)i - 1]) throw Error("corrupt label section: " + path);
			if (n > 0 && lh[0] <= st->row_labels_host_[(size_t)n - 1]) throw Error("corrupt label section: " + path);
			st->next_label_ = lh[0];
			st->row_labels_.alloc((size_t)st->cap_rows_ * sizeof(int64_t));
			if (n > 0)
				LB_CUDA(cudaMemcpy(st->row_labels_.p, st->row_labels_host_.data(), (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice));
		}
		if (!v1 && hdr[6] != 0) st->columns_.load(f, path, n, st->cap_rows_, st->copy_stream_);
	} catch (...) {
		fclose(f);
		throw;
	}
	fclose(f);
	return st.release();
}

} // namespace lb
