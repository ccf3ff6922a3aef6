// store_labels.cu — explicit labels (store.h "Labels"), caller-chosen labels for sharded handles, and compaction.
//
// Reference: labels come from an atomic counter and are a COLUMN of the Lance table (rust_lib/src/lance_manager.rs:214,
// 232), so they survive `compact_files` (lance_manager.rs:553-566, called from LanceIndex::Vacuum,
// src/lance_index.cpp:728-751).  The device store keeps labels implicit (label = base + row) until a compaction or a
// label gap forces them to be stored.
#include <algorithm>
#include <numeric>

#include "flat_gemm.h"
#include "ivfpq.h"
#include "store.h"

namespace lb {

__global__ void relabel_kernel(int64_t *__restrict__ labels, long long n, const int64_t *__restrict__ row_labels,
                               long long n_rows) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const long long r = labels[i];
	if (r >= 0 && r < n_rows) labels[i] = row_labels[r];
}

void launch_relabel(int64_t *labels, long long n, const int64_t *row_labels, long long n_rows, cudaStream_t s) {
	if (n <= 0) return;
	relabel_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(labels, n, row_labels, n_rows);
	LB_CUDA(cudaGetLastError());
}

// out bit r = allow bit label(r): one warp per output word
__global__ void allow_by_row_kernel(const uint32_t *__restrict__ allow, long long allow_bits,
                                    const int64_t *__restrict__ row_labels, long long n_rows, uint32_t *__restrict__ out) {
	const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const long long words = (n_rows + 31) / 32;
	if (w >= words) return;
	const long long r = w * 32 + (threadIdx.x & 31);
	bool ok = false;
	if (r < n_rows) {
		const long long l = row_labels[r];
		ok = l >= 0 && l < allow_bits && ((allow[l >> 5] >> (l & 31)) & 1u);
	}
	const uint32_t word = __ballot_sync(0xffffffffu, ok);
	if ((threadIdx.x & 31) == 0) out[w] = word;
}

void launch_allow_by_row(const uint32_t *allow_by_label, long long allow_bits, const int64_t *row_labels, long long n_rows,
                         uint32_t *out, cudaStream_t s) {
	const long long words = (n_rows + 31) / 32;
	if (words <= 0) return;
	allow_by_row_kernel<<<(unsigned)((words + 7) / 8), 256, 0, s>>>(allow_by_label, allow_bits, row_labels, n_rows, out);
	LB_CUDA(cudaGetLastError());
}

int64_t Store::row_of_label(int64_t label) const {
	if (!explicit_labels_) {
		const int64_t r = label - label_base_;
		return (r >= 0 && r < n_rows_) ? r : -1;
	}
	auto it = std::lower_bound(row_labels_host_.begin(), row_labels_host_.begin() + n_rows_, label);
	if (it == row_labels_host_.begin() + n_rows_ || *it != label) return -1;
	return (int64_t)(it - row_labels_host_.begin());
}

bool Store::labels_explicit() {
	std::shared_lock<std::shared_mutex> lk(mu_);
	return explicit_labels_;
}

void Store::force_explicit_labels() {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	make_labels_explicit();
}

void Store::make_labels_explicit() {
	if (explicit_labels_) return;
	if (index_ && label_base_ != 0) throw Error("cannot switch an indexed store with a label base to explicit labels");
	row_labels_host_.resize((size_t)n_rows_);
	std::iota(row_labels_host_.begin(), row_labels_host_.end(), label_base_);
	row_labels_.alloc((size_t)std::max<int64_t>(cap_rows_, 1) * sizeof(int64_t));
	if (n_rows_ > 0)
		LB_CUDA(cudaMemcpy(row_labels_.p, row_labels_host_.data(), (size_t)n_rows_ * sizeof(int64_t), cudaMemcpyHostToDevice));
	explicit_labels_ = true;
	label_base_ = 0; // from here on the kernels' "labels" are row numbers
	if (flat_gemm_) flat_gemm_->invalidate();
}

void Store::append_labels(int64_t first, int64_t num) {
	if (explicit_labels_) {
		row_labels_host_.resize((size_t)(n_rows_ + num));
		std::iota(row_labels_host_.begin() + n_rows_, row_labels_host_.end(), first);
		LB_CUDA(cudaMemcpy(row_labels_.as<int64_t>() + n_rows_, row_labels_host_.data() + n_rows_, (size_t)num * sizeof(int64_t),
		                   cudaMemcpyHostToDevice));
	}
	next_label_ = first + num;
}

void Store::add_host_labeled(const float *vectors, int64_t num, int64_t first_label) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (num <= 0) return;
	if (n_rows_ == 0 && !explicit_labels_ && !index_) {
		label_base_ = first_label; // an empty store simply starts there: labels stay implicit
		next_label_ = first_label;
	}
	if (first_label < next_label_) throw Error("labels must ascend: " + std::to_string(first_label) + " was already handed out");
	if (first_label != next_label_) make_labels_explicit();
	ensure_capacity(n_rows_ + num);
	LB_CUDA(cudaMemcpyAsync(X_.as<float>() + (size_t)n_rows_ * dim_, vectors, (size_t)num * dim_ * sizeof(float),
	                        cudaMemcpyHostToDevice, copy_stream_));
	LB_CUDA(cudaStreamSynchronize(copy_stream_));
	append_labels(first_label, num);
	n_rows_ += num;
}

void Store::add_device_with_labels(const float *d_vectors, int64_t num, const int64_t *labels, cudaStream_t s) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (num <= 0) return;
	for (int64_t i = 0; i < num; i++)
		if (labels[i] < (i ? labels[i - 1] + 1 : next_label_))
			throw Error("labels must ascend strictly and start at or above " + std::to_string(next_label_));
	make_labels_explicit();
	ensure_capacity(n_rows_ + num);
	LB_CUDA(cudaMemcpyAsync(X_.as<float>() + (size_t)n_rows_ * dim_, d_vectors, (size_t)num * dim_ * sizeof(float),
	                        cudaMemcpyDeviceToDevice, s));
	LB_CUDA(cudaStreamSynchronize(s));
	row_labels_host_.resize((size_t)(n_rows_ + num));
	std::copy(labels, labels + num, row_labels_host_.begin() + n_rows_);
	LB_CUDA(cudaMemcpy(row_labels_.as<int64_t>() + n_rows_, labels, (size_t)num * sizeof(int64_t), cudaMemcpyHostToDevice));
	next_label_ = labels[num - 1] + 1;
	n_rows_ += num;
}

// ------------------------------------------------------------------------------------------------ compaction
int64_t Store::compact() {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (n_deleted_ == 0) return 0;
	LB_CUDA(cudaDeviceSynchronize()); // no search may still be reading the buffers that are about to be replaced
	const int64_t old_n = n_rows_;
	std::vector<int64_t> live;          // old rows that stay, ascending
	std::vector<int64_t> new_row_of((size_t)old_n, -1);
	live.reserve((size_t)(old_n - n_deleted_));
	for (int64_t r = 0; r < old_n; r++)
		if (!((deleted_host_[(size_t)(r >> 5)] >> (r & 31)) & 1u)) {
			new_row_of[(size_t)r] = (int64_t)live.size();
			live.push_back(r);
		}
	const int64_t m = (int64_t)live.size();

	// the index in its exchange layout, filtered and renumbered (ivfpq_host.cuh); installed again below
	CompactedIndex cix;
	if (index_) extract_index_for_compaction(new_row_of, cix);

	std::vector<int64_t> labels((size_t)m);
	for (int64_t i = 0; i < m; i++) labels[(size_t)i] = label_of_row(live[(size_t)i]);

	// gather what stays into fresh buffers (device to device), then swap
	const int64_t ncap = (std::max<int64_t>(m, 4096) + 255) & ~255ll;
	DevBuf d_live, nx, nd, nl;
	int64_t l = 0;
	nx.alloc((size_t)ncap * dim_ * sizeof(float));
	nd.alloc((size_t)(ncap / 32) * sizeof(uint32_t));
	nl.alloc((size_t)ncap * sizeof(int64_t));
	LB_CUDA(cudaMemsetAsync(nd.p, 0, nd.bytes, copy_stream_));
	if (m > 0) {
		d_live.alloc((size_t)m * sizeof(int64_t));
		LB_CUDA(cudaMemcpyAsync(d_live.p, live.data(), (size_t)m * sizeof(int64_t), cudaMemcpyHostToDevice, copy_stream_));
		launch_gather_rows(X_.as<float>(), d_live.as<int64_t>(), m, dim_, nx.as<float>(), copy_stream_, &l);
		LB_CUDA(cudaMemcpyAsync(nl.p, labels.data(), (size_t)m * sizeof(int64_t), cudaMemcpyHostToDevice, copy_stream_));
	}
	LB_CUDA(cudaStreamSynchronize(copy_stream_));
	if (!columns_.empty()) {
		ColumnStore fresh;
		fresh.define_like(columns_);
		if (m > 0) fresh.gather_from(columns_, d_live.as<int64_t>(), m, 0, ncap, copy_stream_, &l);
		fresh.grow(ncap, copy_stream_);
		columns_.swap(fresh);
	}
	std::swap(X_.p, nx.p), std::swap(X_.bytes, nx.bytes);
	std::swap(deleted_.p, nd.p), std::swap(deleted_.bytes, nd.bytes);
	std::swap(row_labels_.p, nl.p), std::swap(row_labels_.bytes, nl.bytes);
	row_labels_host_.swap(labels);
	explicit_labels_ = true;
	label_base_ = 0;
	cap_rows_ = ncap;
	n_rows_ = m;
	n_deleted_ = 0;
	deleted_host_.assign((size_t)(ncap / 32), 0u);
	if (flat_gemm_) flat_gemm_->invalidate();
	index_.reset();
	if (cix.present)
		install_index_locked(cix.P, cix.M, cix.n_indexed, cix.centroids.data(), cix.codebook.data(), cix.part_offsets.data(),
		                     cix.row_ids.data(), cix.codes.data());
	add_launches(l);
	return old_n - m;
}

} // namespace lb
