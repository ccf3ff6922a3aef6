// store_kernels.cu — small bookkeeping kernels of the device store: tombstones, list merge plumbing.
#include "kernels.h"

namespace lb {

// lance_detached_delete_batch (rust_lib/src/lance_manager.rs:461-471: `label IN (...)` delete):
// tombstone bits, applied as a prefilter by every scan.
__global__ void set_bits_kernel(uint32_t *bitmap, const int64_t *rows, int count, long long nrows) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	long long r = rows[i];
	if (r < 0 || r >= nrows) return;
	atomicOr(bitmap + (r >> 5), 1u << (r & 31));
}

void launch_set_bits(uint32_t *bitmap, const int64_t *d_rows, int count, long long nrows, cudaStream_t s,
                     int64_t *launches) {
	if (count <= 0) return;
	set_bits_kernel<<<(count + 255) / 256, 256, 0, s>>>(bitmap, d_rows, count, nrows);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

__global__ void lists_to_keys_kernel(const float *dists, const int32_t *counts, int G, int nq, int k, u64 *keys) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	long long total = (long long)nq * G * k;
	if (t >= total) return;
	int i = (int)(t % k);
	int g = (int)((t / k) % G);
	int q = (int)(t / ((long long)k * G));
	bool ok = i < counts[(size_t)g * nq + q];
	float d = dists[((size_t)g * nq + q) * k + i];
	keys[t] = ok ? make_key(d, (uint32_t)(g * k + i)) : KEY_INVALID;
}

__global__ void gather_labels_kernel(const u64 *sel, const int64_t *labels, int G, int nq, int k, int64_t *out_labels) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (long long)nq * k) return;
	int q = (int)(t / k);
	u64 key = sel[t];
	if (key == KEY_INVALID) {
		out_labels[t] = -1;
		return;
	}
	uint32_t idx = key_row(key);
	int g = idx / k, i = idx % k;
	out_labels[t] = labels[((size_t)g * nq + q) * k + i];
}

void launch_merge_lists(const int64_t *labels, const float *dists, const int32_t *counts, int G, int nq, int k,
                        u64 *scratch_keys, u64 *scratch_sel, int64_t *out_labels, float *out_dists,
                        int32_t *out_counts, cudaStream_t s, int64_t *launches) {
	if (nq <= 0 || k <= 0) return;
	long long total = (long long)nq * G * k;
	lists_to_keys_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(dists, counts, G, nq, k, scratch_keys);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
	launch_topk_merge(scratch_keys, nq, (long long)G * k, k, 0, nullptr, out_dists, out_counts, scratch_sel, s, launches);
	long long outn = (long long)nq * k;
	gather_labels_kernel<<<(unsigned)((outn + 255) / 256), 256, 0, s>>>(scratch_sel, labels, G, nq, k, out_labels);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

} // namespace lb
