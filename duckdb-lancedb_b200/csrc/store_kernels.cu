// store_kernels.cu — small bookkeeping kernels of the device store: tombstones, list merge plumbing.
#include "kernels.h"
#include "topk.cuh"

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

// dst[i][:] = X[rows[i]][:] — one warp per row (index merge: live rows of the source handle, lance_manager.rs:307-345)
__global__ void gather_rows_kernel(const float *__restrict__ X, const int64_t *__restrict__ rows, long long m, int d,
                                   float *__restrict__ dst) {
	const long long i = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (i >= m) return;
	const float *src = X + (size_t)rows[i] * d;
	float *out = dst + (size_t)i * d;
	for (int c = threadIdx.x & 31; c < d; c += 32) out[c] = src[c];
}

void launch_gather_rows(const float *X, const int64_t *d_rows, long long m, int d, float *dst, cudaStream_t s,
                        int64_t *launches) {
	if (m <= 0) return;
	gather_rows_kernel<<<(unsigned)((m + 7) / 8), 256, 0, s>>>(X, d_rows, m, d, dst);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

// Reads `bytes` of device memory once (clean lines): the second half of an L2 flush — after a buffer larger than L2
// was written, this pass replaces its dirty lines, so the next kernel neither hits its table in L2 nor pays write-backs.
__global__ void read_pass_kernel(const uint4 *__restrict__ p, size_t n16, unsigned *sink) {
	unsigned acc = 0;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
		const uint4 v = __ldcg(p + i);
		acc ^= v.x ^ v.y ^ v.z ^ v.w;
	}
	if (acc == 0x9E3779B9u) *sink = acc; // never true in practice: keeps the loads alive
}

void launch_read_pass(const void *p, size_t bytes, unsigned *sink, cudaStream_t s) {
	read_pass_kernel<<<1184, 256, 0, s>>>(reinterpret_cast<const uint4 *>(p), bytes / 16, sink);
	LB_CUDA(cudaGetLastError());
}

// Final select over G per-shard lists in ONE launch, one block per query.  labels/dists/counts of shard g start
// at labels + g*ls, dists + g*ds, counts + g*cs (element strides), so both the [G][nq][k] layout and a gathered
// buffer of per-rank packed records work.  Keys carry the index g*k+i in their low word.
__global__ void __launch_bounds__(256) merge_lists_kernel(const int64_t *__restrict__ labels, const float *__restrict__ dists,
                                                           const int32_t *__restrict__ counts, long long ls, long long ds,
                                                           long long cs, int G, int nq, int k, int cap,
                                                           int64_t *__restrict__ out_labels, float *__restrict__ out_dists,
                                                           int32_t *__restrict__ out_counts) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	int *cnts = reinterpret_cast<int *>(smem_raw);
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw + 64);
	const int nwarps = blockDim.x >> 5, warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int q = blockIdx.x;
	const int m = G * k;
	WarpSelect sel;
	sel.init(bufs + (size_t)warp * cap, cap, k);
	for (int i0 = warp * 32; i0 < m; i0 += nwarps * 32) {
		const int i = i0 + lane;
		bool ok = false;
		u64 key = KEY_INVALID;
		if (i < m) {
			const int g = i / k, j = i - g * k;
			ok = j < counts[(size_t)g * cs + q];
			if (ok) key = make_key(dists[(size_t)g * ds + (size_t)q * k + j], (uint32_t)i);
		}
		sel.push(ok, key);
	}
	sel.finalize();
	if (lane == 0) cnts[warp] = sel.cnt;
	__syncthreads();
	if (warp != 0) return;
	for (int w = 1; w < nwarps; w++) {
		const u64 *s2 = bufs + (size_t)w * cap;
		const int c = cnts[w];
		for (int i0 = 0; i0 < c; i0 += 32) {
			const int i = i0 + lane;
			sel.push(i < c, i < c ? s2[i] : KEY_INVALID);
		}
	}
	sel.finalize();
	for (int i = lane; i < k; i += 32) {
		const bool ok = i < sel.cnt;
		int64_t lab = -1;
		float dd = __int_as_float(0x7F800000);
		if (ok) {
			const u64 key = sel.buf[i];
			const int idx = (int)key_row(key), g = idx / k, j = idx - g * k;
			lab = labels[(size_t)g * ls + (size_t)q * k + j];
			dd = key_dist(key);
		}
		out_labels[(size_t)q * k + i] = lab;
		out_dists[(size_t)q * k + i] = dd;
	}
	if (lane == 0) out_counts[q] = sel.cnt;
}

void launch_merge_lists(const int64_t *labels, const float *dists, const int32_t *counts, long long ls, long long ds,
                        long long cs, int G, int nq, int k, int64_t *out_labels, float *out_dists, int32_t *out_counts,
                        cudaStream_t s, int64_t *launches) {
	if (nq <= 0 || k <= 0) return;
	const int cap = warpsel_cap(k);
	int nwarps = 8;
	while (nwarps > 1 && (nwarps / 2) * 32 >= G * k) nwarps >>= 1;
	while (nwarps > 1 && 64 + (size_t)nwarps * cap * 8 > 200 * 1024) nwarps >>= 1;
	const size_t smem = 64 + (size_t)nwarps * cap * 8;
	if (smem > 200 * 1024) throw Error("k too large for the device merge");
	if (smem > 48 * 1024)
		LB_CUDA(cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	merge_lists_kernel<<<nq, nwarps * 32, smem, s>>>(labels, dists, counts, ls, ds, cs, G, nq, k, cap, out_labels, out_dists,
	                                                  out_counts);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

} // namespace lb
