// topk.cuh — warp-private streaming k-selection on 64-bit (distance,row) keys.
//
// Every scan kernel of this library (flat fp32 scan, ADC scan, candidate re-rank, partial-list merge)
// ends in the same fused selection, so that no distance array is ever written to HBM:
//   * each warp owns a shared-memory buffer of `cap` keys (cap = power of two >= k + 32) and a
//     warp-uniform threshold `tau` (the k-th best key seen so far, KEY_INVALID until k were seen);
//   * push(): lanes whose key beats tau append through a ballot/popc prefix (no atomics);
//   * when the buffer would overflow the warp bitonic-sorts it, keeps the k best and tightens tau.
// After the first few compactions almost nothing passes tau (expected appends ~ k ln(n/k)), so the
// selection costs one compare + ballot per batch of 32 candidates.
#pragma once
#include "common.cuh"

namespace lb {

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }

// smallest legal buffer capacity for a given k
__host__ __device__ inline int warpsel_cap(int k) {
	int c = ceil_pow2(k + 32);
	return c < 64 ? 64 : c;
}

// Ascending bitonic sort of buf[0..cap) by one warp (cap = power of two >= 64).
__device__ inline void warp_bitonic_sort(u64 *buf, int cap) {
	const unsigned lane = lane_id();
	for (int size = 2; size <= cap; size <<= 1) {
		for (int stride = size >> 1; stride > 0; stride >>= 1) {
			for (int i = lane; i < (cap >> 1); i += 32) {
				int pos = 2 * i - (i & (stride - 1));
				int j = pos + stride;
				bool up = ((pos & size) == 0);
				u64 a = buf[pos], b = buf[j];
				if ((a > b) == up) {
					buf[pos] = b;
					buf[j] = a;
				}
			}
			__syncwarp();
		}
	}
}

struct WarpSelect {
	u64 *buf; // shared memory, `cap` entries, private to this warp
	int cap;
	int k;
	int cnt; // warp-uniform
	u64 tau; // warp-uniform: a key is a candidate iff key < tau

	__device__ __forceinline__ void init(u64 *smem, int cap_, int k_) {
		buf = smem;
		cap = cap_;
		k = k_;
		cnt = 0;
		tau = KEY_INVALID;
	}

	// Sort what is buffered, keep the k best, tighten tau.  Afterwards buf[0..cnt) is ascending.
	__device__ inline void compact() {
		const unsigned lane = lane_id();
		__syncwarp();
		for (int i = cnt + lane; i < cap; i += 32) buf[i] = KEY_INVALID;
		__syncwarp();
		warp_bitonic_sort(buf, cap);
		if (cnt > k) cnt = k;
		if (cnt == k) {
			u64 t = buf[k - 1];
			if (t < tau) tau = t;
		}
	}

	// All 32 lanes must call.  `valid` lanes offer `key`.
	__device__ __forceinline__ void push(bool valid, u64 key) {
		bool want = valid && key < tau;
		unsigned m = __ballot_sync(0xFFFFFFFFu, want);
		if (m == 0) return;
		int n = __popc(m);
		if (cnt + n > cap) {
			compact();
			want = valid && key < tau;
			m = __ballot_sync(0xFFFFFFFFu, want);
			if (m == 0) return;
			n = __popc(m);
		}
		if (want) buf[cnt + __popc(m & ((1u << lane_id()) - 1u))] = key;
		cnt += n;
	}

	// Tighten with a bound learnt elsewhere (another warp/block's k-th best): still exact, because
	// any list's k-th best is an upper bound of the global k-th best.
	__device__ __forceinline__ void tighten(u64 bound) {
		if (bound < tau) tau = bound;
	}

	__device__ inline void finalize() { compact(); }

	// Leaves the buffered keys as they are (unsorted, possibly more than k) and fills the rest of the buffer with
	// KEY_INVALID, so that a block-wide sort over all the warps' buffers can take over.
	__device__ inline void pad() {
		__syncwarp();
		for (int i = cnt + lane_id(); i < cap; i += 32) buf[i] = KEY_INVALID;
		__syncwarp();
	}
};

// Ascending bitonic sort of n keys (n = power of two) by the whole block; key i lives at base[map(i)].  All threads
// call it.  Replaces "one warp folds the other warps' lists in" at the end of every scan: with 8 warps x cap keys
// that fold was the longest serial stretch of the small-table scans (ncu: 1/3 of the C1 kernel spent at its barrier).
struct IdentityMap {
	__device__ __forceinline__ int operator()(int i) const { return i; }
};
template <class Map>
__device__ inline void block_bitonic_sort(u64 *base, int n, Map map) {
	__syncthreads();
	// Pair i of a stage touches positions 2i - (i & (stride-1)) and + stride: for stride <= 32 the 32 pairs of a warp
	// stay inside the warp's own 64 positions, so those stages need a warp barrier only; a block barrier is needed
	// after a stage with stride > 32 and before the next one that has (45 stages of a 512-key sort: 10 block barriers).
	for (int size = 2; size <= n; size <<= 1) {
		for (int stride = size >> 1; stride > 0; stride >>= 1) {
			for (int i = threadIdx.x; i < (n >> 1); i += blockDim.x) {
				const int pos = 2 * i - (i & (stride - 1));
				const int pa = map(pos), pb = map(pos + stride);
				const bool up = ((pos & size) == 0);
				const u64 a = base[pa], b = base[pb];
				if ((a > b) == up) {
					base[pa] = b;
					base[pb] = a;
				}
			}
			const int next_stride = stride > 1 ? (stride >> 1) : size; // first stride of the next size
			if (stride > 32 || next_stride > 32)
				__syncthreads();
			else
				__syncwarp();
		}
	}
	__syncthreads();
}

// Final select by a whole block: src[0..m) keys (KEY_INVALID entries ignored, read through L2 because other
// SMs may have written them) -> sorted (label, distance) list of query slot `q`.  All threads of the block
// call it; bufs = nwarps*cap keys of shared memory (cnts is unused, kept for the callers' layouts).
__device__ inline void block_final_select(const u64 *__restrict__ src, long long m, int k, int cap, u64 *bufs, int *cnts,
                                          long long label_base, size_t q, int64_t *__restrict__ out_labels,
                                          float *__restrict__ out_dist, int32_t *__restrict__ out_counts,
                                          u64 *__restrict__ out_keys) {
	(void)cnts;
	const int nwarps = blockDim.x >> 5;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	if (m <= (long long)nwarps * cap) {
		// short list (the usual case for the in-kernel two-level selects): straight into shared memory, one sort
		int n = 64;
		while (n < (int)m || n < k) n <<= 1;
		for (int i = threadIdx.x; i < n; i += blockDim.x) bufs[i] = i < (int)m ? __ldcg(src + i) : KEY_INVALID;
		block_bitonic_sort(bufs, n, IdentityMap());
	} else {
		WarpSelect sel;
		sel.init(bufs + (size_t)warp * cap, cap, k);
		for (long long i0 = (long long)warp * 128; i0 < m; i0 += (long long)nwarps * 128) {
			u64 key[4];
#pragma unroll
			for (int u = 0; u < 4; u++) {
				long long i = i0 + u * 32 + lane;
				key[u] = (i < m) ? __ldcg(src + i) : KEY_INVALID;
			}
#pragma unroll
			for (int u = 0; u < 4; u++) sel.push(key[u] != KEY_INVALID, key[u]);
		}
		sel.pad();
		block_bitonic_sort(bufs, nwarps * cap, IdentityMap());
	}
	if (warp == 0) {
		int cnt = 0;
		for (int i0 = 0; i0 < k; i0 += 32) {
			const int i = i0 + lane;
			const bool ok = i < k && bufs[i] != KEY_INVALID;
			const u64 key = ok ? bufs[i] : KEY_INVALID;
			cnt += __popc(__ballot_sync(0xFFFFFFFFu, ok));
			if (i < k) {
				if (out_labels) out_labels[q * k + i] = ok ? (int64_t)(label_base + (long long)key_row(key)) : -1;
				if (out_dist) out_dist[q * k + i] = ok ? key_dist(key) : __int_as_float(0x7F800000);
				if (out_keys) out_keys[q * k + i] = key;
			}
		}
		if (lane == 0 && out_counts) out_counts[q] = cnt;
	}
	__syncthreads();
}

} // namespace lb
