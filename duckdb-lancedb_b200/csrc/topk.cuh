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
};

// Final select by a whole block: src[0..m) keys (KEY_INVALID entries ignored, read through L2 because other
// SMs may have written them) -> sorted (label, distance) list of query slot `q`.  All threads of the block
// call it; bufs = nwarps*cap keys and cnts = nwarps ints of shared memory.
__device__ inline void block_final_select(const u64 *__restrict__ src, long long m, int k, int cap, u64 *bufs, int *cnts,
                                          long long label_base, size_t q, int64_t *__restrict__ out_labels,
                                          float *__restrict__ out_dist, int32_t *__restrict__ out_counts,
                                          u64 *__restrict__ out_keys) {
	const int nwarps = blockDim.x >> 5;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
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
	sel.finalize();
	if (lane == 0) cnts[warp] = sel.cnt;
	__syncthreads();
	if (warp == 0) {
		for (int w = 1; w < nwarps; w++) {
			const u64 *s2 = bufs + (size_t)w * cap;
			const int c = cnts[w];
			for (int i0 = 0; i0 < c; i0 += 32) {
				int i = i0 + lane;
				u64 key = (i < c) ? s2[i] : KEY_INVALID;
				sel.push(i < c, key);
			}
		}
		sel.finalize();
		for (int i = lane; i < k; i += 32) {
			bool ok = i < sel.cnt;
			u64 key = ok ? sel.buf[i] : KEY_INVALID;
			if (out_labels) out_labels[q * k + i] = ok ? (int64_t)(label_base + (long long)key_row(key)) : -1;
			if (out_dist) out_dist[q * k + i] = ok ? key_dist(key) : __int_as_float(0x7F800000);
			if (out_keys) out_keys[q * k + i] = key;
		}
		if (lane == 0 && out_counts) out_counts[q] = sel.cnt;
	}
	__syncthreads();
}

} // namespace lb
