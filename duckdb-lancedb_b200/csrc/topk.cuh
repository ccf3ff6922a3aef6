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

} // namespace lb
