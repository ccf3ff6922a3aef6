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

// k smallest of m keys by the whole block with an MSB-first radix select (8-bit digits, shared-memory histogram):
// O(m / blockDim) independent loads per pass and 2-4 passes for distance keys, instead of streaming every key through
// the warp selectors (whose compaction sorts serialise) — the tails of the single-query kernels and the threshold
// refresh of large-k batches are latency-bound on exactly this step.  key_at(i) returns key i (i < m); the winners
// land UNSORTED in out[0..k) (KEY_INVALID-padded when fewer than k valid keys exist), `out` must hold k + 32 keys.
// All threads call it; m >= k.
template <class KeyAt>
__device__ inline void block_radix_select(KeyAt key_at, int m, int k, u64 *out) {
	__shared__ int hist[256];
	__shared__ int s_need, s_digit, s_bin, s_nout, s_nbin;
	u64 prefix = 0;
	int bits = 0, need = k;
	__syncthreads();
	while (true) {
		for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
		__syncthreads();
		// four keys per thread and round: the loads (L2 round trips when key_at reads global memory) overlap
		for (int i0 = 0; i0 < m; i0 += 4 * (int)blockDim.x) {
			u64 key[4];
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const int i = i0 + u * (int)blockDim.x + (int)threadIdx.x;
				key[u] = i < m ? key_at(i) : 0ull;
			}
#pragma unroll
			for (int u = 0; u < 4; u++) {
				const int i = i0 + u * (int)blockDim.x + (int)threadIdx.x;
				if (i0 + u * (int)blockDim.x >= m) break; // uniform
				int digit = -1 - (int)(threadIdx.x & 31); // lanes without a key match nobody
				if (i < m && (bits == 0 || (key[u] >> (64 - bits)) == prefix)) digit = (int)((key[u] >> (56 - bits)) & 255u);
				// distance keys share their leading digits: one shared-memory atomic per distinct digit of the warp
				const unsigned peers = __match_any_sync(0xFFFFFFFFu, digit);
				if (digit >= 0 && (int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[digit], __popc(peers));
			}
		}
		__syncthreads();
		if (threadIdx.x < 32) {
			// warp 0: smallest digit whose cumulative count reaches `need`
			const int lane = threadIdx.x;
			int c[8], sum = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) c[j] = hist[lane * 8 + j], sum += c[j];
			int incl = sum;
			for (int o = 1; o < 32; o <<= 1) {
				int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
				if (lane >= o) incl += v;
			}
			int before = incl - sum; // keys in digits below this lane's 8
			const bool mine = before < need && incl >= need;
			if (mine) {
				int j = 0;
				while (before + c[j] < need) before += c[j], j++;
				s_digit = lane * 8 + j;
				s_need = need - before;
				s_bin = c[j];
			}
		}
		__syncthreads();
		prefix = (prefix << 8) | (u64)s_digit;
		bits += 8;
		need = s_need;
		const int bin = s_bin;
		__syncthreads();
		if (bin <= 32 || bits == 64) break;
	}
	// winners: every key below the prefix, plus the `need` smallest of the (<= 32, or all-equal) keys on it
	if (threadIdx.x == 0) s_nout = 0, s_nbin = 0;
	for (int i = threadIdx.x; i < k; i += blockDim.x) out[i] = KEY_INVALID; // fewer than k valid keys: padding
	__syncthreads();
	u64 *binl = out + k;
	for (int i0 = 0; i0 < m; i0 += 4 * (int)blockDim.x) {
		u64 key[4];
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const int i = i0 + u * (int)blockDim.x + (int)threadIdx.x;
			key[u] = i < m ? key_at(i) : KEY_INVALID;
		}
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const int i = i0 + u * (int)blockDim.x + (int)threadIdx.x;
			if (i >= m) continue;
			const u64 top = key[u] >> (64 - bits);
			if (top < prefix)
				out[atomicAdd(&s_nout, 1)] = key[u];
			else if (top == prefix) {
				const int pos = atomicAdd(&s_nbin, 1);
				if (pos < 32) binl[pos] = key[u];
			}
		}
	}
	__syncthreads();
	if (threadIdx.x < 32) {
		// one warp: sort the bin (<= 32 keys) in registers, append its `need` smallest
		const int lane = threadIdx.x, nb = min(s_nbin, 32);
		u64 v = lane < nb ? binl[lane] : KEY_INVALID;
		for (int size = 2; size <= 32; size <<= 1)
			for (int stride = size >> 1; stride > 0; stride >>= 1) {
				const u64 o = __shfl_xor_sync(0xFFFFFFFFu, v, stride);
				const bool up = (lane & size) == 0, lower = (lane & stride) == 0;
				v = (lower == up) ? (v < o ? v : o) : (v < o ? o : v);
			}
		const int base = s_nout;
		__syncwarp();
		if (lane < need) out[base + lane] = v;
	}
	__syncthreads();
}

// Same selection for m <= 16 * blockDim.x with the keys held in registers: every key is loaded ONCE (16 independent
// loads per thread) instead of once per radix pass — the passes of a one-block select over a few thousand keys in
// global memory are otherwise a chain of L2 round trips (IVF probe select: 4 096 centroid keys -> nprobes).
template <class KeyAt>
__device__ inline void block_radix_select_regs(KeyAt key_at, int m, int k, u64 *out) {
	__shared__ int hist[256];
	__shared__ int s_need, s_digit, s_bin, s_nout, s_nbin;
	const int nt = (int)blockDim.x, tid = (int)threadIdx.x;
	u64 key[16];
#pragma unroll
	for (int u = 0; u < 16; u++) key[u] = (u * nt + tid < m) ? key_at(u * nt + tid) : 0ull;
	u64 prefix = 0;
	int bits = 0, need = k;
	__syncthreads();
	while (true) {
		for (int i = tid; i < 256; i += nt) hist[i] = 0;
		__syncthreads();
#pragma unroll
		for (int u = 0; u < 16; u++) {
			if (u * nt >= m) break; // uniform
			int digit = -1 - (tid & 31); // lanes without a key match nobody
			if (u * nt + tid < m && (bits == 0 || (key[u] >> (64 - bits)) == prefix)) digit = (int)((key[u] >> (56 - bits)) & 255u);
			const unsigned peers = __match_any_sync(0xFFFFFFFFu, digit);
			if (digit >= 0 && (tid & 31) == __ffs(peers) - 1) atomicAdd(&hist[digit], __popc(peers));
		}
		__syncthreads();
		if (tid < 32) {
			const int lane = tid;
			int c[8], sum = 0;
#pragma unroll
			for (int j = 0; j < 8; j++) c[j] = hist[lane * 8 + j], sum += c[j];
			int incl = sum;
			for (int o = 1; o < 32; o <<= 1) {
				int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
				if (lane >= o) incl += v;
			}
			int before = incl - sum;
			const bool mine = before < need && incl >= need;
			if (mine) {
				int j = 0;
				while (before + c[j] < need) before += c[j], j++;
				s_digit = lane * 8 + j;
				s_need = need - before;
				s_bin = c[j];
			}
		}
		__syncthreads();
		prefix = (prefix << 8) | (u64)s_digit;
		bits += 8;
		need = s_need;
		const int bin = s_bin;
		__syncthreads();
		if (bin <= 32 || bits == 64) break;
	}
	if (tid == 0) s_nout = 0, s_nbin = 0;
	for (int i = tid; i < k; i += nt) out[i] = KEY_INVALID;
	__syncthreads();
	u64 *binl = out + k;
#pragma unroll
	for (int u = 0; u < 16; u++) {
		if (u * nt + tid >= m) continue;
		const u64 top = key[u] >> (64 - bits);
		if (top < prefix)
			out[atomicAdd(&s_nout, 1)] = key[u];
		else if (top == prefix) {
			const int pos = atomicAdd(&s_nbin, 1);
			if (pos < 32) binl[pos] = key[u];
		}
	}
	__syncthreads();
	if (tid < 32) {
		const int lane = tid, nb = min(s_nbin, 32);
		u64 v = lane < nb ? binl[lane] : KEY_INVALID;
		for (int size = 2; size <= 32; size <<= 1)
			for (int stride = size >> 1; stride > 0; stride >>= 1) {
				const u64 o = __shfl_xor_sync(0xFFFFFFFFu, v, stride);
				const bool up = (lane & size) == 0, lower = (lane & stride) == 0;
				v = (lower == up) ? (v < o ? v : o) : (v < o ? o : v);
			}
		const int base = s_nout;
		__syncwarp();
		if (lane < need) out[base + lane] = v;
	}
	__syncthreads();
}

// Final select by a whole block: src[0..m) keys (KEY_INVALID entries ignored, read through L2 because other
// SMs may have written them) -> sorted (label, distance) list of query slot `q`.  All threads of the block
// call it; bufs = nwarps*cap keys of shared memory (cnts is unused, kept for the callers' layouts).
// REGS: lists of up to 16 keys per thread go through block_radix_select_regs (standalone select kernels only: the
// 32 extra live registers would cost the scan kernels occupancy).
template <bool REGS = false>
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
		// long list: radix-select the k smallest, then sort just those
		if (REGS && m <= 16ll * (long long)blockDim.x)
			block_radix_select_regs([=](int i) { return __ldcg(src + i); }, (int)m, k, bufs);
		else
			block_radix_select([=](int i) { return __ldcg(src + i); }, (int)m, k, bufs);
		int n = 64;
		while (n < k) n <<= 1;
		for (int i = k + threadIdx.x; i < n; i += blockDim.x) bufs[i] = KEY_INVALID;
		block_bitonic_sort(bufs, n, IdentityMap());
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
