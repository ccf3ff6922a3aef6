// flat_gemm_aux.cuh — the small kernels around flat_gemm_kernel (included by flat_gemm.cu):
// bf16 shadow of the table, bf16 query tiles + error margins, threshold refresh / list compaction between
// row segments, and the exact fp32 re-rank that produces the returned (label, distance) lists.
#pragma once
#include "exact.cuh"
#include <cuda_bf16.h>

#include "topk.cuh"

namespace lb {

// Worst-case |bf16 score - exact score| <= ERR_C * |q| * |x|: both operands are rounded to bf16 (2^-9
// relative each, 2^-8 + 2^-18 on a product, Cauchy-Schwarz over the sum) plus the fp32 accumulation in
// the tensor core; 2^-8 = 0.00390625, the rest is slack.
static constexpr float GEMM_ERR_C = 0.0041f;


// ------------------------------------------------------------------ shadow rows
// One warp per row: Xs[row] = bf16 of  x (dot) | x/|x| (cosine) | [x, -|x|^2/2 as three bf16 terms] (l2),
// zero padded to Kpad.  max_norm = max |x| over the converted rows (float bits, non-negative -> uint max).
__global__ void __launch_bounds__(256) shadow_rows_kernel(const float *__restrict__ X, long long row_lo, long long row_hi,
                                                          int d, int metric, int Kpad, __nv_bfloat16 *__restrict__ Xs,
                                                          uint32_t *__restrict__ max_norm_bits) {
	long long row = row_lo + (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (row >= row_hi) return;
	const unsigned lane = lane_id();
	const float *x = X + (size_t)row * d;
	__nv_bfloat16 *o = Xs + (size_t)row * Kpad;
	float s = 0.f;
	for (int i = lane; i < d; i += 32) {
		float v = __ldg(x + i);
		s = fmaf(v, v, s);
	}
	s = warp_sum(s);
	const float nrm = sqrtf(s);
	float scale = 1.f;
	// cosine: a row whose norm is zero or not finite has distance 1 in the exact kernels (their `den > 0` test fails),
	// i.e. similarity 0: its shadow is the zero row
	const bool cos_null = metric == METRIC_COSINE && !(nrm > 0.f && nrm <= 3.0e38f);
	if (metric == METRIC_COSINE) scale = cos_null ? 0.f : 1.0f / nrm;
	for (int i = lane; i < Kpad; i += 32) {
		float v = 0.f;
		if (i < d && !cos_null) v = __ldg(x + i) * scale;
		o[i] = __float2bfloat16_rn(v);
	}
	if (metric == METRIC_L2 && lane == 0) {
		float h = -0.5f * s;
		__nv_bfloat16 h1 = __float2bfloat16_rn(h);
		float r1 = h - __bfloat162float(h1);
		__nv_bfloat16 h2 = __float2bfloat16_rn(r1);
		float r2 = r1 - __bfloat162float(h2);
		__nv_bfloat16 h3 = __float2bfloat16_rn(r2);
		o[d] = h1, o[d + 1] = h2, o[d + 2] = h3;
	}
	// A row with a NaN / Inf component must not poison the margin of every query (max over the finite rows only);
	// its own scores are NaN / Inf and never pass a threshold, so it can only matter when fewer than k finite rows
	// survive — count_flags_kernel sends exactly those queries to the exact scan, which ranks such rows last.
	if (lane == 0 && metric != METRIC_COSINE) {
		if (!(nrm <= 3.0e38f))
			atomicOr(max_norm_bits + 1, 1u);
		else
			atomicMax(max_norm_bits, __float_as_uint(nrm));
	}
}

// ------------------------------------------------------------------ query tiles
// One warp per padded query slot: Qs[q] = bf16(q) (+ three 1.0 columns for l2), margin[q] = 2 * err bound,
// thr[q] = -inf (+inf for padding slots, which then never emit), cnt[q] = 0, flag[q] = 0.
__global__ void __launch_bounds__(256) prep_query_tiles_kernel(const float *__restrict__ Q, int nq, int nslots, int d,
                                                               int metric, int Kpad, const uint32_t *__restrict__ max_norm_bits,
                                                               __nv_bfloat16 *__restrict__ Qs, float *__restrict__ margin,
                                                               float *__restrict__ thr, uint32_t *__restrict__ cnt,
                                                               uint32_t *__restrict__ flag) {
	int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= nslots) return;
	const unsigned lane = lane_id();
	__nv_bfloat16 *o = Qs + (size_t)q * Kpad;
	const bool real = q < nq;
	float s = 0.f;
	for (int i = lane; i < Kpad; i += 32) {
		float v = (real && i < d) ? __ldg(Q + (size_t)q * d + i) : 0.f;
		if (real && metric == METRIC_L2 && i >= d && i < d + 3) v = 1.0f;
		else s = fmaf(v, v, s);
		o[i] = __float2bfloat16_rn(v);
	}
	s = warp_sum(s);
	if (lane == 0) {
		float xn = metric == METRIC_COSINE ? 1.01f : __uint_as_float(*max_norm_bits);
		float qn = sqrtf(s);
		// l2: the folded -|x|^2/2 term is exact to ~2^-24 relative; covered by the absolute slack below
		const float mg = 2.0f * (GEMM_ERR_C * qn * xn + 1e-6f * (1.0f + xn * xn));
		// a query (or table) whose bound is not finite cannot be certified: straight to the exact scan, emit nothing
		const bool certifiable = mg <= 3.0e38f;
		margin[q] = mg;
		thr[q] = (real && certifiable) ? -__int_as_float(0x7F800000) : __int_as_float(0x7F800000);
		cnt[q] = 0;
		flag[q] = (real && !certifiable) ? 1u : 0u;
	}
}

// ------------------------------------------------------------------ threshold refresh between segments
// One block per query.  A_k = k-th best approximate score among the candidates gathered so far (all live
// rows): every true top-k row scores >= A_k - margin, so thr = A_k - margin and the list is compacted to
// the entries that still pass.  A list that overflowed its capacity flags the query (exact re-run).
__global__ void __launch_bounds__(256) refresh_threshold_kernel(uint2 *__restrict__ cand, int CAP, uint32_t *__restrict__ cnt,
                                                                float *__restrict__ thr, const float *__restrict__ margin,
                                                                uint32_t *__restrict__ flag, int k, int cap) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	__shared__ int cnts[8];
	__shared__ float s_thr;
	__shared__ int s_base;
	const int q = blockIdx.x;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const uint32_t c_raw = cnt[q];
	if (flag[q]) return;
	if (c_raw > (uint32_t)CAP) {
		if (threadIdx.x == 0) flag[q] = 1;
		return;
	}
	const int c = (int)c_raw;
	if (c < k) return; // fewer than k live candidates so far: keep everything, threshold stays
	uint2 *list = cand + (size_t)q * CAP;
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw);
	if (c <= 8 * cap && c <= 1024) {
		// short list (every refresh but the one after the seed segment): one block-wide sort instead of per-warp
		// selection plus a serial fold — this kernel runs once per segment and its latency is pure overhead
		int n = 64;
		while (n < c) n <<= 1;
		for (int i = threadIdx.x; i < n; i += 256)
			bufs[i] = i < c ? make_key(-__uint_as_float(list[i].x), (uint32_t)i) : KEY_INVALID;
		block_bitonic_sort(bufs, n, IdentityMap());
		if (threadIdx.x == 0) {
			float ak = -key_dist(bufs[k - 1]);
			float t = ak - margin[q];
			s_thr = fmaxf(t, thr[q]);
			s_base = 0;
		}
	} else {
		// long list (after the seed segment, or large k): radix-select the k best scores; A_k is the worst of them
		__shared__ u64 s_max;
		if (threadIdx.x == 0) s_max = 0;
		block_radix_select([=](int i) { return make_key(-__uint_as_float(list[i].x), (uint32_t)i); }, c, k, bufs);
		u64 mx = 0;
		for (int i = threadIdx.x; i < k; i += 256) mx = bufs[i] > mx ? bufs[i] : mx;
		for (int o = 16; o; o >>= 1) {
			const u64 v = __shfl_xor_sync(0xFFFFFFFFu, mx, o);
			mx = v > mx ? v : mx;
		}
		if (lane == 0) atomicMax(&s_max, mx);
		__syncthreads();
		if (threadIdx.x == 0) {
			float ak = -key_dist(s_max);
			float t = ak - margin[q];
			s_thr = fmaxf(t, thr[q]);
			s_base = 0;
		}
	}
	__syncthreads();
	const float t = s_thr;
	// in-place compaction, chunk by chunk (writes never pass the reads of the same or a later chunk)
	for (int i0 = 0; i0 < c; i0 += 256) {
		int i = i0 + threadIdx.x;
		uint2 e = make_uint2(0, 0);
		bool keep = false;
		if (i < c) {
			e = list[i];
			keep = __uint_as_float(e.x) >= t;
		}
		unsigned m = __ballot_sync(0xFFFFFFFFu, keep);
		if (lane == 0) cnts[warp] = __popc(m);
		__syncthreads();
		int off = s_base;
		for (int w = 0; w < warp; w++) off += cnts[w];
		if (keep) list[off + __popc(m & ((1u << lane) - 1u))] = e;
		__syncthreads();
		if (threadIdx.x == 0) {
			int tot = 0;
			for (int w = 0; w < 8; w++) tot += cnts[w];
			s_base += tot;
		}
		__syncthreads();
	}
	if (threadIdx.x == 0) {
		cnt[q] = (uint32_t)s_base;
		thr[q] = t;
	}
}

// ------------------------------------------------------------------ exact re-rank (warp_exact_distance: exact.cuh)
// One block per query: exact fp32 distance of every surviving candidate, fused top-k, sorted output.
__global__ void __launch_bounds__(256) rerank_kernel(const float *__restrict__ X, int d, int metric,
                                                     const float *__restrict__ Q, const uint2 *__restrict__ cand, int CAP,
                                                     const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ flag,
                                                     int k, int cap, long long label_base, int64_t *__restrict__ out_labels,
                                                     float *__restrict__ out_dist, int32_t *__restrict__ out_counts) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	__shared__ int cnts[8];
	const int q = blockIdx.x;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw);
	const bool bad = flag[q] != 0 || cnt[q] > (uint32_t)CAP;
	const int c = bad ? 0 : (int)cnt[q];
	const uint2 *list = cand + (size_t)q * CAP;
	const float *qv = Q + (size_t)q * d;
	const bool vec4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0) && ((reinterpret_cast<uintptr_t>(Q) & 15) == 0);
	WarpSelect sel;
	sel.init(bufs + (size_t)warp * cap, cap, k);
	for (int i0 = warp * 32; i0 < c; i0 += 256) {
		const int mine = i0 + lane;
		const uint32_t myrow = mine < c ? list[mine].y : 0u;
		float mydist = 0.f;
		const int nb = min(32, c - i0);
		for (int j = 0; j < nb; j += 2) {
			const uint32_t r0 = __shfl_sync(0xFFFFFFFFu, myrow, j);
			const uint32_t r1 = __shfl_sync(0xFFFFFFFFu, myrow, min(j + 1, nb - 1));
			float d0, d1;
			warp_exact_distance_x2(metric, qv, X + (size_t)r0 * d, X + (size_t)r1 * d, d, vec4, d0, d1);
			if ((int)lane == j) mydist = d0;
			if ((int)lane == j + 1) mydist = d1;
		}
		sel.push(mine < c, make_key(mydist, myrow));
	}
	sel.finalize();
	if (lane == 0) cnts[warp] = sel.cnt;
	__syncthreads();
	if (warp != 0) return;
	for (int w = 1; w < 8; w++) {
		const u64 *s2 = bufs + (size_t)w * cap;
		const int cw = cnts[w];
		for (int i0 = 0; i0 < cw; i0 += 32) {
			int i = i0 + lane;
			sel.push(i < cw, i < cw ? s2[i] : KEY_INVALID);
		}
	}
	sel.finalize();
	for (int i = lane; i < k; i += 32) {
		bool ok = i < sel.cnt;
		u64 key = ok ? sel.buf[i] : KEY_INVALID;
		out_labels[(size_t)q * k + i] = ok ? (int64_t)(label_base + (long long)key_row(key)) : -1;
		out_dist[(size_t)q * k + i] = ok ? key_dist(key) : __int_as_float(0x7F800000);
	}
	if (lane == 0) out_counts[q] = sel.cnt;
}

// Small batches: the same re-rank in two launches, so that the exact distances of one query's candidates are
// spread over T blocks instead of being computed by a single one (64 queries would use 64 of the 148 SMs).
// Phase 1, grid (nq, T): candidate i of query q -> (exact distance, row) key, written over the candidate in place.
__global__ void __launch_bounds__(256) rerank_dist_kernel(const float *__restrict__ X, int d, int metric,
                                                          const float *__restrict__ Q, uint2 *__restrict__ cand, int CAP,
                                                          const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ flag) {
	const int q = blockIdx.x;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	if (flag[q] != 0 || cnt[q] > (uint32_t)CAP) return;
	const int c = (int)cnt[q];
	uint2 *list = cand + (size_t)q * CAP;
	const float *qv = Q + (size_t)q * d;
	const bool vec4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0) && ((reinterpret_cast<uintptr_t>(Q) & 15) == 0);
	for (int i = blockIdx.y * 8 + warp; i < c; i += gridDim.y * 8) {
		const uint32_t row = list[i].y;
		const float dist = warp_exact_distance(metric, qv, X + (size_t)row * d, d, vec4);
		if (lane == 0) reinterpret_cast<u64 *>(list)[i] = make_key(dist, row);
	}
}
// Phase 2, one block per query: top-k of the keys.
__global__ void __launch_bounds__(256) rerank_select_kernel(const uint2 *__restrict__ cand, int CAP, const uint32_t *__restrict__ cnt,
                                                            const uint32_t *__restrict__ flag, int k, int cap, long long label_base,
                                                            int64_t *__restrict__ out_labels, float *__restrict__ out_dist,
                                                            int32_t *__restrict__ out_counts) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int q = blockIdx.x;
	const bool bad = flag[q] != 0 || cnt[q] > (uint32_t)CAP;
	const int c = bad ? 0 : (int)cnt[q];
	block_final_select(reinterpret_cast<const u64 *>(cand + (size_t)q * CAP), c, k, cap, reinterpret_cast<u64 *>(smem_raw), nullptr,
	                   label_base, (size_t)q, out_labels, out_dist, out_counts, nullptr);
}

// number of flagged queries -> *total (read by the host to run the exact scan for them).  Flagged: the list
// overflowed, the bound was not finite, or the table holds rows with non-finite components and the query came back
// with fewer than k results (the exact scan ranks such rows last instead of dropping them).
__global__ void count_flags_kernel(const uint32_t *__restrict__ flag, const uint32_t *__restrict__ cnt, int CAP, int nq,
                                   const int32_t *__restrict__ out_counts, int k, const uint32_t *__restrict__ nonfinite,
                                   uint32_t *__restrict__ flag_out, uint32_t *__restrict__ total) {
	int q = blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= nq) return;
	uint32_t f = (flag[q] != 0 || cnt[q] > (uint32_t)CAP) ? 1u : 0u;
	if (!f && *nonfinite != 0u && out_counts[q] < k) f = 1u;
	flag_out[q] = f;
	if (f) atomicAdd(total, 1u);
}

} // namespace lb
