// flat_gemm_aux.cuh — the small kernels around flat_gemm_kernel (included by flat_gemm.cu):
// bf16 shadow of the table, bf16 query tiles + error margins, threshold refresh / list compaction between
// row segments, and the exact fp32 re-rank that produces the returned (label, distance) lists.
#pragma once
#include "exact.cuh"
#include <cuda_bf16.h>

#include "topk.cuh"

namespace lb {

// Error bound of a bf16 score.  With q^ = bf16(q'), x^ = bf16(x') (q', x' = the fp32 rows that are rounded: x, x/|x| or
// [x, -|x|^2/2 split] against q or [q, 1, 1, 1]) the tensor core computes fl(q^.x^) and
//   |q^.x^ - q'.x'| = |(q^-q').x^ + q'.(x^-x')| <= rq |x^| + |q| rx,     rq = |q^-q'|, rx = |x^-x'|  (Cauchy-Schwarz),
// where rq, |q| are MEASURED per query and rx, |x^| per row (their maxima over the table are kept).  Round 1 used
// the a-priori worst case 2^-8 |q||x| of round-to-nearest; the measured residuals are about half of it (a rounding
// error is rarely the full half ulp in every component), and a halved margin means ~30 % fewer candidates to emit,
// refresh and re-rank at C2.  For l2 the constant columns (1.0) and the split of -|x|^2/2 are exact bf16 numbers, so
// both residuals live on the first d columns only.  On top: the fp32 accumulation inside the tensor core (order
// unknown, truncating adders: 2^-22 per term is generous) plus the rounding of the exact fp32 re-rank the candidates
// are compared by (2^-23 per term, worst case), together GEMM_ACC_C K |q^||x^|, and an absolute slack for the residual
// of the three-term split.
static constexpr float GEMM_ACC_C = 3.6e-7f; // per accumulated term

__device__ __forceinline__ void lazy_atomic_max(uint32_t *p, uint32_t v) {
	if (v > __ldca(p)) atomicMax(p, v);
}

// ------------------------------------------------------------------ shadow rows
// One warp per row: Xs[row] = bf16 of  x (dot) | x/|x| (cosine) | [x, -|x|^2/2 as three bf16 terms] (l2),
// zero padded to Kpad.  max_norm = max |x| over the converted rows (float bits, non-negative -> uint max).
__global__ void __launch_bounds__(256) shadow_rows_kernel(const float *__restrict__ X, long long row_lo, long long row_hi,
                                                          int d, int metric, int Kpad, __nv_bfloat16 *__restrict__ Xs,
                                                          uint32_t *__restrict__ max_norm_bits,
                                                          float *__restrict__ bias = nullptr) {
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
	float res2 = 0.f, hat2 = 0.f; // |x^ - x'|^2 and |x^|^2 over the first d columns
	for (int i = lane; i < Kpad; i += 32) {
		float v = 0.f;
		if (i < d && !cos_null) v = __ldg(x + i) * scale;
		const __nv_bfloat16 b = __float2bfloat16_rn(v);
		o[i] = b;
		if (i < d) {
			const float bf = __bfloat162float(b), e = bf - v;
			res2 = fmaf(e, e, res2);
			hat2 = fmaf(bf, bf, hat2);
		}
	}
	res2 = warp_sum(res2);
	hat2 = warp_sum(hat2);
	if (metric == METRIC_L2 && bias) {
		if (lane == 0) bias[row] = -0.5f * s; // bias epilogue (K-streamed kernel): the term stays in fp32
	} else if (metric == METRIC_L2 && lane == 0) {
		float h = -0.5f * s;
		__nv_bfloat16 h1 = __float2bfloat16_rn(h);
		float r1 = h - __bfloat162float(h1);
		__nv_bfloat16 h2 = __float2bfloat16_rn(r1);
		float r2 = r1 - __bfloat162float(h2);
		__nv_bfloat16 h3 = __float2bfloat16_rn(r2);
		o[d] = h1, o[d + 1] = h2, o[d + 2] = h3;
	}
	// max_norm words: [0] max |x| (dot, l2) | [1] "a row with a non-finite norm exists" | [2] max |x^ - x'| | [3] max |x^|
	// (float bits of non-negative values order like unsigned integers).  A row with a NaN / Inf component must not
	// poison the margin of every query (maxima over the finite rows only); its own scores are NaN / Inf and never pass
	// a threshold, so it can only matter when fewer than k finite rows survive — count_flags_kernel sends exactly
	// those queries to the exact scan, which ranks such rows last.
	if (lane == 0) {
		if (metric != METRIC_COSINE && !(nrm <= 3.0e38f)) {
			atomicOr(max_norm_bits + 1, 1u);
		} else {
			// a running maximum grows O(log n) times: look at the (L1-cached, possibly stale = smaller) word first instead of
			// sending three same-address atomics per row to one L2 slice (100 M rows: 101 ms of atomics around 13 ms of copying)
			if (metric != METRIC_COSINE) lazy_atomic_max(max_norm_bits, __float_as_uint(nrm));
			lazy_atomic_max(max_norm_bits + 2, __float_as_uint(sqrtf(res2)));
			lazy_atomic_max(max_norm_bits + 3, __float_as_uint(sqrtf(hat2)));
		}
	}
}

// ------------------------------------------------------------------ query tiles
// One warp per padded query slot: Qs[q] = bf16(q) (+ three 1.0 columns for l2), margin[q] = 2 * err bound,
// thr[q] = -inf (+inf for padding slots, which then never emit), cnt[q] = 0, flag[q] = 0.
__global__ void __launch_bounds__(256) prep_query_tiles_kernel(const float *__restrict__ Q, int nq, int nslots, int d,
                                                               int metric, int Kpad, const uint32_t *__restrict__ max_norm_bits,
                                                               __nv_bfloat16 *__restrict__ Qs, float *__restrict__ margin,
                                                               float *__restrict__ thr, uint32_t *__restrict__ cnt,
                                                               uint32_t *__restrict__ flag, int bias_mode = 0) {
	int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= nslots) return;
	const unsigned lane = lane_id();
	__nv_bfloat16 *o = Qs + (size_t)q * Kpad;
	const bool real = q < nq;
	float s = 0.f, res2 = 0.f;
	for (int i = lane; i < Kpad; i += 32) {
		float v = (real && i < d) ? __ldg(Q + (size_t)q * d + i) : 0.f;
		if (real && metric == METRIC_L2 && !bias_mode && i >= d && i < d + 3) v = 1.0f; // exact in bf16: no residual
		else s = fmaf(v, v, s);
		const __nv_bfloat16 b = __float2bfloat16_rn(v);
		o[i] = b;
		const float e = __bfloat162float(b) - v;
		res2 = fmaf(e, e, res2);
	}
	s = warp_sum(s);
	res2 = warp_sum(res2);
	if (lane == 0) {
		const float xn = metric == METRIC_COSINE ? 1.01f : __uint_as_float(max_norm_bits[0]);
		const float rx = __uint_as_float(max_norm_bits[2]), xh = __uint_as_float(max_norm_bits[3]);
		const float qn = sqrtf(s), rq = sqrtf(res2);
		// full operand norms for the accumulation term: l2 rows carry -|x|^2/2, the query three ones
		// (bias epilogue: the contraction is the plain inner product, the fp32 term adds one rounding of its own size)
		const bool l2cols = metric == METRIC_L2 && !bias_mode;
		const float qfull = l2cols ? sqrtf(s + 3.0f) : qn;
		const float xfull = l2cols ? sqrtf(xh * xh + 0.25f * xn * xn * xn * xn) * 1.01f : xh;
		const float err = 1.02f * (rq * xh + qn * rx) + GEMM_ACC_C * (float)Kpad * qfull * xfull + 1e-6f * (1.0f + xn * xn);
		const float mg = 2.0f * err;
		// a query (or table) whose bound is not finite cannot be certified: straight to the exact scan, emit nothing
		const bool certifiable = mg <= 3.0e38f;
		margin[q] = mg;
		thr[q] = (real && certifiable) ? -__int_as_float(0x7F800000) : __int_as_float(0x7F800000);
		cnt[q] = 0;
		flag[q] = (real && !certifiable) ? 1u : 0u;
	}
}

__global__ void fill_u32_kernel(uint32_t *__restrict__ dst, int n, uint32_t v) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) dst[i] = v;
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

// ------------------------------------------------------------------ nearest row from the seeding pass's maxima
// FlatGemm::nearest: entries[q][0..E) = (best bf16 score, its row) of every (half) tile of the table.  One warp per point:
// the four best entries (one per lane at most), the exact fp32 distance of the point to their rows, argmin (ties to the
// lower row, as nearest_center_kernel).
__global__ void __launch_bounds__(256) nearest_from_maxima_kernel(const uint2 *__restrict__ entries, int stride, int E,
                                                                  const float *__restrict__ table, int d, int metric,
                                                                  const float *__restrict__ Q, int nq, int32_t *__restrict__ out) {
	const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
	if (q >= nq) return;
	const unsigned lane = lane_id();
	const uint2 *list = entries + (size_t)q * stride;
	float bs = -__int_as_float(0x7F800000);
	uint32_t br = 0xFFFFFFFFu;
	for (int e = lane; e < E; e += 32) {
		const uint2 c = list[e];
		const float sc = __uint_as_float(c.x);
		if (sc > bs || (sc == bs && c.y < br)) bs = sc, br = c.y;
	}
	const float *qv = Q + (size_t)q * d;
	const bool vec4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(table) & 15) == 0) && ((reinterpret_cast<uintptr_t>(Q) & 15) == 0);
	float best_d = __int_as_float(0x7F800000);
	uint32_t best_row = 0xFFFFFFFFu;
	for (int t = 0; t < 4; t++) {
		float ms = bs;
		uint32_t mr = br;
		for (int o = 16; o; o >>= 1) {
			const float os = __shfl_xor_sync(0xFFFFFFFFu, ms, o);
			const uint32_t orow = __shfl_xor_sync(0xFFFFFFFFu, mr, o);
			if (os > ms || (os == ms && orow < mr)) ms = os, mr = orow;
		}
		if (mr == 0xFFFFFFFFu) break; // warp-uniform: no entry left
		const float dist = warp_exact_distance(metric, qv, table + (size_t)mr * d, d, vec4);
		if (dist < best_d || (dist == best_d && mr < best_row)) best_d = dist, best_row = mr;
		if (br == mr) bs = -__int_as_float(0x7F800000), br = 0xFFFFFFFFu; // rows are distinct across entries
	}
	if (lane == 0) out[q] = best_row == 0xFFFFFFFFu ? 0 : (int32_t)best_row;
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
	// A short list (k = 10: a few dozen survivors) is dealt to ALL eight warps, ceil(c / 8) candidates each: the block's
	// latency is one warp's chain of dependent row gathers, and 40 candidates on two warps were 16 of them in a row.
	const int per = c <= 256 ? (c + 7) >> 3 : 32;
	for (int base = 0; base < c; base += 8 * per) {
		const int i0 = base + warp * per;
		const int nb = max(0, min(per, c - i0));
		const uint32_t myrow = (int)lane < nb ? list[i0 + lane].y : 0u;
		float mydist = 0.f;
		for (int j = 0; j < nb; j += 2) {
			const uint32_t r0 = __shfl_sync(0xFFFFFFFFu, myrow, j);
			const uint32_t r1 = __shfl_sync(0xFFFFFFFFu, myrow, min(j + 1, nb - 1));
			float d0, d1;
			warp_exact_distance_x2(metric, qv, X + (size_t)r0 * d, X + (size_t)r1 * d, d, vec4, d0, d1);
			if ((int)lane == j) mydist = d0;
			if ((int)lane == j + 1) mydist = d1;
		}
		sel.push((int)lane < nb, make_key(mydist, myrow));
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
