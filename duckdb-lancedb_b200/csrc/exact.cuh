// exact.cuh — exact fp32 distance of one (query, row) pair by one warp, in the reference's reported units (same
// formulas as flat_scan.cu): used by the re-rank of the tcgen05 flat path and by the IVF_PQ refine step (a14).
#pragma once
#include "topk.cuh"

namespace lb {

__device__ __forceinline__ float warp_sum(float v) {
	for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
	return v;
}

// exact distance of (q, x) by one warp, in the reference's reported units (same formulas as flat_scan.cu)
// QLDG = false: q may live in shared memory (plain loads instead of the read-only global path); same arithmetic
template <bool QLDG = true>
__device__ __forceinline__ float warp_exact_distance(int metric, const float *__restrict__ q, const float *__restrict__ x,
                                                     int d, bool vec4) {
	const unsigned lane = lane_id();
	float acc = 0.f, nx = 0.f, nq = 0.f;
	if (vec4) {
		const float4 *q4 = reinterpret_cast<const float4 *>(q), *x4 = reinterpret_cast<const float4 *>(x);
#pragma unroll 8
		for (int c = lane; c < (d >> 2); c += 32) {
			float4 a = QLDG ? __ldg(q4 + c) : q4[c], b = __ldg(x4 + c);
			if (metric == METRIC_L2) {
				float e0 = a.x - b.x, e1 = a.y - b.y, e2 = a.z - b.z, e3 = a.w - b.w;
				acc = fmaf(e0, e0, acc), acc = fmaf(e1, e1, acc), acc = fmaf(e2, e2, acc), acc = fmaf(e3, e3, acc);
			} else {
				acc = fmaf(a.x, b.x, acc), acc = fmaf(a.y, b.y, acc), acc = fmaf(a.z, b.z, acc), acc = fmaf(a.w, b.w, acc);
				if (metric == METRIC_COSINE) {
					nx = fmaf(b.x, b.x, nx), nx = fmaf(b.y, b.y, nx), nx = fmaf(b.z, b.z, nx), nx = fmaf(b.w, b.w, nx);
					nq = fmaf(a.x, a.x, nq), nq = fmaf(a.y, a.y, nq), nq = fmaf(a.z, a.z, nq), nq = fmaf(a.w, a.w, nq);
				}
			}
		}
	} else {
		for (int c = lane; c < d; c += 32) {
			float a = QLDG ? __ldg(q + c) : q[c], b = __ldg(x + c);
			if (metric == METRIC_L2) {
				float e = a - b;
				acc = fmaf(e, e, acc);
			} else {
				acc = fmaf(a, b, acc);
				nx = fmaf(b, b, nx);
				nq = fmaf(a, a, nq);
			}
		}
	}
	acc = warp_sum(acc);
	if (metric == METRIC_L2) return acc;
	if (metric == METRIC_DOT) return 1.0f - acc;
	nx = warp_sum(nx);
	nq = warp_sum(nq);
	float den = sqrtf(nq) * sqrtf(nx);
	if (!(den > 0.f)) return 1.0f;
	float sim = fminf(1.f, fmaxf(-1.f, acc / den));
	return 1.0f - sim;
}

// Two rows at once (same arithmetic per row, in the same order): twice the loads in flight per warp — the re-rank
// reads scattered 4·d-byte rows and is bound by their latency (ncu: 60 % of its stall samples wait on these loads).
__device__ __forceinline__ void warp_exact_distance_x2(int metric, const float *__restrict__ q, const float *__restrict__ x0,
                                                       const float *__restrict__ x1, int d, bool vec4, float &o0, float &o1) {
	if (!vec4) {
		o0 = warp_exact_distance(metric, q, x0, d, false);
		o1 = warp_exact_distance(metric, q, x1, d, false);
		return;
	}
	const unsigned lane = lane_id();
	float acc[2] = {0.f, 0.f}, nx[2] = {0.f, 0.f}, nq = 0.f;
	const float4 *q4 = reinterpret_cast<const float4 *>(q);
	const float4 *xa = reinterpret_cast<const float4 *>(x0), *xb = reinterpret_cast<const float4 *>(x1);
#pragma unroll 4
	for (int c = lane; c < (d >> 2); c += 32) {
		const float4 a = __ldg(q4 + c);
		const float4 b[2] = {__ldg(xa + c), __ldg(xb + c)};
#pragma unroll
		for (int r = 0; r < 2; r++) {
			if (metric == METRIC_L2) {
				float e0 = a.x - b[r].x, e1 = a.y - b[r].y, e2 = a.z - b[r].z, e3 = a.w - b[r].w;
				acc[r] = fmaf(e0, e0, acc[r]), acc[r] = fmaf(e1, e1, acc[r]), acc[r] = fmaf(e2, e2, acc[r]), acc[r] = fmaf(e3, e3, acc[r]);
			} else {
				acc[r] = fmaf(a.x, b[r].x, acc[r]), acc[r] = fmaf(a.y, b[r].y, acc[r]), acc[r] = fmaf(a.z, b[r].z, acc[r]),
				acc[r] = fmaf(a.w, b[r].w, acc[r]);
				if (metric == METRIC_COSINE)
					nx[r] = fmaf(b[r].x, b[r].x, nx[r]), nx[r] = fmaf(b[r].y, b[r].y, nx[r]), nx[r] = fmaf(b[r].z, b[r].z, nx[r]),
					nx[r] = fmaf(b[r].w, b[r].w, nx[r]);
			}
		}
		if (metric == METRIC_COSINE) nq = fmaf(a.x, a.x, nq), nq = fmaf(a.y, a.y, nq), nq = fmaf(a.z, a.z, nq), nq = fmaf(a.w, a.w, nq);
	}
	float out[2];
	if (metric == METRIC_COSINE) nq = warp_sum(nq);
#pragma unroll
	for (int r = 0; r < 2; r++) {
		const float s = warp_sum(acc[r]);
		if (metric == METRIC_L2)
			out[r] = s;
		else if (metric == METRIC_DOT)
			out[r] = 1.0f - s;
		else {
			const float den = sqrtf(nq) * sqrtf(warp_sum(nx[r]));
			out[r] = !(den > 0.f) ? 1.0f : 1.0f - fminf(1.f, fmaxf(-1.f, s / den));
		}
	}
	o0 = out[0], o1 = out[1];
}

} // namespace lb
