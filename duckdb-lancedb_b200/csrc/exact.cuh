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
__device__ __forceinline__ float warp_exact_distance(int metric, const float *__restrict__ q, const float *__restrict__ x,
                                                     int d, bool vec4) {
	const unsigned lane = lane_id();
	float acc = 0.f, nx = 0.f, nq = 0.f;
	if (vec4) {
		const float4 *q4 = reinterpret_cast<const float4 *>(q), *x4 = reinterpret_cast<const float4 *>(x);
#pragma unroll 8
		for (int c = lane; c < (d >> 2); c += 32) {
			float4 a = __ldg(q4 + c), b = __ldg(x4 + c);
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
			float a = __ldg(q + c), b = __ldg(x + c);
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


} // namespace lb
