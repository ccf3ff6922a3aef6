// scan_common.cuh — device helpers shared by the fp32 streaming scans (flat_scan.cu: 1..4 queries per pass;
// flat_scan1.cuh: the single-query latency path): 128-bit row loads, per-metric accumulation, the transposed warp
// reduction and the distance finish.  Both kernels use the same per-lane accumulation order, so they return
// bit-identical distances.
#pragma once
#include "kernels.h"
#include "topk.cuh"

namespace lb {

// Transposed warp reduction: v[i] = this lane's partial sum for row i (8 rows).  Returns the full
// 32-lane sum of row (lane>>2)&7.  7 + 2 shuffles.
__device__ __forceinline__ float reduce8(float (&v)[8]) {
	const unsigned lane = lane_id();
	{
		const bool up = lane & 16;
#pragma unroll
		for (int i = 0; i < 4; i++) {
			float send = up ? v[i] : v[i + 4];
			float keep = up ? v[i + 4] : v[i];
			v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 16);
		}
	}
	{
		const bool up = lane & 8;
#pragma unroll
		for (int i = 0; i < 2; i++) {
			float send = up ? v[i] : v[i + 2];
			float keep = up ? v[i + 2] : v[i];
			v[i] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 8);
		}
	}
	{
		const bool up = lane & 4;
		float send = up ? v[0] : v[1];
		float keep = up ? v[1] : v[0];
		v[0] = keep + __shfl_xor_sync(0xFFFFFFFFu, send, 4);
	}
	float s = v[0];
	s += __shfl_xor_sync(0xFFFFFFFFu, s, 2);
	s += __shfl_xor_sync(0xFFFFFFFFu, s, 1);
	return s;
}

template <int VEC>
struct XVec;
template <>
struct XVec<4> {
	float4 v;
	__device__ __forceinline__ void zero() { v = make_float4(0.f, 0.f, 0.f, 0.f); }
	__device__ __forceinline__ void load(const float *p) { v = __ldg(reinterpret_cast<const float4 *>(p)); }
	__device__ __forceinline__ void load_shared(const float *p) { v = *reinterpret_cast<const float4 *>(p); }
};
template <>
struct XVec<1> {
	float v;
	__device__ __forceinline__ void zero() { v = 0.f; }
	__device__ __forceinline__ void load(const float *p) { v = __ldg(p); }
	__device__ __forceinline__ void load_shared(const float *p) { v = *p; }
};

template <int METRIC>
__device__ __forceinline__ void accum(float &acc, const XVec<4> &q, const XVec<4> &x) {
	if (METRIC == METRIC_L2) {
		float a = q.v.x - x.v.x, b = q.v.y - x.v.y, c = q.v.z - x.v.z, e = q.v.w - x.v.w;
		acc = fmaf(a, a, acc);
		acc = fmaf(b, b, acc);
		acc = fmaf(c, c, acc);
		acc = fmaf(e, e, acc);
	} else {
		acc = fmaf(q.v.x, x.v.x, acc);
		acc = fmaf(q.v.y, x.v.y, acc);
		acc = fmaf(q.v.z, x.v.z, acc);
		acc = fmaf(q.v.w, x.v.w, acc);
	}
}
template <int METRIC>
__device__ __forceinline__ void accum(float &acc, const XVec<1> &q, const XVec<1> &x) {
	if (METRIC == METRIC_L2) {
		float a = q.v - x.v;
		acc = fmaf(a, a, acc);
	} else {
		acc = fmaf(q.v, x.v, acc);
	}
}
__device__ __forceinline__ void accum_norm(float &acc, const XVec<4> &x) {
	acc = fmaf(x.v.x, x.v.x, acc);
	acc = fmaf(x.v.y, x.v.y, acc);
	acc = fmaf(x.v.z, x.v.z, acc);
	acc = fmaf(x.v.w, x.v.w, acc);
}
__device__ __forceinline__ void accum_norm(float &acc, const XVec<1> &x) { acc = fmaf(x.v, x.v, acc); }

// distance in the reference's reported units from the accumulated sums
template <int METRIC>
__device__ __forceinline__ float finish_distance(float acc, float qnorm2, float xnorm2) {
	if (METRIC == METRIC_L2) return acc;
	if (METRIC == METRIC_DOT) return 1.0f - acc;
	float den = sqrtf(qnorm2) * sqrtf(xnorm2);
	if (!(den > 0.f)) return 1.0f;
	float sim = acc / den;
	sim = fminf(1.f, fmaxf(-1.f, sim));
	return 1.0f - sim;
}

// liveness of the 8 rows [row0, row0+8), row0 % 8 == 0
__device__ __forceinline__ unsigned live_mask8(long long row0, long long row_begin, long long n,
                                               const uint32_t *allow, long long allow_bits,
                                               const uint32_t *deleted) {
	unsigned m = 0xFFu;
	long long left = n - row0;
	if (left <= 0) return 0u;
	if (left < 8) m = (1u << (int)left) - 1u;
	if (row0 < row_begin) {
		if (row_begin - row0 >= 8) return 0u;
		m &= ~((1u << (int)(row_begin - row0)) - 1u);
	}
	const int sh = (int)(row0 & 31);
	if (deleted) m &= ~(__ldg(deleted + (row0 >> 5)) >> sh);
	if (allow) {
		unsigned a = 0;
		if (row0 < allow_bits) {
			a = (__ldg(allow + (row0 >> 5)) >> sh) & 0xFFu;
			long long aleft = allow_bits - row0;
			if (aleft < 8) a &= (1u << (int)aleft) - 1u;
		}
		m &= a;
	}
	return m & 0xFFu;
}

} // namespace lb
