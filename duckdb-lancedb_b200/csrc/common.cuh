// common.cuh — shared device/host helpers of liblance_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdexcept>
#include <string>

namespace lb {

typedef unsigned long long u64;

enum Metric : int { METRIC_L2 = 0, METRIC_COSINE = 1, METRIC_DOT = 2 };

// Thrown inside the library, always caught at the C boundary (ffi.cpp) and turned into err_buf text.
struct Error : std::runtime_error {
	explicit Error(const std::string &m) : std::runtime_error(m) {}
};

#define LB_CUDA(expr)                                                                                   \
	do {                                                                                                \
		cudaError_t _e = (expr);                                                                        \
		if (_e != cudaSuccess)                                                                          \
			throw ::lb::Error(std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + __FILE__ + ":" + \
			                  std::to_string(__LINE__) + " (" #expr ")");                              \
	} while (0)

static constexpr u64 KEY_INVALID = ~0ull;

// ---- order-preserving (distance, row) -> u64 key.  Smaller key = better rank; ties on distance break on
// the smaller row, the tie rule the oracle uses (oracle/oracle.c ent_less).
__host__ __device__ inline uint32_t float_to_ordered(float f) {
#ifdef __CUDA_ARCH__
	f = f + 0.0f; // -0 -> +0
	uint32_t u = __float_as_uint(f);
#else
	union { float f; uint32_t u; } c; c.f = f + 0.0f; uint32_t u = c.u;
#endif
	if (f != f) return 0xFFFFFFFEu; // NaN sorts last (but before the invalid key)
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float ordered_to_float(uint32_t o) {
	uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#ifdef __CUDA_ARCH__
	return __uint_as_float(u);
#else
	union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ inline u64 make_key(float dist, uint32_t row) {
	return ((u64)float_to_ordered(dist) << 32) | (u64)row;
}
__host__ __device__ inline float key_dist(u64 key) { return ordered_to_float((uint32_t)(key >> 32)); }
__host__ __device__ inline uint32_t key_row(u64 key) { return (uint32_t)(key & 0xFFFFFFFFull); }

__host__ __device__ inline int ceil_pow2(int v) {
	int p = 1;
	while (p < v) p <<= 1;
	return p;
}

// bit (i & 31) of word (i >> 5)
__device__ __forceinline__ bool bit_test(const uint32_t *bm, long long i) { return (bm[i >> 5] >> (i & 31)) & 1u; }

} // namespace lb
