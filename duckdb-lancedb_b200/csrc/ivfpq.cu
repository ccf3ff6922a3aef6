// ivfpq.cu — IVF_PQ on the device: probe -> per-query lookup table -> ADC scan -> refine, plus the
// trainer behind lance_detached_create_index.
//
// What the reference delegates to lancedb's IvfPq index (rust_lib/src/lance_manager.rs:483-515 build,
// :411-419 query with .nprobes(n).refine_factor(rf)); SURVEY.md 8a rows a12-a14.  The arithmetic follows
// oracle/oracle.c (oracle_ivf_probe / oracle_pq_lut / oracle_adc_distance / oracle_ivfpq_search):
//   probe : the nprobes centroids nearest to q' (q' = q, or q/|q| for cosine), L2 (l2, cosine) or 1-dot;
//   LUT   : residual PQ, entry(m,c) = |(q' - centroid_p)_m - cb[m][c]|^2   (dot: -(q_m . cb[m][c]), no residual);
//   ADC   : l2 = sum, cosine = sum/2, dot = 1 + sum;   refine: exact distance of the k*rf best, re-sorted.
//
// B200 mapping.  The residual LUT is never materialised per (query, partition).  Expanding the square,
//   sum_m |(q'-c_p)_m - cb_m|^2 = sum_m |q'_m - cb_m|^2  +  (|q'-c_p|^2 - |q'|^2)  +  2 sum_m <c_p,m , cb_m>
//                                 `-- per-QUERY table T1 --'  `-- per (q,p) scalar, --'  `-- per ROW, stored --'
//                                                              known from the probe      at index build
// so one 96 KB table per query (M=96) serves all probed partitions, the codebook is read once per query
// batch instead of once per (query, partition), and the scan kernel streams nothing but codes
// (M bytes/row), a 4-byte row term and a 4-byte row id.  Codes are stored interleaved in blocks of 32
// rows ([block][M/16][32 lanes][16 B]) so that every warp load is one fully coalesced 512-byte
// request of 128-bit words; every partition is padded to a multiple of 32 rows.
// Roofline of the scan: HBM (codes + row terms + ids), in practice bounded by shared-memory lookups
// (M random LDS per row); see DESIGN.md.
#include "ivfpq.h"

#include <cooperative_groups.h>
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

#include "exact.cuh"
#include "flat_gemm.h"
#include "flat_scan1.cuh"
#include "tc_ptx.cuh"
#include "topk.cuh"

namespace lb {

// ------------------------------------------------------------------ query preparation
// q' = q (l2, dot) or q/|q| (cosine); qn2 = |q'|^2.  One warp per query.
__global__ void prep_queries_kernel(const float *__restrict__ Q, int nq, int d, int normalise, float *__restrict__ Qp,
                                    float *__restrict__ qn2) {
	int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (q >= nq) return;
	const unsigned lane = lane_id();
	const float *src = Q + (size_t)q * d;
	float s = 0.f;
	for (int i = lane; i < d; i += 32) s = fmaf(src[i], src[i], s);
	for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
	float r = 1.f;
	if (normalise) r = s > 0.f ? rsqrtf(s) : 1.f;
	float s2 = 0.f;
	for (int i = lane; i < d; i += 32) {
		float v = src[i] * r;
		Qp[(size_t)q * d + i] = v;
		s2 = fmaf(v, v, s2);
	}
	for (int o = 16; o; o >>= 1) s2 += __shfl_xor_sync(0xFFFFFFFFu, s2, o);
	if (lane == 0) qn2[q] = s2;
}

// ------------------------------------------------------------------ batched probe (a12), nq >= 8
// keys[q][p] = (d(q', c_p), p) for a 64 x 64 tile of (queries, centroids) per CTA, fp32, same per-element formula
// as the streaming scan ((a-b)^2 summed, or 1 - a.b).  The centroid table is small (P x d, L2-resident), so this
// is a register-tiled fp32 contraction (4 x 4 outputs per thread) rather than a pass of the table per 4 queries.
template <int IS_DOT>
__global__ void __launch_bounds__(256) probe_tile_kernel(const float *__restrict__ Qp, int nq, const float *__restrict__ C,
                                                         int P, int d, u64 *__restrict__ keys) {
	__shared__ __align__(16) float As[16][68];
	__shared__ __align__(16) float Bs[16][68];
	const int tq = threadIdx.x >> 4, tp = threadIdx.x & 15;
	const int q0 = blockIdx.y * 64, p0 = blockIdx.x * 64;
	float acc[4][4];
#pragma unroll
	for (int i = 0; i < 4; i++)
#pragma unroll
		for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
	for (int k0 = 0; k0 < d; k0 += 16) {
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const int i = threadIdx.x + u * 256;
			const int r = i >> 4, c = i & 15;
			const bool kin = k0 + c < d;
			As[c][r] = (kin && q0 + r < nq) ? __ldg(Qp + (size_t)(q0 + r) * d + k0 + c) : 0.f;
			Bs[c][r] = (kin && p0 + r < P) ? __ldg(C + (size_t)(p0 + r) * d + k0 + c) : 0.f;
		}
		__syncthreads();
#pragma unroll
		for (int kk = 0; kk < 16; kk++) {
			const float4 a4 = *reinterpret_cast<const float4 *>(&As[kk][tq * 4]);
			const float4 b4 = *reinterpret_cast<const float4 *>(&Bs[kk][tp * 4]);
			const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
			for (int i = 0; i < 4; i++)
#pragma unroll
				for (int j = 0; j < 4; j++) {
					if (IS_DOT)
						acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
					else {
						const float e = a[i] - b[j];
						acc[i][j] = fmaf(e, e, acc[i][j]);
					}
				}
		}
		__syncthreads();
	}
#pragma unroll
	for (int i = 0; i < 4; i++) {
		const int q = q0 + tq * 4 + i;
		if (q >= nq) continue;
#pragma unroll
		for (int j = 0; j < 4; j++) {
			const int pp = p0 + tp * 4 + j;
			if (pp < P) keys[(size_t)q * P + pp] = make_key(IS_DOT ? 1.0f - acc[i][j] : acc[i][j], (uint32_t)pp);
		}
	}
}

// ------------------------------------------------------------------ probe of a few queries (a12), nq < 8
// One warp per centroid: the row is read once (128-bit loads when d % 4 == 0) and scored against all nq queries held
// in shared memory; keys[q][p] = (d(q', c_p), p).  The select (launch_topk_merge) follows.  A lone query used to go
// through the streaming scan, whose two-level in-kernel select of k = nprobes keys over hundreds of tiny blocks cost
// more than reading the 12 MB table.
template <int IS_DOT>
__global__ void __launch_bounds__(256) probe_rows_kernel(const float *__restrict__ Qp, int nq, const float *__restrict__ C,
                                                         int P, int d, u64 *__restrict__ keys) {
	extern __shared__ __align__(16) float probe_qs[]; // [nq][d]
	for (int i = threadIdx.x; i < nq * d; i += blockDim.x) probe_qs[i] = Qp[i];
	__syncthreads();
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int row = blockIdx.x * (blockDim.x >> 5) + warp;
	if (row >= P) return;
	float acc[8];
#pragma unroll
	for (int j = 0; j < 8; j++) acc[j] = 0.f;
	const float *x = C + (size_t)row * d;
	if ((d & 3) == 0 && ((reinterpret_cast<uintptr_t>(C) & 15) == 0)) {
		const float4 *x4 = reinterpret_cast<const float4 *>(x);
		const float4 *q4 = reinterpret_cast<const float4 *>(probe_qs);
		const int n4 = d >> 2;
#pragma unroll 4
		for (int c = lane; c < n4; c += 32) {
			const float4 b = __ldg(x4 + c);
#pragma unroll
			for (int j = 0; j < 8; j++) {
				if (j < nq) {
					const float4 a = q4[j * n4 + c];
					if (IS_DOT) {
						acc[j] = fmaf(a.x, b.x, acc[j]), acc[j] = fmaf(a.y, b.y, acc[j]), acc[j] = fmaf(a.z, b.z, acc[j]),
						acc[j] = fmaf(a.w, b.w, acc[j]);
					} else {
						const float e0 = a.x - b.x, e1 = a.y - b.y, e2 = a.z - b.z, e3 = a.w - b.w;
						acc[j] = fmaf(e0, e0, acc[j]), acc[j] = fmaf(e1, e1, acc[j]), acc[j] = fmaf(e2, e2, acc[j]),
						acc[j] = fmaf(e3, e3, acc[j]);
					}
				}
			}
		}
	} else {
		for (int c = lane; c < d; c += 32) {
			const float b = __ldg(x + c);
#pragma unroll
			for (int j = 0; j < 8; j++) {
				if (j < nq) {
					const float a = probe_qs[j * d + c];
					if (IS_DOT)
						acc[j] = fmaf(a, b, acc[j]);
					else {
						const float e = a - b;
						acc[j] = fmaf(e, e, acc[j]);
					}
				}
			}
		}
	}
#pragma unroll
	for (int j = 0; j < 8; j++) {
		if (j < nq) {
			float v = acc[j];
			for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
			if (lane == 0) keys[(size_t)j * P + row] = make_key(IS_DOT ? 1.0f - v : v, (uint32_t)row);
		}
	}
}

// ------------------------------------------------------------------ per-query lookup table T1
// lut[q][m][c] = |q'_m - cb[m][c]|^2 (l2/cosine) or -(q'_m . cb[m][c]) (dot).
// grid (M, ceil(nq / QPB)); 256 threads = one codeword each; the codebook slice of sub-quantizer m is
// read once per block and reused for QPB queries.
static const int LUT_QPB = 16;
__global__ void __launch_bounds__(256) lut_kernel(const float *__restrict__ Qp, int nq, int d, int M, int dsub,
                                                  const float *__restrict__ codebook, int is_dot, int transposed,
                                                  float *__restrict__ lut) {
	extern __shared__ float lut_sm[]; // cb slice [256][dsub+1] + q sub-vectors [QPB][dsub]
	const int m = blockIdx.x;
	const int q0 = blockIdx.y * LUT_QPB;
	const int c = threadIdx.x;
	const int ld = dsub + 1;
	float *cbs = lut_sm;
	float *qs = lut_sm + 256 * ld;
	const float *cb = codebook + (size_t)m * 256 * dsub;
	for (int i = threadIdx.x; i < 256 * dsub; i += 256) cbs[(i / dsub) * ld + (i % dsub)] = cb[i];
	for (int i = threadIdx.x; i < LUT_QPB * dsub; i += 256) {
		int qq = q0 + i / dsub;
		qs[i] = qq < nq ? Qp[(size_t)qq * d + m * dsub + (i % dsub)] : 0.f;
	}
	__syncthreads();
	for (int j = 0; j < LUT_QPB && q0 + j < nq; j++) {
		float acc = 0.f;
		if (is_dot) {
			for (int i = 0; i < dsub; i++) acc = fmaf(-qs[j * dsub + i], cbs[c * ld + i], acc);
		} else {
			for (int i = 0; i < dsub; i++) {
				float e = qs[j * dsub + i] - cbs[c * ld + i];
				acc = fmaf(e, e, acc);
			}
		}
		// transposed ([q][256][M]): the layout of the bank-conflict-free scan (adc_scan_kernel<32>)
		if (transposed)
			lut[((size_t)(q0 + j) * 256 + c) * M + m] = acc;
		else
			lut[((size_t)(q0 + j) * M + m) * 256 + c] = acc;
	}
}

// ------------------------------------------------------------------ ADC scan
// One CTA per (probe slice s, query q): loads T1[q] into shared memory, then for each of its probed
// partitions streams the interleaved codes; lane r of a warp owns row r of a 32-row block.
template <int GRAN>
struct CodeWord;
template <>
struct CodeWord<16> {
	typedef uint4 T;
};
template <>
struct CodeWord<32> { // 16-byte granule, sub-quantizers XOR-rotated by the lane inside every block of 32
	typedef uint4 T;
};
template <>
struct CodeWord<4> {
	typedef uint32_t T;
};
template <>
struct CodeWord<1> {
	typedef uint8_t T;
};

__device__ __forceinline__ float lut4(const float *lut_m, uint32_t w) {
	// four consecutive sub-quantizers packed in one 32-bit word (little endian)
	return lut_m[w & 0xFFu] + lut_m[256 + ((w >> 8) & 0xFFu)] + lut_m[512 + ((w >> 16) & 0xFFu)] + lut_m[768 + (w >> 24)];
}

// Rotated layout (M % 32 == 0): stored position i of a 32-wide sub-quantizer block holds, for the row of lane l, the
// code of sub-quantizer (i ^ l), and the table is transposed ([256][M]).  The 32 lanes of a lookup then read the 32
// different sub-quantizers of the block: word index c*M + 32*jj + (i ^ l) = bank (i ^ l), whatever the codes are.
// `base` = table + 128*jj bytes, `i0` = first stored position of the four bytes of w, M4 = 4*M, lane4 = 4*lane.
template <uint32_t M4>
__device__ __forceinline__ float lut4_rot(const unsigned char *base, uint32_t w, int i0, uint32_t lane4) {
	const uint32_t c0 = w & 0xFFu, c1 = __byte_perm(w, 0, 0x4441), c2 = __byte_perm(w, 0, 0x4442), c3 = w >> 24;
	const float a = *reinterpret_cast<const float *>(base + c0 * M4 + ((uint32_t)((i0 + 0) << 2) ^ lane4));
	const float b = *reinterpret_cast<const float *>(base + c1 * M4 + ((uint32_t)((i0 + 1) << 2) ^ lane4));
	const float c = *reinterpret_cast<const float *>(base + c2 * M4 + ((uint32_t)((i0 + 2) << 2) ^ lane4));
	const float d = *reinterpret_cast<const float *>(base + c3 * M4 + ((uint32_t)((i0 + 3) << 2) ^ lane4));
	return (a + b) + (c + d);
}

// exact distance of (query qv, row of `key`) by one warp -> (distance, row) key on every lane (a14)
__device__ __forceinline__ u64 warp_exact_key(const float *__restrict__ X, int d, const float *__restrict__ qv, int metric,
                                              u64 key) {
	if (key == KEY_INVALID) return KEY_INVALID;
	const uint32_t row = key_row(key);
	const bool vec4 = (d % 4 == 0) && (((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(qv)) & 15) == 0);
	const float dist = warp_exact_distance(metric, qv, X + (size_t)row * d, d, vec4);
	return make_key(dist, row);
}

struct AdcParams {
	const float *lut_g; // [nq][M][256] per-query tables
	int M;
	const u64 *probe_keys; // [nq][probe_stride] (probe distance, partition) keys
	int nprobes_eff, probe_stride;
	const float *qn2;
	int metric;
	const long long *pad_offsets;
	const int32_t *slot_rows;
	const uint8_t *codes_il;
	const float *row_term;
	const uint32_t *allow;
	long long allow_bits;
	const uint32_t *deleted;
	int kc, cap, PS;
	u64 *partials;      // [nq][S][kc]
	unsigned *tickets;  // [nq], zero before the launch; the last CTA of a query zeroes its ticket again
	u64 *cand;          // [nq][kc] the kc best ADC keys (ascending)
	// refine + final select, run by the last CTA of a query
	const float *X;     // raw vectors
	int d;
	const float *Q;     // the caller's queries (not normalised)
	int refine;         // 0: ADC distances are final
	u64 *fin;           // [nq][fin_stride], first kc entries written here
	int fin_stride;
	int fuse_final;     // 1: also write the sorted top-k (no unindexed tail to merge)
	int k;
	long long label_base;
	int64_t *out_labels;
	float *out_dist;
	int32_t *out_counts;
};

// NB = M / 32 for the rotated layout (GRAN 32; compile-time so that every table offset is an immediate), else 0
template <int GRAN, int NB>
__global__ void __launch_bounds__(512) adc_scan_kernel(const __grid_constant__ AdcParams p) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	const int nwarps = blockDim.x >> 5;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int q = blockIdx.y;
	const int M = p.M, kc = p.kc, cap = p.cap, PS = p.PS;
	// slice s = (probe slot ps, sub-slice j): probes ps, ps+PS, .. and the j-th of J equal parts of each list,
	// so that a single query still fills the machine (32 probed lists -> PS*J CTAs)
	const int S = gridDim.x, s = blockIdx.x;
	const int ps = s % PS, j = s / PS, J = S / PS;
	float *lut = reinterpret_cast<float *>(smem_raw);                                 // [M][256]
	int *cnts = reinterpret_cast<int *>(smem_raw + (size_t)M * 1024);                 // [nwarps] (+ flag at [8])
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw + (size_t)M * 1024 + 64);            // [nwarps][cap]
	const uint32_t bar = tc::smem_u32(smem_raw + (size_t)M * 1024 + 48);

	// the query's table comes in as one bulk copy (TMA engine), 32 KB pieces on one mbarrier
	if (threadIdx.x == 0) {
		tc::mbar_init(bar, 1);
		tc::fence_barrier_init();
		tc::mbar_expect_tx(bar, (uint32_t)M * 1024u);
		const unsigned char *src = reinterpret_cast<const unsigned char *>(p.lut_g + (size_t)q * M * 256);
		for (uint32_t off = 0; off < (uint32_t)M * 1024u; off += 32768u) {
			uint32_t n = min(32768u, (uint32_t)M * 1024u - off);
			tc::bulk_load(tc::smem_u32(lut) + off, src + off, n, bar);
		}
	}
	__syncthreads();

	WarpSelect sel;
	sel.init(bufs + (size_t)warp * cap, cap, kc);
	const float qq = p.qn2[q];
	const int metric = p.metric;
	constexpr int GBYTES = GRAN == 32 ? 16 : GRAN;
	const int chunks = M / GBYTES;
	const uint32_t lane4 = lane << 2;
	typedef typename CodeWord<GRAN>::T word_t;
	bool lut_ready = false;

	if constexpr (GRAN == 32 && NB <= 3) {
		// Software-pipelined form of the loop below: a warp walks its blocks of all its lists as one sequence and
		// loads the code words of block i+1 before it looks up block i, so the HBM latency overlaps the lookups.
		constexpr int CH = NB * 2;
		int pi = ps - PS;
		long long b = 0, b1 = 0; // next block of the current list, end of the list
		float cst_list = 0.f;
		// advance to the warp's next block; false when the lists are exhausted
		auto next_block = [&](long long &blk, float &cst) -> bool {
			while (b >= b1) {
				pi += PS;
				if (pi >= p.nprobes_eff) return false;
				const u64 pk = p.probe_keys[(size_t)q * p.probe_stride + pi];
				if (pk == KEY_INVALID) continue;
				const int part = (int)key_row(pk);
				cst_list = (metric == METRIC_DOT) ? 1.0f : key_dist(pk) - qq;
				const long long pb0 = p.pad_offsets[part] >> 5, pnb = (p.pad_offsets[part + 1] >> 5) - pb0;
				b = pb0 + pnb * j / J + warp;
				b1 = pb0 + pnb * (j + 1) / J;
			}
			blk = b, cst = cst_list;
			b += nwarps;
			return true;
		};
		uint4 wn[CH];
		int row_n = -1;
		float rterm_n = 0.f, cst_n = 0.f;
		auto fetch = [&](long long blk) {
			const long long slot = (blk << 5) + lane;
			row_n = __ldg(p.slot_rows + slot);
			rterm_n = (metric == METRIC_DOT) ? 0.f : __ldg(p.row_term + slot);
			const uint4 *cp = reinterpret_cast<const uint4 *>(p.codes_il) + (size_t)blk * CH * 32 + lane;
#pragma unroll
			for (int u = 0; u < CH; u++) wn[u] = __ldg(cp + (size_t)u * 32);
		};
		long long blk;
		bool more = next_block(blk, cst_n);
		if (more) fetch(blk);
		tc::mbar_wait(bar, 0);
		lut_ready = true;
		while (more) {
			uint4 w[CH];
#pragma unroll
			for (int u = 0; u < CH; u++) w[u] = wn[u];
			const int row = row_n;
			const float rterm = rterm_n, cst = cst_n;
			more = next_block(blk, cst_n);
			if (more) fetch(blk);
			float a0 = 0.f, a1 = 0.f;
#pragma unroll
			for (int u = 0; u < CH; u++) {
				const unsigned char *lb = reinterpret_cast<const unsigned char *>(lut) + (u >> 1) * 128;
				const int i0 = (u & 1) * 16;
				a0 += lut4_rot<NB * 128u>(lb, w[u].x, i0, lane4) + lut4_rot<NB * 128u>(lb, w[u].y, i0 + 4, lane4);
				a1 += lut4_rot<NB * 128u>(lb, w[u].z, i0 + 8, lane4) + lut4_rot<NB * 128u>(lb, w[u].w, i0 + 12, lane4);
			}
			const float acc = a0 + a1;
			float dist;
			if (metric == METRIC_DOT)
				dist = 1.0f + acc;
			else {
				dist = acc + rterm + cst;
				if (metric == METRIC_COSINE) dist *= 0.5f;
			}
			bool live = row >= 0;
			if (live && p.deleted) live = !bit_test(p.deleted, row);
			if (live && p.allow) live = (row < p.allow_bits) && bit_test(p.allow, row);
			sel.push(live, make_key(dist, (uint32_t)row));
		}
	} else
	for (int pi = ps; pi < p.nprobes_eff; pi += PS) {
		const u64 pk = p.probe_keys[(size_t)q * p.probe_stride + pi];
		if (pk == KEY_INVALID) continue;
		const int part = (int)key_row(pk);
		const float cst = (metric == METRIC_DOT) ? 1.0f : key_dist(pk) - qq;
		const long long pb0 = p.pad_offsets[part] >> 5, pnb = (p.pad_offsets[part + 1] >> 5) - pb0;
		const long long b0 = pb0 + pnb * j / J, b1 = pb0 + pnb * (j + 1) / J;
		for (long long b = b0 + warp; b < b1; b += nwarps) {
			const long long slot = (b << 5) + lane;
			const int row = __ldg(p.slot_rows + slot);
			const float rterm = (metric == METRIC_DOT) ? 0.f : __ldg(p.row_term + slot);
			const word_t *cp = reinterpret_cast<const word_t *>(p.codes_il) + (size_t)b * chunks * 32 + lane;
			float acc = 0.f;
			if (GRAN == 32) {
				// chunk ch = stored positions 16*(ch&1) .. +15 of sub-quantizer block ch/2
				constexpr int CH = NB * 2;
				float a0 = 0.f, a1 = 0.f;
#pragma unroll
				for (int c0 = 0; c0 < CH; c0 += 6) {
					uint4 w[6];
#pragma unroll
					for (int u = 0; u < 6; u++)
						if (c0 + u < CH) w[u] = __ldg(reinterpret_cast<const uint4 *>(cp) + (size_t)(c0 + u) * 32);
					if (!lut_ready) {
						tc::mbar_wait(bar, 0);
						lut_ready = true;
					}
#pragma unroll
					for (int u = 0; u < 6; u++)
						if (c0 + u < CH) {
							const unsigned char *lb = reinterpret_cast<const unsigned char *>(lut) + ((c0 + u) >> 1) * 128;
							const int i0 = ((c0 + u) & 1) * 16;
							a0 += lut4_rot<NB * 128u>(lb, w[u].x, i0, lane4) + lut4_rot<NB * 128u>(lb, w[u].y, i0 + 4, lane4);
							a1 += lut4_rot<NB * 128u>(lb, w[u].z, i0 + 8, lane4) + lut4_rot<NB * 128u>(lb, w[u].w, i0 + 12, lane4);
						}
				}
				acc = a0 + a1;
			} else if (GRAN == 16) {
				// issue all loads of the row first (M/16 independent 128-bit loads)
				for (int c0 = 0; c0 < chunks; c0 += 6) {
					uint4 w[6];
#pragma unroll
					for (int u = 0; u < 6; u++)
						if (c0 + u < chunks) w[u] = __ldg(reinterpret_cast<const uint4 *>(cp) + (size_t)(c0 + u) * 32);
					if (!lut_ready) {
						tc::mbar_wait(bar, 0);
						lut_ready = true;
					}
#pragma unroll
					for (int u = 0; u < 6; u++)
						if (c0 + u < chunks) {
							const float *lm = lut + (size_t)(c0 + u) * 16 * 256;
							acc += lut4(lm, w[u].x) + lut4(lm + 1024, w[u].y) + lut4(lm + 2048, w[u].z) +
							       lut4(lm + 3072, w[u].w);
						}
				}
			} else {
				if (!lut_ready) {
					tc::mbar_wait(bar, 0);
					lut_ready = true;
				}
				if (GRAN == 4) {
					for (int c = 0; c < chunks; c++) {
						uint32_t w = __ldg(reinterpret_cast<const uint32_t *>(cp) + (size_t)c * 32);
						acc += lut4(lut + (size_t)c * 4 * 256, w);
					}
				} else {
					for (int c = 0; c < chunks; c++) {
						uint8_t w = __ldg(reinterpret_cast<const uint8_t *>(cp) + (size_t)c * 32);
						acc += lut[c * 256 + w];
					}
				}
			}
			float dist;
			if (metric == METRIC_DOT)
				dist = 1.0f + acc;
			else {
				dist = acc + rterm + cst;
				if (metric == METRIC_COSINE) dist *= 0.5f;
			}
			bool live = row >= 0;
			if (live && p.deleted) live = !bit_test(p.deleted, row);
			if (live && p.allow) live = (row < p.allow_bits) && bit_test(p.allow, row);
			sel.push(live, make_key(dist, (uint32_t)row));
		}
	}
	if (!lut_ready) tc::mbar_wait(bar, 0); // the copy must land before the CTA may exit
	sel.pad();
	block_bitonic_sort(bufs, nwarps * cap, IdentityMap()); // the CTA's kc best = the first kc keys
	if (warp == 0) {
		u64 *dst = p.partials + ((size_t)q * S + s) * kc;
		for (int i = lane; i < kc; i += 32) dst[i] = bufs[i];
	}
	// the last CTA of the query selects its kc best ADC candidates, refines them and (no unindexed tail) sorts the
	// final top-k: one launch instead of scan + merge + refine + merge
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) cnts[8] = (atomicAdd(p.tickets + q, 1u) == (unsigned)S - 1u);
	__syncthreads();
	if (!cnts[8]) return;
	if (threadIdx.x == 0) p.tickets[q] = 0u; // every other CTA of the query has arrived: ready for the next launch
	__threadfence();
	block_final_select(p.partials + (size_t)q * S * kc, (long long)S * kc, kc, cap, bufs, cnts, 0, (size_t)q, nullptr, nullptr,
	                   nullptr, p.cand);
	u64 *fin = p.fin + (size_t)q * p.fin_stride;
	for (int i = warp; i < kc; i += nwarps) {
		const u64 key = p.cand[(size_t)q * kc + i];
		const u64 res = p.refine ? warp_exact_key(p.X, p.d, p.Q + (size_t)q * p.d, metric, key) : key;
		if (lane == 0) fin[i] = res;
	}
	__syncthreads();
	if (p.fuse_final)
		block_final_select(fin, kc, p.k, cap, bufs, cnts, p.label_base, (size_t)q, p.out_labels, p.out_dist, p.out_counts,
		                   nullptr);
}

// ------------------------------------------------------------------ refine (a14): exact distance of the candidates
// One warp per candidate key; writes (exact distance, row) keys.  Same distance formulas as flat_scan.cu.
__global__ void refine_kernel(const float *__restrict__ X, int d, const float *__restrict__ Q, int metric,
                              const u64 *__restrict__ cand, int kc, int nq, u64 *__restrict__ out, int out_stride) {
	long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (w >= (long long)nq * kc) return;
	const int q = (int)(w / kc), i = (int)(w % kc);
	const u64 res = warp_exact_key(X, d, Q + (size_t)q * d, metric, cand[(size_t)q * kc + i]);
	if (lane_id() == 0) out[(size_t)q * out_stride + i] = res;
}

__global__ void copy_keys_strided_kernel(const u64 *__restrict__ src, int src_stride, int n_per, int nq,
                                         u64 *__restrict__ dst, int dst_stride, int dst_off) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (long long)nq * n_per) return;
	int q = (int)(t / n_per), i = (int)(t % n_per);
	dst[(size_t)q * dst_stride + dst_off + i] = src[(size_t)q * src_stride + i];
}

__global__ void keys_to_debug_kernel(const u64 *__restrict__ keys, int n, long long label_base, int64_t *labels,
                                     float *dists, int32_t *count) {
	int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i == 0) {
		int c = 0;
		while (c < n && keys[c] != KEY_INVALID) c++;
		*count = c;
	}
	if (i >= n) return;
	u64 k = keys[i];
	labels[i] = k == KEY_INVALID ? -1 : label_base + (long long)key_row(k);
	dists[i] = k == KEY_INVALID ? __int_as_float(0x7F800000) : key_dist(k);
}

// ------------------------------------------------------------------ index install: exchange layout -> scan layout
// slot -> (source position in the partition-major exchange arrays, row); padding slots get -1.
__global__ void build_slots_kernel(const long long *__restrict__ part_offsets, const long long *__restrict__ pad_offsets,
                                   int P, const int64_t *__restrict__ row_ids, long long label_base, long long n_pad,
                                   int32_t *__restrict__ slot_rows, long long *__restrict__ slot_src) {
	long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n_pad) return;
	int lo = 0, hi = P; // partition p with pad_offsets[p] <= s < pad_offsets[p+1]
	while (hi - lo > 1) {
		int mid = (lo + hi) >> 1;
		if (pad_offsets[mid] <= s) lo = mid; else hi = mid;
	}
	long long src = part_offsets[lo] + (s - pad_offsets[lo]);
	if (src < part_offsets[lo + 1]) {
		slot_src[s] = src;
		slot_rows[s] = (int32_t)(row_ids[src] - label_base);
	} else {
		slot_src[s] = -1;
		slot_rows[s] = -1;
	}
}

// codes[n][M] (exchange) -> interleaved [n_pad/32][M/gran][32][gran]; gran 32 = 16-byte granules with the
// sub-quantizers of every block of 32 XOR-rotated by the row's lane (lut4_rot)
__global__ void interleave_codes_kernel(const uint8_t *__restrict__ codes, const long long *__restrict__ slot_src, int M,
                                        int gran, long long n_pad, uint8_t *__restrict__ codes_il) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n_pad * M) return;
	long long s = t / M;
	int m = (int)(t % M);
	long long src = slot_src[s];
	uint8_t v = src >= 0 ? codes[(size_t)src * M + m] : 0;
	long long b = s >> 5;
	int r = (int)(s & 31);
	if (gran == 32) {
		m = (m & ~31) | ((m & 31) ^ r);
		gran = 16;
	}
	codes_il[(((size_t)b * (M / gran) + m / gran) * 32 + r) * gran + (m % gran)] = v;
}

// row_term[slot] = 2 * sum_m <centroid_p[m], cb[m][code_m]>  (l2 / cosine); one warp per slot
__global__ void row_term_kernel(const uint8_t *__restrict__ codes, const long long *__restrict__ slot_src,
                                const long long *__restrict__ pad_offsets, int P, const float *__restrict__ centroids,
                                const float *__restrict__ codebook, int d, int M, int dsub, long long n_pad,
                                float *__restrict__ row_term) {
	long long s = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (s >= n_pad) return;
	const unsigned lane = lane_id();
	long long src = slot_src[s];
	float acc = 0.f;
	if (src >= 0) {
		int lo = 0, hi = P;
		while (hi - lo > 1) {
			int mid = (lo + hi) >> 1;
			if (pad_offsets[mid] <= s) lo = mid; else hi = mid;
		}
		const float *cen = centroids + (size_t)lo * d;
		for (int i = lane; i < d; i += 32) {
			int m = i / dsub, j = i - m * dsub;
			int c = codes[(size_t)src * M + m];
			acc = fmaf(cen[i], codebook[((size_t)m * 256 + c) * dsub + j], acc);
		}
		for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
	}
	if (lane == 0) row_term[s] = 2.0f * acc;
}

// ------------------------------------------------------------------ trainer kernels
// Nearest centre of every row (k-means assignment / PQ encoding).  score = |c|^2 - 2 x.c (l2) or -x.c (dot);
// argmin, ties to the lower index.  blockIdx.z selects a sub-space: x columns [z*dsub, z*dsub+dd), centres
// C + z*nc*dd.  Register-tiled SIMT GEMM: 64 rows x 64 centres per block step, 256 threads x (4x4).
static const int NC_BM = 64, NC_BN = 64, NC_BK = 16;
__global__ void __launch_bounds__(256) nearest_center_kernel(const float *__restrict__ X, long long n, int ldx, int dd,
                                                              const float *__restrict__ C, int nc, int is_dot,
                                                              const float *__restrict__ sub_base, // nullable: subtract
                                                              const int32_t *__restrict__ sub_idx,  // sub_base[idx[row]]
                                                              int32_t *__restrict__ assign, long long assign_stride) {
	__shared__ float Xs[NC_BK][NC_BM + 1];
	__shared__ float Cs[NC_BK][NC_BN + 1];
	__shared__ float best_s[NC_BM][17];
	__shared__ int best_i[NC_BM][17];
	const int z = blockIdx.z;
	const int xoff = z * dd;
	const float *Cz = C + (size_t)z * nc * dd;
	const long long row0 = (long long)blockIdx.x * NC_BM;
	const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4; // tx: centre group, ty: row group
	float best[4];
	int bidx[4];
#pragma unroll
	for (int i = 0; i < 4; i++) best[i] = 3.4e38f, bidx[i] = 0;
	for (int c0 = 0; c0 < nc; c0 += NC_BN) {
		float acc[4][4];
		float cn[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
		for (int i = 0; i < 4; i++)
#pragma unroll
			for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
		for (int k0 = 0; k0 < dd; k0 += NC_BK) {
			for (int t = threadIdx.x; t < NC_BK * NC_BM; t += 256) {
				int r = t / NC_BK, kk = t % NC_BK;
				long long row = row0 + r;
				float v = 0.f;
				if (row < n && k0 + kk < dd) {
					v = X[(size_t)row * ldx + xoff + k0 + kk];
					if (sub_base) v -= sub_base[(size_t)sub_idx[row] * ldx + xoff + k0 + kk];
				}
				Xs[kk][r] = v;
			}
			for (int t = threadIdx.x; t < NC_BK * NC_BN; t += 256) {
				int c = t / NC_BK, kk = t % NC_BK;
				Cs[kk][c] = (c0 + c < nc && k0 + kk < dd) ? Cz[(size_t)(c0 + c) * dd + k0 + kk] : 0.f;
			}
			__syncthreads();
#pragma unroll
			for (int kk = 0; kk < NC_BK; kk++) {
				float xv[4], cv[4];
#pragma unroll
				for (int i = 0; i < 4; i++) xv[i] = Xs[kk][ty * 4 + i];
#pragma unroll
				for (int j = 0; j < 4; j++) cv[j] = Cs[kk][tx * 4 + j];
#pragma unroll
				for (int j = 0; j < 4; j++) cn[j] = fmaf(cv[j], cv[j], cn[j]);
#pragma unroll
				for (int i = 0; i < 4; i++)
#pragma unroll
					for (int j = 0; j < 4; j++) acc[i][j] = fmaf(xv[i], cv[j], acc[i][j]);
			}
			__syncthreads();
		}
#pragma unroll
		for (int j = 0; j < 4; j++) {
			int c = c0 + tx * 4 + j;
			if (c >= nc) continue;
#pragma unroll
			for (int i = 0; i < 4; i++) {
				float sc = is_dot ? -acc[i][j] : cn[j] - 2.0f * acc[i][j];
				if (sc < best[i]) best[i] = sc, bidx[i] = c;
			}
		}
	}
#pragma unroll
	for (int i = 0; i < 4; i++) {
		best_s[ty * 4 + i][tx] = best[i];
		best_i[ty * 4 + i][tx] = bidx[i];
	}
	__syncthreads();
	if (threadIdx.x < NC_BM) {
		int r = threadIdx.x;
		float b = best_s[r][0];
		int bi = best_i[r][0];
		for (int t = 1; t < 16; t++) {
			float s = best_s[r][t];
			int si = best_i[r][t];
			if (s < b || (s == b && si < bi)) b = s, bi = si;
		}
		if (row0 + r < n) assign[(size_t)z * assign_stride + row0 + r] = bi;
	}
}

// centre sums by atomics: sums[z][assign][dd] += x ; counts[z][assign]++
__global__ void accumulate_centers_kernel(const float *__restrict__ X, long long n, int ldx, int dd,
                                          const float *__restrict__ sub_base, const int32_t *__restrict__ sub_idx,
                                          const int32_t *__restrict__ assign, long long assign_stride, int nc,
                                          float *__restrict__ sums, int *__restrict__ counts) {
	const int z = blockIdx.z;
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n * dd) return;
	long long row = t / dd;
	int j = (int)(t % dd);
	int a = assign[(size_t)z * assign_stride + row];
	float v = X[(size_t)row * ldx + z * dd + j];
	if (sub_base) v -= sub_base[(size_t)sub_idx[row] * ldx + z * dd + j];
	atomicAdd(sums + ((size_t)z * nc + a) * dd + j, v);
	if (j == 0) atomicAdd(counts + (size_t)z * nc + a, 1);
}

__global__ void finish_centers_kernel(float *__restrict__ C, const float *__restrict__ sums, const int *__restrict__ counts,
                                      long long total, int dd) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= total) return;
	int cnt = counts[t / dd];
	if (cnt > 0) C[t] = sums[t] / (float)cnt; // an empty cluster keeps its previous centre (oracle_kmeans)
}

// rows picked at a stride -> contiguous sample (optionally L2-normalised: the cosine index lives on unit rows)
__global__ void gather_rows_kernel(const float *__restrict__ X, int d, long long n_out, long long stride, int normalise,
                                   float *__restrict__ out) {
	long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (r >= n_out) return;
	const unsigned lane = lane_id();
	const float *src = X + (size_t)(r * stride) * d;
	float s = 0.f;
	if (normalise) {
		for (int i = lane; i < d; i += 32) s = fmaf(src[i], src[i], s);
		for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
	}
	float sc = (normalise && s > 0.f) ? rsqrtf(s) : 1.f;
	for (int i = lane; i < d; i += 32) out[(size_t)r * d + i] = src[i] * sc;
}

__global__ void pack_codes_kernel(const int32_t *__restrict__ sub_assign, long long n, int M, uint8_t *__restrict__ codes) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n * M) return;
	long long row = t / M;
	int m = (int)(t % M);
	codes[t] = (uint8_t)sub_assign[(size_t)m * n + row];
}

__global__ void iota_kernel(int32_t *p, long long n, int32_t base) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t < n) p[t] = base + (int32_t)t;
}

// partition-major gather after the sort: codes_x[pos] = codes_rm[order[pos]], row_ids[pos] = base + order[pos]
__global__ void gather_sorted_kernel(const uint8_t *__restrict__ codes_rm, const int32_t *__restrict__ order, long long n,
                                     int M, long long label_base, uint8_t *__restrict__ codes_x,
                                     int64_t *__restrict__ row_ids) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= n * M) return;
	long long pos = t / M;
	int m = (int)(t % M);
	int32_t row = order[pos];
	codes_x[t] = codes_rm[(size_t)row * M + m];
	if (m == 0) row_ids[pos] = label_base + row;
}

__global__ void histogram_kernel(const int32_t *__restrict__ assign, long long n, int *__restrict__ hist) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t < n) atomicAdd(hist + assign[t], 1);
}

#include "ivfpq1.cuh"
#include "ivfpq_host.cuh"
#include "ivfpq_train.cuh"

} // namespace lb
