// flat_gemm_ts_kernel.cuh — query-stationary-in-TMEM variant of the candidate kernel, for K <= 256.
//
// Same contract as flat_gemm_kernel (flat_gemm_kernel.cuh), different operand placement: the QT resident query
// tiles live in TENSOR MEMORY (bf16 pairs, lane = query, 32-bit column = two consecutive K elements) and feed
// tcgen05.mma as the A operand from TMEM, so an MMA reads only the X tile from shared memory (4 KB per 64-clk
// instruction = 64 B/clk instead of the 128 B/clk of the all-shared form, which sits exactly on the
// shared-memory bandwidth limit).  Shared memory then holds nothing but X stages (up to 6 x 128 rows x K).
// TMEM map (512 columns): [0,128) accumulator 0 | [128,256) accumulator 1 | [256, 256 + QT*K/2) query tiles.
//
// Warp roles as in the all-shared kernel; the 8 epilogue warps first copy their queries global -> registers ->
// TMEM (tcgen05.st), then drain accumulators: group e owns accumulator buffer e.
#pragma once
#include "flat_gemm_kernel.cuh"

namespace lb {

// SEED = true is the threshold seeding pass (flat_gemm.cu "seeding"): nothing is emitted; the epilogue reduces every
// 128-row tile to the best score of its live rows and stores it with its row at cand[q][tile - tile_lo], so the pass runs at
// the speed of its MMAs.  The k-th largest of those maxima is the score of k distinct rows, i.e. a lower bound of the k-th
// best score over the seeded rows, and almost always equal to it (the k best of 131 072 rows rarely share a tile).
template <int NKB, bool SEED>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
flat_gemm_ts_kernel(const __grid_constant__ CUtensorMap tmX, const GemmParams p) {
	extern __shared__ uint8_t smem_raw[];
	const int warp = threadIdx.x >> 5;
	const int lane = threadIdx.x & 31;
	const int g = blockIdx.x % p.NG;
	const int r = blockIdx.x / p.NG;
	constexpr uint32_t stage_bytes = NKB * GEMM_KB_BYTES;
	const int KC = NKB * 32; // TMEM columns per query tile

	const uint32_t sbase = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
	uint8_t *sgen = smem_raw + (sbase - tc::smem_u32(smem_raw));
	const uint32_t Xs = sbase;
	const uint32_t ctl_off = (uint32_t)p.S * stage_bytes;
	const uint32_t bars = sbase + ctl_off;
	// barrier slots (8 B each): full[S] | empty[S] | qready | tfull[2] | tempty[2]
	const uint32_t bar_full = bars, bar_empty = bars + 8u * p.S, bar_q = bars + 16u * p.S;
	const uint32_t bar_tfull = bar_q + 8, bar_tempty = bar_tfull + 16;
	uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sgen + ctl_off + 16u * p.S + 8 + 32);
	float *thr_s = reinterpret_cast<float *>(sgen + ctl_off + 16u * p.S + 8 + 32 + 16);
	// emit buffers: counters [2][QT][128] then slots [2][QT][EB][128]
	int *ecnt = reinterpret_cast<int *>(thr_s + p.QT * 128);
	uint2 *eslots = reinterpret_cast<uint2 *>(ecnt + 2 * p.QT * 128);
	// parked-chunk slots of the 8 epilogue warps (gemm_park_chunk), behind the emit buffers
	uint32_t *pend_all = reinterpret_cast<uint32_t *>(eslots + (size_t)2 * p.QT * p.EB * 128);
	for (int i = threadIdx.x; i < 8 * GEMM_PEND_WORDS; i += GEMM_THREADS) pend_all[i] = 0u;
	for (int i = threadIdx.x; i < 2 * p.QT * 128; i += GEMM_THREADS) ecnt[i] = 0;
	const bool has_work = p.tile_lo + r < p.tile_hi;

	if (warp == 0 && lane == 0) {
		tc::tma_prefetch_desc(&tmX);
		for (int s = 0; s < p.S; s++) {
			tc::mbar_init(bar_full + 8u * s, 1);
			tc::mbar_init(bar_empty + 8u * s, 1);
		}
		tc::mbar_init(bar_q, 8);
		for (int b = 0; b < 2; b++) {
			tc::mbar_init(bar_tfull + 8u * b, 1);
			tc::mbar_init(bar_tempty + 8u * b, 4);
		}
		tc::fence_barrier_init();
	}
	if (warp == 1) {
		tc::tmem_alloc(tc::smem_u32(tmem_slot), 512);
		tc::tmem_relinquish();
	}
	for (int i = threadIdx.x; i < p.QT * 128; i += GEMM_THREADS) thr_s[i] = p.thr[(size_t)g * p.QT * 128 + i];
	tc::tc_fence_before();
	__syncthreads();
	tc::tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;
	const uint32_t tmem_q = tmem_base + 256;

	if (warp == 0) {
		// ---------------- TMA producer: X tiles only (whole warp converged, one elected lane issues)
		uint32_t it = 0;
		for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R, it++) {
			const uint32_t s = it % p.S, ph = (it / p.S) & 1u;
			tc::mbar_wait(bar_empty + 8u * s, ph ^ 1u);
			if (tc::elect_one()) {
				tc::mbar_expect_tx(bar_full + 8u * s, stage_bytes);
#pragma unroll
				for (int kb = 0; kb < NKB; kb++)
					tc::tma_load_2d(Xs + s * stage_bytes + (uint32_t)kb * GEMM_KB_BYTES, &tmX, kb * 64, (int)(tile * GEMM_BN),
					                bar_full + 8u * s);
			}
			__syncwarp();
		}
	} else if (warp == 1) {
		if (has_work) {
			// ---------------- MMA issuer: D[acc] = Q[t] (TMEM) x X[stage]^T (smem); the warp stays converged, one
			// elected lane issues the instructions
			const uint32_t idesc = tc::umma_idesc_bf16(128, GEMM_BN);
			tc::mbar_wait(bar_q, 0);
			tc::tc_fence_after();
			uint32_t it = 0, acc = 0;
			for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R, it++) {
				const uint32_t s = it % p.S, ph = (it / p.S) & 1u;
				tc::mbar_wait(bar_full + 8u * s, ph);
				tc::tc_fence_after();
				const uint64_t bdesc = tc::umma_smem_desc(Xs + s * stage_bytes);
				for (int t = 0; t < p.QT; t++, acc++) {
					const uint32_t b = acc & 1u, n = acc >> 1;
					tc::mbar_wait(bar_tempty + 8u * b, (n & 1u) ^ 1u);
					tc::tc_fence_after();
					if (tc::elect_one()) {
						const uint32_t d = tmem_base + b * GEMM_BN;
						const uint32_t qa = tmem_q + (uint32_t)(t * KC);
#pragma unroll
						for (int kb = 0; kb < NKB; kb++) {
#pragma unroll
							for (int k4 = 0; k4 < 4; k4++) {
								if (kb == NKB - 1 && k4 >= p.ksteps_last) break;
								tc::umma_bf16_ts(d, qa + kb * 32 + k4 * 8, bdesc + (uint64_t)((kb * GEMM_KB_BYTES + k4 * 32) >> 4),
								                 idesc, (uint32_t)((kb | k4) != 0));
							}
						}
						tc::umma_commit(bar_tfull + 8u * b);
						if (t == p.QT - 1) tc::umma_commit(bar_empty + 8u * s);
					}
					__syncwarp();
				}
			}
		}
	} else {
		const int e = (warp - 2) >> 2;
		const int quad = warp & 3;
		const int lrow = quad * 32 + lane;
		const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
		if (has_work) {
			// ---------------- queries: global bf16 -> registers -> TMEM (this thread's lane = its query)
			for (int t = e; t < p.QT; t += 2) {
				const uint4 *src = reinterpret_cast<const uint4 *>(reinterpret_cast<const uint16_t *>(p.Qs) +
				                                                   (size_t)((g * p.QT + t) * 128 + lrow) * p.Kpad);
				for (int c0 = 0; c0 < KC; c0 += 32) {
					uint32_t v[32];
#pragma unroll
					for (int j = 0; j < 8; j++) {
						uint4 w = __ldg(src + (c0 >> 2) + j);
						v[4 * j] = w.x, v[4 * j + 1] = w.y, v[4 * j + 2] = w.z, v[4 * j + 3] = w.w;
					}
					tc::tmem_st32(tmem_q + lane_base + (uint32_t)(t * KC + c0), v);
				}
			}
			tc::tmem_st_wait();
			tc::tc_fence_before();
			__syncwarp();
			if (lane == 0) tc::mbar_arrive(bar_q);
		}
		// ---------------- epilogue: group e drains accumulator buffer e
		uint32_t acc = 0;
		for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R) {
			for (int t = 0; t < p.QT; t++, acc++) {
				if ((int)(acc & 1u) != e) continue;
				const uint32_t n = acc >> 1;
				tc::mbar_wait(bar_tfull + 8u * e, n & 1u);
				tc::tc_fence_after();
				const float thr = thr_s[t * 128 + lrow];
				const int q = (g * p.QT + t) * 128 + lrow;
				const uint32_t taddr = tmem_base + lane_base + (uint32_t)e * GEMM_BN;
				const long long row0 = tile * GEMM_BN;
				const EmitBuf eb{p.EB ? eslots + (size_t)((e * p.QT + t) * p.EB) * 128 + lrow : nullptr,
				                 ecnt + (e * p.QT + t) * 128 + lrow};
				uint32_t *pend = pend_all + (warp - 2) * GEMM_PEND_WORDS;
				uint32_t va[32], vb[32];
				if constexpr (SEED) {
					float best = -INFINITY;
					uint32_t best_row = 0u;
					tc::tmem_ld32(taddr, va);
					tmem_wait_regs(va);
					tc::tmem_ld32(taddr + 32, vb);
					gemm_chunk_best(p, va, row0, best, best_row);
					tmem_wait_regs(vb);
					tc::tmem_ld32(taddr + 64, va);
					gemm_chunk_best(p, vb, row0 + 32, best, best_row);
					tmem_wait_regs(va);
					tc::tmem_ld32(taddr + 96, vb);
					gemm_chunk_best(p, va, row0 + 64, best, best_row);
					tmem_wait_regs(vb);
					tc::tc_fence_before();
					__syncwarp();
					if (lane == 0) tc::mbar_arrive(bar_tempty + 8u * e);
					gemm_chunk_best(p, vb, row0 + 96, best, best_row);
					p.cand[(size_t)q * p.CAP + (size_t)(tile - p.tile_lo)] = make_uint2(__float_as_uint(best), best_row);
					continue;
				}
				tc::tmem_ld32(taddr, va);
				tmem_wait_regs(va);
				tc::tmem_ld32(taddr + 32, vb);
				gemm_park_chunk(p, va, thr, q, row0, eb, pend, 0);
				tmem_wait_regs(vb);
				tc::tmem_ld32(taddr + 64, va);
				gemm_park_chunk(p, vb, thr, q, row0 + 32, eb, pend, 1);
				tmem_wait_regs(va);
				tc::tmem_ld32(taddr + 96, vb);
				gemm_park_chunk(p, va, thr, q, row0 + 64, eb, pend, 2);
				tmem_wait_regs(vb);
				tc::tc_fence_before();
				__syncwarp();
				if (lane == 0) tc::mbar_arrive(bar_tempty + 8u * e);
				// the accumulator is free again: everything below overlaps the next MMAs
				gemm_process_chunk(p, vb, thr, q, row0 + 96, eb);
				gemm_drain_pending(p, thr, q, row0, eb, pend);
			}
		}
		if (p.EB) {
			for (int t = 0; t < p.QT; t++) {
				const EmitBuf eb{eslots + (size_t)((e * p.QT + t) * p.EB) * 128 + lrow, ecnt + (e * p.QT + t) * 128 + lrow};
				gemm_flush(p, (g * p.QT + t) * 128 + lrow, eb);
			}
		}
	}
	tc::tc_fence_before();
	__syncthreads();
	if (warp == 1) {
		tc::tc_fence_after();
		tc::tmem_dealloc(tmem_base, 512);
	}
}

static inline size_t gemm_ts_smem_bytes(int QT, int nkb, int S, int EB) {
	return 1024 + (size_t)S * nkb * GEMM_KB_BYTES + 16u * S + 8 + 32 + 16 + (size_t)QT * 128 * 4 +
	       (size_t)2 * QT * 128 * 4 + (size_t)2 * QT * EB * 128 * 8 + (size_t)8 * GEMM_PEND_WORDS * 4;
}

} // namespace lb
