// flat_gemm_tsk_kernel.cuh — large-K variant (256 < K <= 768) of the TMEM-resident candidate kernel.
//
// One query tile (128 queries x K bf16 = K/2 TMEM columns, up to 384) stays in tensor memory for the CTA's
// life; X streams through shared memory in stages of KS 64-wide k-blocks.
//
// TMEM map: [0,128) the ONE 128 x 128 accumulator | [128, 128 + K/2) the query tile.  With 384 columns taken by the
// queries there is no room for a second N=128 accumulator, so the tile's MMAs (12 x 4 x 64 clk) and the drain
// alternate; the drain is kept short by having all 8 epilogue warps pull their 64 columns into registers at once and
// release the accumulator BEFORE filtering.  Round-2 experiments on C2, B = 4096 (LANCE_B200_TSK_DBG):
// MMAs alone 3.9 ms per batch, + TMA traffic 4.3 ms, + threshold filter 5.75 ms.  Tried and measured slower:
//   * a hybrid operand placement (k-blocks 0..7 of the queries in TMEM, the rest in shared memory as SS-form MMAs) to
//     make room for a second accumulator: the SS MMAs read 128 B/clk of shared memory next to the TMA's 64 B/clk
//     writes — 13.5 ms per batch (scripts/dev/attic/flat_gemm_tsk_hybrid_kernel.cuh.txt);
//   * 64-row tiles with two N=64 accumulators in the same 128 columns: an N=64 MMA re-reads the A operand from TMEM
//     twice as often per flop — 8.7 ms per batch (scripts/dev/attic/flat_gemm_tsk_n64_double_acc.cuh.txt).
// What did help: the issue loop (one elected lane, running descriptors: a k-block is four tcgen05.mma and a few
// uniform adds instead of ~36 instructions) and, marginally, multicast pairs.
//
// MC = true: launched as clusters of two CTAs that hold two DIFFERENT query tiles (g, g+1) and walk the SAME X tiles.
// Each CTA fetches half of every 128-row box (64 rows, tmX is then the 64-row map) and the TMA engine multicasts it
// into both CTAs' shared memory: every X byte leaves the L2 once per PAIR (measured: +1 %; the L2 was not the limit).
#pragma once
#include "flat_gemm_kernel.cuh"

namespace lb {

static constexpr int TSK_BN = GEMM_BN;                // rows of an X tile
static constexpr int TSK_KB_BYTES = GEMM_KB_BYTES;    // one 64-wide k-block of a 128-row tile (16 KB)

// BIAS: the l2 score is <q, x> - |x|^2/2; instead of three extra bf16 K columns (which push D = 766..768 to a 13th
// k-block that does not fit the 512 TMEM columns, and cost a 7th k-block at D = 384) the fp32 term is added to the
// accumulator columns in the epilogue: 8 broadcast 128-bit loads + 32 adds per 32-column chunk, off the MMA's path.
template <bool MC, bool BIAS = false>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
flat_gemm_tsk_kernel(const __grid_constant__ CUtensorMap tmX, const GemmParams p) {
	extern __shared__ uint8_t smem_raw[];
	const int warp = threadIdx.x >> 5;
	const int lane = threadIdx.x & 31;
	const uint32_t crank = MC ? tc::cluster_ctarank() : 0u;
	// MC: cluster c = blockIdx.x / 2 serves query groups 2*(c % (NG/2)) + {0,1} on row slice r = c / (NG/2)
	const int g = MC ? 2 * (int)((blockIdx.x >> 1) % (p.NG >> 1)) + (int)crank : (int)(blockIdx.x % p.NG);
	const int r = MC ? (int)((blockIdx.x >> 1) / (p.NG >> 1)) : (int)(blockIdx.x / p.NG);
	const int KC = p.nkb * 32; // TMEM columns of the query tile
	const uint32_t TSK_STAGE_BYTES = (uint32_t)p.KS * TSK_KB_BYTES;

	const uint32_t sbase = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
	uint8_t *sgen = smem_raw + (sbase - tc::smem_u32(smem_raw));
	const uint32_t Xs = sbase;
	const uint32_t ctl_off = (uint32_t)p.S * TSK_STAGE_BYTES;
	const uint32_t bars = sbase + ctl_off;
	// barrier slots (8 B each): full[S] | empty[S] | qready | tfull | (unused) | tempty | (unused)
	const uint32_t bar_full = bars, bar_empty = bars + 8u * p.S, bar_q = bars + 16u * p.S;
	const uint32_t bar_tfull = bar_q + 8, bar_tempty = bar_tfull + 16;
	uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sgen + ctl_off + 16u * p.S + 8 + 32);
	float *thr_s = reinterpret_cast<float *>(sgen + ctl_off + 16u * p.S + 8 + 32 + 16);
	int *ecnt = reinterpret_cast<int *>(thr_s + 128);               // [2][128]
	uint2 *eslots = reinterpret_cast<uint2 *>(ecnt + 2 * 128);      // [2][EB][128]
	for (int i = threadIdx.x; i < 2 * 128; i += GEMM_THREADS) ecnt[i] = 0;
	const bool has_work = p.tile_lo + r < p.tile_hi;

	if (warp == 0 && lane == 0) {
		tc::tma_prefetch_desc(&tmX);
		for (int s = 0; s < p.S; s++) {
			tc::mbar_init(bar_full + 8u * s, 1);
			tc::mbar_init(bar_empty + 8u * s, MC ? 2 : 1); // MC: a stage is free when BOTH CTAs' MMAs have read it
		}
		tc::mbar_init(bar_q, 8);
		tc::mbar_init(bar_tfull, 1);
		tc::mbar_init(bar_tempty, 8);
		tc::fence_barrier_init();
	}
	if (warp == 1) {
		tc::tmem_alloc(tc::smem_u32(tmem_slot), 512);
		tc::tmem_relinquish();
	}
	for (int i = threadIdx.x; i < 128; i += GEMM_THREADS) thr_s[i] = p.thr[(size_t)g * 128 + i];
	tc::tc_fence_before();
	__syncthreads();
	if (MC) tc::cluster_sync(); // the peer's barriers exist before anything is multicast at them
	tc::tc_fence_after();
	const uint32_t tmem_base = *tmem_slot;
	const uint32_t tmem_q = tmem_base + 128;

	if (warp == 0) {
		// ---------------- TMA producer: stages of KS boxes (128 rows x 64 columns each).  ONE thread runs the whole
		// loop (stage / phase kept as counters: no division, no per-stage election)
		if (lane == 0) {
			uint32_t s = 0, ph = 0;
			bool first = true;
			for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R) {
				const int row0 = (int)(tile * GEMM_BN);
				for (int kb0 = 0; kb0 < p.nkb; kb0 += p.KS) {
					const int nk = min(p.KS, p.nkb - kb0);
					const uint32_t full = bar_full + 8u * s, dst = Xs + s * TSK_STAGE_BYTES;
					tc::mbar_wait(bar_empty + 8u * s, ph ^ 1u);
					if ((p.dbg & 2) && !(first && ph == 0)) { // experiment: no TMA traffic after the first fill
						tc::mbar_arrive(full);
					} else {
						tc::mbar_expect_tx(full, (uint32_t)nk * TSK_KB_BYTES);
						for (int kk = 0; kk < nk; kk++) {
							if (MC) // my half of the box (rows 64*crank ..), into both CTAs
								tc::tma_load_2d_mc(dst + (uint32_t)kk * TSK_KB_BYTES + crank * (TSK_KB_BYTES / 2), &tmX, (kb0 + kk) * 64,
								                   row0 + (TSK_BN / 2) * (int)crank, full, (uint16_t)3);
							else
								tc::tma_load_2d(dst + (uint32_t)kk * TSK_KB_BYTES, &tmX, (kb0 + kk) * 64, row0, full);
						}
					}
					if (++s == (uint32_t)p.S) s = 0, ph ^= 1u, first = false;
				}
			}
		}
		__syncwarp();
	} else if (warp == 1) {
		// ---------------- MMA issuer: D (128 x 128) = Q (TMEM) x Xtile^T (smem), K streamed.  ncu (round 2) showed this warp
		// BUSY, not waiting: ~36 instructions of descriptor arithmetic, election and warp barriers per k-block.  Now the
		// warp stays converged (all values below are warp-uniform, so they live in uniform registers and an MMA is ONE
		// instruction with running descriptors) and one elected lane issues
		if (has_work) {
			const uint32_t idesc = tc::umma_idesc_bf16(128, TSK_BN);
			tc::mbar_wait(bar_q, 0);
			tc::tc_fence_after();
			const uint32_t d = tmem_base;
			const uint64_t desc0 = tc::umma_smem_desc(Xs);
			const uint64_t stage_step = (uint64_t)(TSK_STAGE_BYTES >> 4), kb_step = (uint64_t)(TSK_KB_BYTES >> 4);
			const int last_kb = p.nkb - 1, nk_last = p.ksteps_last;
			uint32_t s = 0, ph = 0, acc = 0;
			uint64_t sdesc = desc0;
			for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R, acc++) {
				tc::mbar_wait(bar_tempty, (acc & 1u) ^ 1u);
				tc::tc_fence_after();
				uint32_t qa = tmem_q;
				for (int kb0 = 0; kb0 < p.nkb; kb0 += p.KS) {
					const int nkk = min(p.KS, p.nkb - kb0);
					tc::mbar_wait(bar_full + 8u * s, ph);
					tc::tc_fence_after();
					if (tc::elect_one()) {
						uint64_t bd = sdesc;
						uint32_t qk = qa;
						for (int kk = 0; kk < nkk; kk++, bd += kb_step, qk += 32) {
							const int kb = kb0 + kk;
							if (kb != last_kb || nk_last == 4) {
								tc::umma_bf16_ts(d, qk, bd, idesc, (uint32_t)(kb != 0));
								tc::umma_bf16_ts(d, qk + 8, bd + 2, idesc, 1u);
								tc::umma_bf16_ts(d, qk + 16, bd + 4, idesc, 1u);
								tc::umma_bf16_ts(d, qk + 24, bd + 6, idesc, 1u);
							} else {
								for (int k4 = 0; k4 < nk_last; k4++)
									tc::umma_bf16_ts(d, qk + k4 * 8, bd + (uint64_t)(2 * k4), idesc, (uint32_t)((kb | k4) != 0));
							}
						}
						if (MC) tc::umma_commit_mc(bar_empty + 8u * s, (uint16_t)3);
						else tc::umma_commit(bar_empty + 8u * s);
						if (kb0 + nkk >= p.nkb) tc::umma_commit(bar_tfull);
					}
					__syncwarp();
					qa += (uint32_t)nkk * 32;
					if (++s == (uint32_t)p.S) s = 0, ph ^= 1u, sdesc = desc0;
					else sdesc += stage_step;
				}
			}
		}
	} else {
		const int e = (warp - 2) >> 2;
		const int quad = warp & 3;
		const int lrow = quad * 32 + lane;
		const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
		const int q = g * 128 + lrow;
		if (has_work) {
			// ---------------- queries: global bf16 -> registers -> TMEM; the two groups split the columns
			const uint4 *src = reinterpret_cast<const uint4 *>(reinterpret_cast<const uint16_t *>(p.Qs) + (size_t)q * p.Kpad);
			for (int c0 = e * 32; c0 < KC; c0 += 64) {
				uint32_t v[32];
#pragma unroll
				for (int j = 0; j < 8; j++) {
					uint4 w = __ldg(src + (c0 >> 2) + j);
					v[4 * j] = w.x, v[4 * j + 1] = w.y, v[4 * j + 2] = w.z, v[4 * j + 3] = w.w;
				}
				tc::tmem_st32(tmem_q + lane_base + (uint32_t)c0, v);
			}
			tc::tmem_st_wait();
			tc::tc_fence_before();
			__syncwarp();
			if (lane == 0) tc::mbar_arrive(bar_q);
		}
		// ---------------- epilogue: every warp pulls 64 columns (group e: [64e, 64e+64)) of its lane quadrant
		const float thr = thr_s[lrow];
		const EmitBuf eb{p.EB ? eslots + (size_t)(e * p.EB) * 128 + lrow : nullptr, ecnt + e * 128 + lrow};
		const uint32_t taddr = tmem_base + lane_base + (uint32_t)e * 64;
		uint32_t acc = 0;
		for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R, acc++) {
			tc::mbar_wait(bar_tfull, acc & 1u);
			tc::tc_fence_after();
			const long long row0 = tile * GEMM_BN + e * 64;
			uint32_t va[32], vb[32];
			if (!(p.dbg & 4)) {
				tc::tmem_ld32(taddr, va);
				tc::tmem_ld32(taddr + 32, vb);
				tmem_wait_regs(va);
				tmem_wait_regs(vb);
			}
			tc::tc_fence_before();
			__syncwarp();
			if (lane == 0) tc::mbar_arrive(bar_tempty);
			if (BIAS) {
				// rows beyond the table read as zero scores + whatever the padded bias holds (0): they are filtered by row
				const float4 *b4 = reinterpret_cast<const float4 *>(p.bias + row0);
#pragma unroll
				for (int j = 0; j < 8; j++) {
					const float4 ba = __ldg(b4 + j), bb = __ldg(b4 + 8 + j);
					va[4 * j] = __float_as_uint(__uint_as_float(va[4 * j]) + ba.x);
					va[4 * j + 1] = __float_as_uint(__uint_as_float(va[4 * j + 1]) + ba.y);
					va[4 * j + 2] = __float_as_uint(__uint_as_float(va[4 * j + 2]) + ba.z);
					va[4 * j + 3] = __float_as_uint(__uint_as_float(va[4 * j + 3]) + ba.w);
					vb[4 * j] = __float_as_uint(__uint_as_float(vb[4 * j]) + bb.x);
					vb[4 * j + 1] = __float_as_uint(__uint_as_float(vb[4 * j + 1]) + bb.y);
					vb[4 * j + 2] = __float_as_uint(__uint_as_float(vb[4 * j + 2]) + bb.z);
					vb[4 * j + 3] = __float_as_uint(__uint_as_float(vb[4 * j + 3]) + bb.w);
				}
			}
			if (p.seed) {
				// threshold seeding pass (flat_gemm.cu "Seeding"): nothing is emitted, the best live score of this group's
				// 64 rows and that row go to cand[q][2 (tile - tile_lo) + e]
				float best = -INFINITY;
				uint32_t best_row = 0u;
				gemm_chunk_best(p, va, row0, best, best_row);
				gemm_chunk_best(p, vb, row0 + 32, best, best_row);
				p.cand[(size_t)q * p.CAP + (size_t)(2 * (tile - p.tile_lo) + e)] = make_uint2(__float_as_uint(best), best_row);
				continue;
			}
			if (!(p.dbg & 5)) {
				gemm_process_chunk_inline(p, va, thr, q, row0, eb);
				gemm_process_chunk_inline(p, vb, thr, q, row0 + 32, eb);
			}
		}
		if (p.EB) gemm_flush(p, q, eb);
	}
	tc::tc_fence_before();
	__syncthreads();
	if (MC) tc::cluster_sync(); // nobody leaves while the peer may still multicast into this CTA or arrive on its barriers
	if (warp == 1) {
		tc::tc_fence_after();
		tc::tmem_dealloc(tmem_base, 512);
	}
}

static inline size_t gemm_tsk_smem_bytes(int S, int KS, int EB) {
	return 1024 + (size_t)S * KS * TSK_KB_BYTES + 16u * S + 8 + 32 + 16 + 128 * 4 + 2 * 128 * 4 + (size_t)2 * EB * 128 * 8;
}

} // namespace lb
