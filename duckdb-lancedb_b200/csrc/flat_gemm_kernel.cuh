// flat_gemm_kernel.cuh — the tcgen05 candidate kernel of the batched flat search (included by flat_gemm.cu).
//
// score[q][row] = <Qs[q], Xs[row]> over bf16 shadow operands (fp32 accumulate in TMEM); every (q,row) whose
// score reaches the query's threshold is appended to that query's candidate list.  No score matrix is ever
// written: the epilogue reads each 128x128 accumulator out of TMEM once, takes a 3-input max over 32-column
// chunks and only walks a chunk when its max passes the threshold (rare once the threshold is warm).
//
// Mapping ("query-stationary"): a CTA keeps QT query tiles (QT x 128 queries x K, bf16, 128B-swizzled) resident
// in shared memory for its whole life and streams X tiles of BN=128 rows through S TMA stages, so each X byte
// fetched from L2/HBM is used for QT*128 queries.  UMMA: A = queries (M=128 -> TMEM lanes = queries),
// B = rows (N=128 -> TMEM columns = rows): an epilogue thread owns ONE query, its threshold sits in a register.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..9 = two epilogue groups of 4 warps (warp%4 = TMEM lane quadrant); accumulators are
// quadruple-buffered in TMEM (4 x 128 columns), group e drains accumulators with (index & 1) == e.
#pragma once
#include "tc_ptx.cuh"

namespace lb {

struct GemmParams {
	int nkb;          // 64-wide k-blocks per row (Kpad / 64)
	int ksteps_last;  // MMA k-steps (16 wide) to issue in the last k-block (1..4); the rest is zero padding
	int QT;           // resident query tiles per CTA
	int KS;           // k-blocks per X stage
	int S;            // X stages
	int NG;           // query groups (QT*128 queries each); gridDim.x = NG * R
	int R;            // CTAs per query group
	long long tile_lo, tile_hi; // X tiles [tile_lo, tile_hi) of 128 rows
	long long n_rows;
	const float *thr;     // [NG*QT*128] a score >= thr passes
	uint32_t *cnt;        // [NG*QT*128] candidates appended (may exceed CAP: overflow)
	uint2 *cand;          // [NG*QT*128][CAP] {score bits, row}
	int CAP;
	const uint32_t *deleted; // tombstones or null
	const uint32_t *allow;   // prefilter bitmap or null
	long long allow_bits;
	const void *Qs;          // bf16 [NG*QT*128][Kpad] query tiles in global memory (TMEM-resident variant)
	int Kpad;
	int EB;                  // per-(epilogue group, query) shared-memory emit buffer entries (0 = append directly)
	const float *bias;       // K-streamed kernel, l2 with the bias epilogue: -|x|^2/2 per row (fp32), else null
	int seed;                // K-streamed kernel: 1 = threshold seeding pass (per-half-tile maxima into cand, nothing emitted)
	int dbg;                 // LANCE_B200_TSK_DBG (ceiling experiments): 1 = epilogue drains but does not filter, 2 = X stages
	                         // are loaded once and re-used (no TMA traffic), 4 = epilogue does not even read the accumulator
};

static constexpr int GEMM_BN = 128;
static constexpr int GEMM_NBUF = 4;
static constexpr int GEMM_THREADS = 320;
static constexpr int GEMM_KB_BYTES = 128 * 128; // one 64-wide k-block of a 128-row tile

// Thread-private emit buffer in shared memory: slot-major ([slot][128 threads]) so that the lanes of a warp
// never collide.  A hit costs one shared store; the global append (one atomicAdd for the whole buffer, whose
// latency would otherwise stall the epilogue warp on every hit) happens when the buffer fills and at kernel end.
struct EmitBuf {
	uint2 *slots; // &buf[0][thread] or null
	int *count;   // &cnt[thread]
};

__device__ __noinline__ void gemm_flush(const GemmParams &p, int q, const EmitBuf &eb) {
	const int n = *eb.count;
	if (n == 0) return;
	const uint32_t pos = atomicAdd(p.cnt + q, (uint32_t)n);
	for (int i = 0; i < n; i++)
		if (pos + i < (uint32_t)p.CAP) p.cand[(size_t)q * p.CAP + pos + i] = eb.slots[i * 128];
	*eb.count = 0;
}

// rows [row0, row0+32) (row0 % 32 == 0) that may be emitted: inside the table, not deleted, allowed by the prefilter
__device__ __forceinline__ uint32_t gemm_chunk_valid(const GemmParams &p, long long row0) {
	const long long left = p.n_rows - row0; // beyond the table TMA filled zeros
	if (left <= 0) return 0u;
	uint32_t m = left < 32 ? (1u << (int)left) - 1u : 0xFFFFFFFFu;
	if (p.deleted) m &= ~__ldg(p.deleted + (row0 >> 5));
	if (p.allow) {
		uint32_t a = 0u;
		if (row0 < p.allow_bits) {
			a = __ldg(p.allow + (row0 >> 5));
			const long long al = p.allow_bits - row0;
			if (al < 32) a &= (1u << (int)al) - 1u;
		}
		m &= a;
	}
	return m;
}

// Bulk emit of one 32-column chunk for the lanes (queries) that have hits (`hits` = passing columns, already
// masked by gemm_chunk_valid).  Out of line and fed through a local-memory copy of the chunk, so that the hot
// loop keeps its registers; the work per lane is proportional to its own hits (a loop over the set bits), and the
// buffered + new hits go to the query's list with ONE atomicAdd when the thread's shared-memory buffer is full.
__device__ __noinline__ void gemm_emit_chunk(const GemmParams &p, const uint32_t *r, uint32_t hits, int q, long long row0,
                                             uint2 *slots, int *count) {
	const int h = __popc(hits);
	const uint32_t rbase = (uint32_t)row0;
	int n = slots ? *count : 0;
	if (slots && n + h <= p.EB) {
		while (hits) {
			const int c = __ffs(hits) - 1;
			hits &= hits - 1;
			slots[n * 128] = make_uint2(r[c], rbase + c);
			n++;
		}
		*count = n;
		return;
	}
	uint32_t pos = atomicAdd(p.cnt + q, (uint32_t)(n + h));
	uint2 *dst = p.cand + (size_t)q * p.CAP;
	for (int i = 0; i < n; i++, pos++)
		if (pos < (uint32_t)p.CAP) dst[pos] = slots[i * 128];
	if (slots) *count = 0;
	while (hits) {
		const int c = __ffs(hits) - 1;
		hits &= hits - 1;
		if (pos < (uint32_t)p.CAP) dst[pos] = make_uint2(r[c], rbase + c);
		pos++;
	}
}

// One hit of one lane (the sparse case: a warm threshold passes a few scores per query and segment).
__device__ __noinline__ void gemm_emit(const GemmParams &p, int q, long long row, float score, uint2 *slots, int *count) {
	if (row >= p.n_rows) return; // TMA zero-fill beyond the table
	if (p.deleted && bit_test(p.deleted, row)) return;
	if (p.allow && !(row < p.allow_bits && bit_test(p.allow, row))) return;
	const uint2 e = make_uint2(__float_as_uint(score), (uint32_t)row);
	if (!slots) {
		uint32_t pos = atomicAdd(p.cnt + q, 1u);
		if (pos < (uint32_t)p.CAP) p.cand[(size_t)q * p.CAP + pos] = e;
		return;
	}
	int n = *count;
	if (n == p.EB) {
		const uint32_t pos = atomicAdd(p.cnt + q, (uint32_t)n);
		for (int i = 0; i < n; i++)
			if (pos + i < (uint32_t)p.CAP) p.cand[(size_t)q * p.CAP + pos + i] = slots[i * 128];
		n = 0;
	}
	slots[n * 128] = e;
	*count = n + 1;
}

__device__ __forceinline__ void tmem_wait_regs(uint32_t (&r)[32]) {
	// tcgen05.wait::ld with the destination registers as in/out operands, so that no use is scheduled above it
	asm volatile("tcgen05.wait::ld.sync.aligned;"
	             : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
	               "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
	               "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]),
	               "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]),
	               "+r"(r[30]), "+r"(r[31])::"memory");
}

__device__ __forceinline__ void gemm_process_chunk(const GemmParams &p, const uint32_t (&r)[32], float thr, int q,
                                                   long long row0, const EmitBuf &eb) {
	float a[10];
#pragma unroll
	for (int i = 0; i < 10; i++)
		a[i] = tc::fmax3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
	float m1 = tc::fmax3(a[0], a[1], a[2]), m2 = tc::fmax3(a[3], a[4], a[5]), m3 = tc::fmax3(a[6], a[7], a[8]);
	float m4 = tc::fmax3(a[9], __uint_as_float(r[30]), __uint_as_float(r[31]));
	float m = fmaxf(tc::fmax3(m1, m2, m3), m4);
	if (m >= thr) {
		// rare once the threshold is warm.  Few lanes here (the steady state: one query, one score): walk the triple
		// maxima, a handful of instructions, because this epilogue is the kernel's critical path.  Many lanes (cold
		// threshold, seed segment): the branch-free bulk routine, whose cost does not grow with the number of hits.
		if (__popc(__activemask()) <= 4) {
#pragma unroll
			for (int i = 0; i < 10; i++) {
				if (a[i] >= thr) {
#pragma unroll
					for (int j = 0; j < 3; j++) {
						const float v = __uint_as_float(r[3 * i + j]);
						if (v >= thr) gemm_emit(p, q, row0 + 3 * i + j, v, eb.slots, eb.count);
					}
				}
			}
#pragma unroll
			for (int i = 30; i < 32; i++) {
				const float v = __uint_as_float(r[i]);
				if (v >= thr) gemm_emit(p, q, row0 + i, v, eb.slots, eb.count);
			}
		} else {
			uint32_t hits = 0u;
#pragma unroll
			for (int c = 0; c < 32; c++) hits |= (__uint_as_float(r[c]) >= thr ? 1u : 0u) << c;
			hits &= gemm_chunk_valid(p, row0);
			if (hits) {
				uint32_t tmp[32];
#pragma unroll
				for (int c = 0; c < 32; c++) tmp[c] = r[c];
				gemm_emit_chunk(p, tmp, hits, q, row0, eb.slots, eb.count);
			}
		}
	}
}

// Deferred form for the kernel whose accumulator release sits on the critical path (flat_gemm_ts_kernel: one
// accumulator's MMAs take 512 clocks, about what its drain takes, so whatever a hit costs BEFORE the release stalls
// the tensor pipe — measured: emission cost 0.6 ms of a 3.1 ms batch on a 12.5 M-row shard).  A chunk with a hit is
// only parked: the hitting lanes copy their 32 scores into one of the warp's GEMM_NPEND shared-memory slots
// ([33 words]: scores + (lane | chunk << 8)); gemm_drain_pending emits them after the accumulator has been released.
// Returns false when the slots are full (the caller falls back to gemm_process_chunk).
static constexpr int GEMM_NPEND = 4;
static constexpr int GEMM_PEND_WORDS = GEMM_NPEND * 33 + 1; // per epilogue warp, last word = count

__device__ __forceinline__ float gemm_chunk_max(const uint32_t (&r)[32]) {
	float a[10];
#pragma unroll
	for (int i = 0; i < 10; i++)
		a[i] = tc::fmax3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
	float m1 = tc::fmax3(a[0], a[1], a[2]), m2 = tc::fmax3(a[3], a[4], a[5]), m3 = tc::fmax3(a[6], a[7], a[8]);
	float m4 = tc::fmax3(a[9], __uint_as_float(r[30]), __uint_as_float(r[31]));
	return fmaxf(tc::fmax3(m1, m2, m3), m4);
}

// Seeding pass: the best live score of a 32-row chunk and the row that holds it (lowest row on equal scores), folded
// into the running (best, best_row) of the tile.
__device__ __forceinline__ void gemm_chunk_best(const GemmParams &p, const uint32_t (&v)[32], long long row0, float &best,
                                                uint32_t &best_row) {
	const uint32_t live = gemm_chunk_valid(p, row0);
	float m;
	if (live == 0xFFFFFFFFu) {
		m = gemm_chunk_max(v);
	} else {
		m = -INFINITY;
#pragma unroll
		for (int j = 0; j < 32; j++)
			if ((live >> j) & 1u) m = fmaxf(m, __uint_as_float(v[j]));
	}
	if (m > best) {
		int idx = 0;
#pragma unroll
		for (int j = 31; j >= 0; j--)
			if (((live >> j) & 1u) && __uint_as_float(v[j]) == m) idx = j;
		best = m, best_row = (uint32_t)row0 + (uint32_t)idx;
	}
}

__device__ __forceinline__ void gemm_park_chunk(const GemmParams &p, const uint32_t (&r)[32], float thr, int q, long long row0,
                                                const EmitBuf &eb, uint32_t *pend, int chunk) {
	const bool hit = gemm_chunk_max(r) >= thr;
	const unsigned hm = __ballot_sync(0xFFFFFFFFu, hit);
	if (hm == 0u) return; // the common case: one vote
	const int lane = threadIdx.x & 31;
	const int base = (int)pend[GEMM_NPEND * 33];
	const int nh = __popc(hm);
	if (base + nh <= GEMM_NPEND) {
		if (hit) {
			uint32_t *dst = pend + (base + __popc(hm & ((1u << lane) - 1u))) * 33;
#pragma unroll
			for (int c = 0; c < 32; c++) dst[c] = r[c];
			dst[32] = (uint32_t)lane | ((uint32_t)chunk << 8);
		}
		__syncwarp();
		if (lane == 0) pend[GEMM_NPEND * 33] = (uint32_t)(base + nh);
		__syncwarp();
	} else {
		gemm_process_chunk(p, r, thr, q, row0, eb); // slots full (cold threshold): emit now
	}
}

// after the accumulator's release: every parked chunk is emitted by the lane that owns it
__device__ __forceinline__ void gemm_drain_pending(const GemmParams &p, float thr, int q, long long tile_row0,
                                                   const EmitBuf &eb, uint32_t *pend) {
	const int np = (int)pend[GEMM_NPEND * 33];
	if (np == 0) return;
	const int lane = threadIdx.x & 31;
	for (int s = 0; s < np; s++) {
		const uint32_t *v = pend + s * 33;
		const uint32_t meta = v[32];
		if ((int)(meta & 0xFFu) == lane) {
			const long long row0 = tile_row0 + 32 * (int)(meta >> 8);
			uint32_t hits = 0u;
#pragma unroll
			for (int c = 0; c < 32; c++) hits |= (__uint_as_float(v[c]) >= thr ? 1u : 0u) << c;
			hits &= gemm_chunk_valid(p, row0);
			if (hits) gemm_emit_chunk(p, v, hits, q, row0, eb.slots, eb.count);
		}
	}
	__syncwarp();
	if (lane == 0) pend[GEMM_NPEND * 33] = 0u;
	__syncwarp();
}

// Variant for the K-streamed kernel (flat_gemm_tsk_kernel), whose epilogue has registers to spare and whose
// workloads (large k) emit two orders of magnitude more candidates: the slow path is taken by the whole warp
// together and stays in registers (measured on C2, B=4096: 8.9 ms per batch against 13.2 ms with the call).
__device__ __forceinline__ void gemm_process_chunk_inline(const GemmParams &p, const uint32_t (&r)[32], float thr, int q,
                                                          long long row0, const EmitBuf &eb) {
	float a[10];
#pragma unroll
	for (int i = 0; i < 10; i++)
		a[i] = tc::fmax3(__uint_as_float(r[3 * i]), __uint_as_float(r[3 * i + 1]), __uint_as_float(r[3 * i + 2]));
	float m1 = tc::fmax3(a[0], a[1], a[2]), m2 = tc::fmax3(a[3], a[4], a[5]), m3 = tc::fmax3(a[6], a[7], a[8]);
	float m4 = tc::fmax3(a[9], __uint_as_float(r[30]), __uint_as_float(r[31]));
	float m = fmaxf(tc::fmax3(m1, m2, m3), m4);
	if (!__any_sync(0xFFFFFFFFu, m >= thr)) return;
	uint32_t hits = 0u;
#pragma unroll
	for (int c = 0; c < 32; c++) hits |= (__uint_as_float(r[c]) >= thr ? 1u : 0u) << c;
	if (hits) hits &= gemm_chunk_valid(p, row0);
	const int h = __popc(hits);
	const uint32_t rbase = (uint32_t)row0;
	int n = eb.slots ? *eb.count : 0;
	// Flushing is WARP-WIDE: a lane whose buffer cannot take its hits makes every lane of the warp flush in the same
	// instruction (32 atomics to 32 lists, one L2 round trip), instead of each lane stalling the warp on its own
	// atomic some other time.  At k = 100 a lane overflows its 8 slots ~150 times per large segment; flushing lane by
	// lane that was ~4 dependent atomic round trips per warp and tile — as long as the tile's MMAs (round-2 timers).
	const bool overflow = !eb.slots || n + h > p.EB;
	if (!__any_sync(0xFFFFFFFFu, overflow)) {
#pragma unroll
		for (int c = 0; c < 32; c++)
			if ((hits >> c) & 1u) {
				eb.slots[n * 128] = make_uint2(r[c], rbase + c);
				n++;
			}
		*eb.count = n;
		return;
	}
	if (n + h == 0) return;
	uint32_t pos = atomicAdd(p.cnt + q, (uint32_t)(n + h));
	uint2 *dst = p.cand + (size_t)q * p.CAP;
	for (int i = 0; i < n; i++, pos++)
		if (pos < (uint32_t)p.CAP) dst[pos] = eb.slots[i * 128];
	if (eb.slots) *eb.count = 0;
#pragma unroll
	for (int c = 0; c < 32; c++)
		if ((hits >> c) & 1u) {
			if (pos < (uint32_t)p.CAP) dst[pos] = make_uint2(r[c], rbase + c);
			pos++;
		}
}

__global__ void __launch_bounds__(GEMM_THREADS, 1)
flat_gemm_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmX, const GemmParams p) {
	extern __shared__ uint8_t smem_raw[];
	const int warp = threadIdx.x >> 5;
	const int lane = threadIdx.x & 31;
	const int g = blockIdx.x % p.NG;
	const int r = blockIdx.x / p.NG;
	const int nst = p.nkb / p.KS;
	const uint32_t stage_bytes = (uint32_t)p.KS * GEMM_KB_BYTES;

	const uint32_t sbase = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
	uint8_t *sgen = smem_raw + (sbase - tc::smem_u32(smem_raw));
	const uint32_t Qs = sbase;
	const uint32_t Xs = Qs + (uint32_t)p.QT * p.nkb * GEMM_KB_BYTES;
	const uint32_t ctl_off = (uint32_t)p.QT * p.nkb * GEMM_KB_BYTES + (uint32_t)p.S * stage_bytes;
	const uint32_t bars = sbase + ctl_off;
	// barrier slots (8 B each): full[S] | empty[S] | qfull | tfull[4] | tempty[4]
	const uint32_t bar_full = bars, bar_empty = bars + 8u * p.S, bar_q = bars + 16u * p.S;
	const uint32_t bar_tfull = bar_q + 8, bar_tempty = bar_tfull + 8 * GEMM_NBUF;
	uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sgen + ctl_off + 16u * p.S + 8 + 16 * GEMM_NBUF);
	float *thr_s = reinterpret_cast<float *>(sgen + ctl_off + 16u * p.S + 8 + 16 * GEMM_NBUF + 16);

	if (warp == 0 && lane == 0) {
		tc::tma_prefetch_desc(&tmQ);
		tc::tma_prefetch_desc(&tmX);
		for (int s = 0; s < p.S; s++) {
			tc::mbar_init(bar_full + 8u * s, 1);
			tc::mbar_init(bar_empty + 8u * s, 1);
		}
		tc::mbar_init(bar_q, 1);
		for (int b = 0; b < GEMM_NBUF; b++) {
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
	const bool has_work = p.tile_lo + r < p.tile_hi;

	if (warp == 0) {
		if (has_work) {
			// ---------------- TMA producer (whole warp converged, one elected lane issues)
			if (tc::elect_one()) {
				tc::mbar_expect_tx(bar_q, (uint32_t)p.QT * p.nkb * GEMM_KB_BYTES);
				for (int t = 0; t < p.QT; t++)
					for (int kb = 0; kb < p.nkb; kb++)
						tc::tma_load_2d(Qs + (uint32_t)(t * p.nkb + kb) * GEMM_KB_BYTES, &tmQ, kb * 64, (g * p.QT + t) * 128,
						                bar_q);
			}
			__syncwarp();
			uint32_t it = 0;
			for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R) {
				for (int c = 0; c < nst; c++, it++) {
					const uint32_t s = it % p.S, ph = (it / p.S) & 1u;
					tc::mbar_wait(bar_empty + 8u * s, ph ^ 1u);
					if (tc::elect_one()) {
						tc::mbar_expect_tx(bar_full + 8u * s, stage_bytes);
						for (int kk = 0; kk < p.KS; kk++)
							tc::tma_load_2d(Xs + s * stage_bytes + (uint32_t)kk * GEMM_KB_BYTES, &tmX, (c * p.KS + kk) * 64,
							                (int)(tile * GEMM_BN), bar_full + 8u * s);
					}
					__syncwarp();
				}
			}
		}
	} else if (warp == 1) {
		if (has_work) {
			// ---------------- MMA issuer: the warp stays converged, one elected lane issues the instructions
			const uint32_t idesc = tc::umma_idesc_bf16(128, GEMM_BN);
			tc::mbar_wait(bar_q, 0);
			tc::tc_fence_after();
			uint32_t it = 0, acc = 0;
			for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R) {
				for (int t = 0; t < p.QT; t++, acc++) {
					const uint32_t b = acc % GEMM_NBUF, n = acc / GEMM_NBUF;
					tc::mbar_wait(bar_tempty + 8u * b, (n & 1u) ^ 1u);
					tc::tc_fence_after();
					const uint32_t d = tmem_base + b * GEMM_BN;
					for (int c = 0; c < nst; c++) {
						const uint32_t ic = it + c, s = ic % p.S, ph = (ic / p.S) & 1u;
						if (t == 0) {
							tc::mbar_wait(bar_full + 8u * s, ph);
							tc::tc_fence_after();
						}
						if (tc::elect_one()) {
							for (int kk = 0; kk < p.KS; kk++) {
								const int kb = c * p.KS + kk;
								const int nk = (kb == p.nkb - 1) ? p.ksteps_last : 4;
								// descriptors built once per k-block; a 16-element K step adds 32 B = 2 to the address field
								const uint64_t a0 = tc::umma_smem_desc(Qs + (uint32_t)(t * p.nkb + kb) * GEMM_KB_BYTES);
								const uint64_t b0 = tc::umma_smem_desc(Xs + s * stage_bytes + (uint32_t)kk * GEMM_KB_BYTES);
#pragma unroll
								for (int k4 = 0; k4 < 4; k4++) {
									if (k4 >= nk) break;
									tc::umma_bf16(d, a0 + (uint64_t)(2 * k4), b0 + (uint64_t)(2 * k4), idesc,
									              (uint32_t)((kb | k4) != 0));
								}
							}
							if (t == p.QT - 1) tc::umma_commit(bar_empty + 8u * s);
							if (c == nst - 1) tc::umma_commit(bar_tfull + 8u * b);
						}
						__syncwarp();
					}
				}
				it += nst;
			}
		}
	} else {
		// ---------------- epilogue: TMEM -> registers -> threshold filter -> candidate lists
		const int e = (warp - 2) >> 2;
		const int quad = warp & 3;
		const int lrow = quad * 32 + lane;
		uint32_t acc = 0;
		for (long long tile = p.tile_lo + r; tile < p.tile_hi; tile += p.R) {
			for (int t = 0; t < p.QT; t++, acc++) {
				if ((int)(acc & 1u) != e) continue;
				const uint32_t b = acc % GEMM_NBUF, n = acc / GEMM_NBUF;
				tc::mbar_wait(bar_tfull + 8u * b, n & 1u);
				tc::tc_fence_after();
				const float thr = thr_s[t * 128 + lrow];
				const int q = (g * p.QT + t) * 128 + lrow;
				const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + b * GEMM_BN;
				const long long row0 = tile * GEMM_BN;
				const EmitBuf eb{nullptr, nullptr};
				uint32_t va[32], vb[32];
				tc::tmem_ld32(taddr, va);
				tmem_wait_regs(va);
				tc::tmem_ld32(taddr + 32, vb);
				gemm_process_chunk(p, va, thr, q, row0, eb);
				tmem_wait_regs(vb);
				tc::tmem_ld32(taddr + 64, va);
				gemm_process_chunk(p, vb, thr, q, row0 + 32, eb);
				tmem_wait_regs(va);
				tc::tmem_ld32(taddr + 96, vb);
				gemm_process_chunk(p, va, thr, q, row0 + 64, eb);
				tmem_wait_regs(vb);
				// the accumulator is in registers: hand the TMEM buffer back before the last chunk is examined
				tc::tc_fence_before();
				__syncwarp();
				if (lane == 0) tc::mbar_arrive(bar_tempty + 8u * b);
				gemm_process_chunk(p, vb, thr, q, row0 + 96, eb);
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

static inline size_t gemm_smem_bytes(int QT, int nkb, int KS, int S) {
	return 1024 /*alignment slack*/ + (size_t)QT * nkb * GEMM_KB_BYTES + (size_t)S * KS * GEMM_KB_BYTES + 16u * S + 8 +
	       16 * GEMM_NBUF + 16 + (size_t)QT * 128 * 4;
}

} // namespace lb
