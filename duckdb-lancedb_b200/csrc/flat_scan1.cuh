// flat_scan1.cuh — the single-query latency path of the exact fp32 flat scan.
//
// What the SQL surface issues: lance_search(...) and the ORDER BY array_distance LIMIT k rewrite search with ONE
// query (src/lance_search.cpp:54-82, src/lance_optimizer.cpp:60-82 -> LanceIndex::Search, src/lance_index.cpp:442-465).
// A 100 K x 128 table is 51 MB = 8 us of HBM time, so the call is bound by everything AROUND the arithmetic.  Measured
// on B200 with the register-fed scan (flat_scan.cu, in-kernel globaltimer stamps): 12 dependent DRAM round trips per
// warp (15 us), a 512-key block sort (5 us), a two-level cross-block select (8 us).  This kernel keeps
// flat_scan_kernel's arithmetic (scan_common.cuh: same accumulation order, bit-identical distances) and removes the
// latency around it:
//   * rows are STAGED through shared memory by the TMA engine (cp.async.bulk on per-warp mbarrier rings, 16 warps x
//     12 KB): a warp keeps 8-12 KB in flight without holding a register for it, so a 51 MB table streams at HBM
//     speed instead of one round trip per 8 rows, and a 1 % prefilter gathers its live rows (one bulk copy each) at
//     the same rate;
//   * rows are taken in 32-row chunks (one word of the tombstone / prefilter bitmaps, fetched one group ahead);
//     live rows are batched 8 at a time ACROSS chunks, so sparse filters still issue full batches;
//   * the query can ride in the kernel parameters (<= 1024 floats): a host call needs no H2D copy;
//   * k <= 32 (every configuration of BASELINE.json but C2): selection lives in REGISTERS — a warp's best 32 keys
//     are one key per lane, kept sorted by shuffle networks (sort 32 pending keys, bitonic-merge them in: ~20
//     dependent shuffles, no shared-memory sort); the block result is a 4-level tree of such merges over the 16
//     warps, and the last block of the grid (one ticket per block) merges the blocks' sorted lists the same way.
//     No bitonic sort over shared memory, no radix pass, no second level of tickets;
//   * k > 32: per-warp shared-memory selection as in flat_scan.cu, block result by radix select, last block
//     gathers all slots into the idle stage memory and radix-selects;
//   * an optional completion word (mapped pinned host memory) is written behind a system-scope fence, so the host
//     polls one word instead of a stream synchronise.
#pragma once
#include "scan_common.cuh"
#include "tc_ptx.cuh"

namespace lb {

static constexpr int SCAN1_QINLINE_MAX = 1024; // floats of an inline query
static constexpr int SCAN1_HDR_WORDS = 4;      // u64: [0] bound (all-ones), [1] ticket (0)
static constexpr int SCAN1_WARPS = 16;
static constexpr int SCAN1_RB = 8;                 // rows per batch
static constexpr int SCAN1_SLAB = 128;             // floats of a row per stage (512 B)
static constexpr int SCAN1_RING_BYTES = 12 * 1024; // stage ring of one warp (k <= 32; 8 KB above, the selection buffers need the room)
static constexpr int SCAN1_MAX_STAGES = 6;
static constexpr int SCAN1_KREG = 32;              // k up to this selects in registers

#ifdef SCAN1_TRACE
__device__ __forceinline__ unsigned long long scan1_now() {
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}
#define SCAN1_T(slot) do { if (p.trace && threadIdx.x == 0) p.trace[(size_t)blockIdx.x * 8 + (slot)] = scan1_now(); } while (0)
#else
#define SCAN1_T(slot) do { } while (0)
#endif

struct Scan1Params {
	const float *X;
	long long n, row_begin;
	int d, k, cap;
	const float *q_dev; // the query in device memory, or null: it is in qinline
	const uint32_t *allow;
	long long allow_bits;
	const uint32_t *deleted;
	u64 *hdr;   // SCAN1_HDR_WORDS, persistent per workspace: the last block hands it back clean
	u64 *slots; // [grid][slot_keys] block results (slot_keys = 32 for k <= 32, else k)
	long long label_base;
	int64_t *out_labels;
	float *out_dist;
	int32_t *out_counts;
	u64 *out_keys;
	int unit_shift;               // 3: the batch generator walks 8-row units | 5: whole 32-row bitmap words (scan1_unit_shift)
	volatile uint32_t *done_flag; // nullable; set to done_value once the outputs are visible system-wide
	uint32_t done_value;
	unsigned long long *trace; // dev harness only (SCAN1_TRACE): [grid][8] globaltimer stamps
	long long *trace2;         // dev harness only: [grid][warps][4] loop cycles, wait cycles, first-stage latency, stages
	float4 qinline[SCAN1_QINLINE_MAX / 4];
};

// shared-memory layout of one CTA (bytes), everything 16-byte aligned
struct Scan1Smem {
	int stage_bytes, nstages, ring_bytes; // per warp
	size_t off_q, off_ctl, off_bars, off_rows, off_sel, off_ring, total;
};
// 32-row units once every warp of the grid gets at least eight of them; 8-row units keep smaller tables balanced
__host__ __device__ inline int scan1_unit_shift(long long rows, int grid) {
	return rows >= 32ll * 8 * grid * SCAN1_WARPS ? 5 : 3;
}
__host__ __device__ inline int scan1_slot_keys(int k) { return k <= SCAN1_KREG ? SCAN1_KREG : k; }
// per-warp selection buffer entries: k <= 32 keeps 32 pending keys (the best 32 live in registers)
__host__ __device__ inline int scan1_cap(int k) { return k <= SCAN1_KREG ? 32 : warpsel_cap(k); }
__host__ __device__ inline Scan1Smem scan1_smem(int d, int k) {
	Scan1Smem s;
	const int dpad = (d + 3) & ~3;
	const int slab = dpad < SCAN1_SLAB ? dpad : SCAN1_SLAB;
	s.stage_bytes = SCAN1_RB * slab * 4;
	s.ring_bytes = k <= SCAN1_KREG ? SCAN1_RING_BYTES : 8 * 1024;
	s.nstages = s.ring_bytes / s.stage_bytes;
	if (s.nstages > SCAN1_MAX_STAGES) s.nstages = SCAN1_MAX_STAGES;
	s.off_q = 0;
	s.off_ctl = (size_t)dpad * 4;
	s.off_bars = s.off_ctl + 64;
	s.off_rows = s.off_bars + (size_t)SCAN1_WARPS * SCAN1_MAX_STAGES * 8;
	s.off_sel = s.off_rows + (size_t)SCAN1_WARPS * SCAN1_MAX_STAGES * (SCAN1_RB + 1) * 4;
	s.off_sel = (s.off_sel + 15) & ~(size_t)15;
	s.off_ring = s.off_sel + (size_t)SCAN1_WARPS * scan1_cap(k) * 8;
	s.off_ring = (s.off_ring + 127) & ~(size_t)127;
	s.total = s.off_ring + (size_t)SCAN1_WARPS * s.ring_bytes;
	return s;
}

// liveness of the 32 rows of chunk c (rows [32c, 32c+32)): inside [row_begin, n), not deleted, allowed
__device__ __forceinline__ uint32_t scan1_live_word(const Scan1Params &p, long long c) {
	const long long row0 = c << 5;
	uint32_t m = 0xFFFFFFFFu;
	const long long left = p.n - row0;
	if (left <= 0) return 0u;
	if (left < 32) m = (1u << (int)left) - 1u;
	if (row0 < p.row_begin) {
		if (p.row_begin - row0 >= 32) return 0u;
		m &= ~((1u << (int)(p.row_begin - row0)) - 1u);
	}
	if (p.deleted) m &= ~__ldg(p.deleted + c);
	if (p.allow) {
		uint32_t a = 0u;
		if (row0 < p.allow_bits) {
			a = __ldg(p.allow + c);
			const long long al = p.allow_bits - row0;
			if (al < 32) a &= (1u << (int)al) - 1u;
		}
		m &= a;
	}
	return m;
}

// ---- 32-key lists, one key per lane, sorted ascending across the lanes
__device__ __forceinline__ u64 warp_sort32(u64 v) {
	const unsigned lane = lane_id();
#pragma unroll
	for (int size = 2; size <= 32; size <<= 1)
#pragma unroll
		for (int stride = size >> 1; stride > 0; stride >>= 1) {
			const u64 o = __shfl_xor_sync(0xFFFFFFFFu, v, stride);
			const bool up = (lane & size) == 0, lower = (lane & stride) == 0;
			v = (lower == up) ? (v < o ? v : o) : (v < o ? o : v);
		}
	return v;
}
// the 32 smallest of two sorted lists, sorted: min(a[i], b[31-i]) is bitonic and holds them; 5 half-cleaner stages
__device__ __forceinline__ u64 warp_merge32(u64 a, u64 b) {
	const unsigned lane = lane_id();
	const u64 br = __shfl_sync(0xFFFFFFFFu, b, 31 - lane);
	u64 v = a < br ? a : br;
#pragma unroll
	for (int stride = 16; stride > 0; stride >>= 1) {
		const u64 o = __shfl_xor_sync(0xFFFFFFFFu, v, stride);
		v = ((lane & stride) == 0) ? (v < o ? v : o) : (v < o ? o : v);
	}
	return v;
}

// Streaming selection of the 32 best keys of a warp: `best` (one key per lane, sorted) + up to 32 pending keys in shared
// memory; a full pending buffer is sorted and merged in with shuffles.  tau = the k-th best so far (KEY_INVALID
// until k keys were seen): only keys below it are kept.
struct WarpSelect32 {
	u64 *pend; // shared memory, 32 entries, private to the warp
	u64 best;
	u64 tau;
	int cnt, k;
	__device__ __forceinline__ void init(u64 *smem, int k_) {
		pend = smem, k = k_, cnt = 0, best = KEY_INVALID, tau = KEY_INVALID;
	}
	__device__ __forceinline__ void flush() {
		__syncwarp();
		const u64 v = (int)lane_id() < cnt ? pend[lane_id()] : KEY_INVALID;
		__syncwarp();
		best = warp_merge32(best, warp_sort32(v));
		const u64 t = __shfl_sync(0xFFFFFFFFu, best, k - 1);
		if (t < tau) tau = t;
		cnt = 0;
	}
	__device__ __forceinline__ void push(bool valid, u64 key) {
		bool want = valid && key < tau;
		unsigned m = __ballot_sync(0xFFFFFFFFu, want);
		if (m == 0) return;
		int n = __popc(m);
		if (cnt + n > 32) {
			flush();
			want = valid && key < tau;
			m = __ballot_sync(0xFFFFFFFFu, want);
			if (m == 0) return;
			n = __popc(m);
		}
		if (want) pend[cnt + __popc(m & ((1u << lane_id()) - 1u))] = key;
		cnt += n;
	}
	__device__ __forceinline__ void tighten(u64 bound) {
		if (bound < tau) tau = bound;
	}
};

template <int METRIC>
__global__ void __launch_bounds__(SCAN1_WARPS * 32, 1) flat_scan1_kernel(const __grid_constant__ Scan1Params p) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	constexpr int RB = SCAN1_RB;
	constexpr int NT = SCAN1_WARPS * 32;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int d = p.d, k = p.k;
	const int dpad = (d + 3) & ~3;
	const Scan1Smem L = scan1_smem(d, k);
	float *qs = reinterpret_cast<float *>(smem_raw + L.off_q);
	int *ctl = reinterpret_cast<int *>(smem_raw + L.off_ctl);
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw + L.off_sel);
	const int S = L.nstages;
	const uint32_t stage_bytes = (uint32_t)L.stage_bytes;
	const uint32_t bar0 = tc::smem_u32(smem_raw + L.off_bars) + (uint32_t)warp * SCAN1_MAX_STAGES * 8u;
	uint32_t *srows = reinterpret_cast<uint32_t *>(smem_raw + L.off_rows) + (size_t)warp * SCAN1_MAX_STAGES * (RB + 1);
	unsigned char *ring = smem_raw + L.off_ring + (size_t)warp * L.ring_bytes;
	const uint32_t ring_u32 = tc::smem_u32(ring);
	const bool kreg = k <= SCAN1_KREG;

	SCAN1_T(0);
	if (lane == 0) {
		for (int s = 0; s < S; s++) tc::mbar_init(bar0 + 8u * s, 1);
		tc::fence_barrier_init();
	}
	// ---- the query: from the launch parameters (warp-uniform constant loads, each warp its share) or from memory
	if (p.q_dev) {
		for (int i = threadIdx.x; i < dpad; i += NT) qs[i] = i < d ? __ldg(p.q_dev + i) : 0.f;
	} else {
		const int n4 = dpad >> 2;
		for (int i = warp; i < n4; i += SCAN1_WARPS) {
			const float4 v = p.qinline[i]; // index is warp-uniform: one constant-bank transaction
			if (lane == 0) reinterpret_cast<float4 *>(qs)[i] = v;
		}
	}
	__syncthreads();
	float qn = 0.f;
	if (METRIC == METRIC_COSINE) {
		for (int c = lane; c < d; c += 32) qn = fmaf(qs[c], qs[c], qn);
		for (int o = 16; o; o >>= 1) qn += __shfl_xor_sync(0xFFFFFFFFu, qn, o);
	}
	SCAN1_T(1);

	WarpSelect sel;      // k > 32
	WarpSelect32 sel32;  // k <= 32
	if (kreg) sel32.init(bufs + (size_t)warp * 32, k);
	else sel.init(bufs + (size_t)warp * p.cap, p.cap, k);
	u64 *g_tau = p.hdr;
	const long long gstride = (long long)gridDim.x * SCAN1_WARPS;
	const uint32_t row_bytes = (uint32_t)d * 4u;
	const int nslab = (d + SCAN1_SLAB - 1) / SCAN1_SLAB;
	const unsigned char *Xb = reinterpret_cast<const unsigned char *>(p.X);

	// ---- batch generator: walks this warp's 8-row units (ub, ub + gstride, ..; 32 of them per bitmap fetch) and hands
	// out the live rows RB at a time.  Units of 8 rows (one byte of the tombstone / prefilter words) rather than whole
	// 32-row words keep a small table balanced over the 2 368 warps of the grid (100 K rows = 5 or 6 units per warp).
	// Everything is warp-uniform except byte_l / byte_n (lane l = the l-th unit of the group) and myrow (lane i = the
	// i-th row of the batch being built).
	// unit_shift 5 (tables that give every warp at least eight words): a unit is a whole 32-row bitmap word, so a group
	// fetch covers 1 024 rows — under a 1 % prefilter the generator otherwise spends its time on dependent bitmap loads
	// (one 256-row group holds 2.5 live rows: 3 fetches per batch, 2 860 busy cycles per stage against 200 of arithmetic)
	const int ush = p.unit_shift;
	const long long nunits = (p.n + (1ll << ush) - 1) >> ush;
	long long ub = (p.row_begin >> ush) + (long long)blockIdx.x * SCAN1_WARPS + warp;
	uint32_t byte_l = 0u, byte_n = 0u, w = 0u, row0 = 0u;
	unsigned todo = 0u;
	bool group_loaded = false, contig = false;
	auto fetch_group = [&](long long base) -> uint32_t {
		const long long ul = base + (long long)lane * gstride;
		if (base >= nunits || ul >= nunits) return 0u;
		if (ush == 5) return scan1_live_word(p, ul);
		return (scan1_live_word(p, ul >> 2) >> ((int)(ul & 3) * 8)) & 0xFFu;
	};
	byte_n = fetch_group(ub);
	uint32_t myrow = 0u;
	// returns the number of rows of the next batch (0 = this warp's rows are exhausted)
	auto next_batch = [&]() -> int {
		int n = 0;
		contig = false;
		while (n < RB) {
			if (w == 0u) {
				if (todo == 0u) {
					if (group_loaded) ub += 32 * gstride;
					if (ub >= nunits) break;
					byte_l = byte_n;
					group_loaded = true;
					byte_n = fetch_group(ub + 32 * gstride); // one group ahead: its latency hides behind this group's rows
					todo = __ballot_sync(0xFFFFFFFFu, byte_l != 0u);
					continue;
				}
				const int src = __ffs(todo) - 1;
				todo &= todo - 1;
				w = __shfl_sync(0xFFFFFFFFu, byte_l, src);
				row0 = (uint32_t)((ub + (long long)src * gstride) << ush);
			}
			if (n == 0) {
				// 8 consecutive live rows from the first live one on: the common (dense) case, no per-row bookkeeping
				const int p0 = __ffs(w) - 1;
				if (((w >> p0) & 0xFFu) == 0xFFu) {
					myrow = row0 + p0 + lane;
					w &= ~(0xFFu << p0);
					contig = true;
					return RB;
				}
			}
			const int r0 = __ffs(w) - 1;
			w &= w - 1;
			if ((int)lane == n) myrow = row0 + r0;
			n++;
		}
		return n;
	};

	// ---- producer side of the per-warp ring: one (batch, slab) item per stage; the rows of the item's batch sit in
	// srows[stage][0..RB), its row count and slab index in srows[stage][RB]
	int st_issue = 0, st_cons = 0;
	uint32_t ph_cons = 0u;
	int in_flight = 0;
	int cur_n = next_batch(), cur_slab = 0;
	auto issue_next = [&]() {
		if (cur_n == 0) return;
		const uint32_t bar = bar0 + 8u * st_issue;
		const int f0 = cur_slab * SCAN1_SLAB;
		const uint32_t sbytes = (uint32_t)(((d - f0) < SCAN1_SLAB ? (d - f0) : SCAN1_SLAB) * 4);
		const uint32_t dst = ring_u32 + (uint32_t)st_issue * stage_bytes;
		uint32_t *sr = srows + st_issue * (RB + 1);
		if (lane < (unsigned)RB) sr[lane] = myrow;
		if (lane == 0) {
			sr[RB] = (uint32_t)cur_n | ((uint32_t)cur_slab << 8);
			tc::mbar_expect_tx(bar, sbytes * (uint32_t)cur_n);
		}
		__syncwarp();
		if (contig && nslab == 1) {
			if (lane == 0) tc::bulk_load(dst, Xb + (size_t)myrow * row_bytes, sbytes * (uint32_t)cur_n, bar);
		} else if ((int)lane < cur_n) {
			tc::bulk_load(dst + lane * sbytes, Xb + (size_t)myrow * row_bytes + (size_t)f0 * 4, sbytes, bar);
		}
		if (++st_issue == S) st_issue = 0;
		in_flight++;
		if (++cur_slab == nslab) {
			cur_slab = 0;
			cur_n = next_batch();
		}
	};
	for (int s = 0; s < S; s++) issue_next();

	float acc[RB], nx[RB];
	int iter = 0;
#ifdef SCAN1_TRACE
	long long tr_wait = 0, tr_first = 0, tr_items = 0;
	const long long tr_loop0 = clock64();
#endif
	while (in_flight > 0) {
#ifdef SCAN1_TRACE
		const long long tw0 = clock64();
#endif
		tc::mbar_wait(bar0 + 8u * st_cons, ph_cons);
#ifdef SCAN1_TRACE
		tr_wait += clock64() - tw0;
		if (tr_items++ == 0) tr_first = clock64() - tr_loop0;
#endif
		const uint32_t *sr = srows + st_cons * (RB + 1);
		const uint32_t meta = sr[RB];
		const int nrows = (int)(meta & 0xFFu), slab = (int)(meta >> 8);
		const int f0 = slab * SCAN1_SLAB;
		const int sl4 = ((d - f0) < SCAN1_SLAB ? (d - f0) : SCAN1_SLAB) >> 2; // float4 per row in this stage
		if (slab == 0) {
#pragma unroll
			for (int r = 0; r < RB; r++) acc[r] = 0.f, nx[r] = 0.f;
		}
		const float4 *st4 = reinterpret_cast<const float4 *>(ring + (size_t)st_cons * stage_bytes);
		if ((int)lane < sl4) { // one float4 per lane and row: a slab is at most 32 float4 wide
			XVec<4> xv[RB];
#pragma unroll
			for (int r = 0; r < RB; r++) xv[r].v = st4[(r < nrows ? r : 0) * sl4 + lane];
			XVec<4> qv;
			qv.v = reinterpret_cast<const float4 *>(qs + f0)[lane];
#pragma unroll
			for (int r = 0; r < RB; r++) {
				accum<METRIC>(acc[r], qv, xv[r]);
				if (METRIC == METRIC_COSINE) accum_norm(nx[r], xv[r]);
			}
		}
		const bool last_slab = slab == nslab - 1;
		uint32_t rid = 0u;
		if (last_slab) rid = sr[(lane >> 2) & 7];
		__syncwarp(); // every lane is done with the stage: it may be refilled
		if (++st_cons == S) st_cons = 0, ph_cons ^= 1u;
		in_flight--;
		issue_next();
		if (last_slab) {
			const float s = reduce8(acc);
			float xn = 0.f;
			if (METRIC == METRIC_COSINE) xn = reduce8(nx);
			const float dist = finish_distance<METRIC>(s, qn, xn);
			const bool mine = (lane & 3u) == 0u && (int)((lane >> 2) & 7) < nrows;
			const u64 key = make_key(dist, rid);
			if (kreg) sel32.push(mine, key);
			else sel.push(mine, key);
			// blocks publish / adopt the best k-th key seen anywhere: late rows are rejected with one compare
			if ((++iter & 63) == 0) {
				const u64 gt = *reinterpret_cast<volatile u64 *>(g_tau);
				const u64 mt = kreg ? sel32.tau : sel.tau;
				if (mt < gt) {
					if (lane == 0) atomicMin(g_tau, mt);
				} else if (kreg)
					sel32.tighten(gt);
				else
					sel.tighten(gt);
			}
		}
	}
#ifdef SCAN1_TRACE
	if (p.trace2 && lane == 0) {
		long long *t2 = p.trace2 + ((size_t)blockIdx.x * SCAN1_WARPS + warp) * 4;
		t2[0] = clock64() - tr_loop0, t2[1] = tr_wait, t2[2] = tr_first, t2[3] = tr_items;
	}
#endif
	SCAN1_T(2);

	const int slot_keys = scan1_slot_keys(k);
	u64 *slot = p.slots + (size_t)blockIdx.x * slot_keys;
	u64 *xch = reinterpret_cast<u64 *>(smem_raw + L.off_ring); // exchange area: the rings are idle from here on
	if (kreg) {
		// ---- block result: 4-level tree of register merges over the 16 warps' sorted lists
		sel32.flush();
		u64 best = sel32.best;
		__syncthreads();
#pragma unroll
		for (int step = 1; step < SCAN1_WARPS; step <<= 1) {
			if ((warp & (2 * step - 1)) == step) xch[(size_t)warp * 32 + lane] = best;
			__syncthreads();
			if ((warp & (2 * step - 1)) == 0) best = warp_merge32(best, xch[(size_t)(warp + step) * 32 + lane]);
			__syncthreads();
		}
		if (warp == 0) {
			slot[lane] = best;
			const u64 kth = __shfl_sync(0xFFFFFFFFu, best, k - 1);
			if (lane == 0 && kth != KEY_INVALID) atomicMin(g_tau, kth);
		}
	} else {
		// ---- block result: the k best keys of the block (unsorted) by radix select over the warps' buffers
		sel.pad();
		__syncthreads();
		block_radix_select([=](int i) { return bufs[i]; }, SCAN1_WARPS * p.cap, k, xch);
		u64 mx = 0;
		bool full = true;
		for (int i = threadIdx.x; i < k; i += NT) {
			const u64 v = xch[i];
			if (v == KEY_INVALID) full = false;
			else mx = v > mx ? v : mx;
			slot[i] = v;
		}
		for (int o = 16; o; o >>= 1) {
			const u64 v = __shfl_xor_sync(0xFFFFFFFFu, mx, o);
			mx = v > mx ? v : mx;
		}
		u64 *wmax = xch + k + 32; // [SCAN1_WARPS]
		if (lane == 0) wmax[warp] = mx;
		const int any_short = __syncthreads_or(full ? 0 : 1);
		// the block's k-th best (the maximum of its k winners) bounds the global k-th best from above
		if (threadIdx.x == 0 && !any_short) {
			u64 b = 0;
			for (int w2 = 0; w2 < SCAN1_WARPS; w2++) b = wmax[w2] > b ? wmax[w2] : b;
			atomicMin(g_tau, b);
		}
	}
	SCAN1_T(3);
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) ctl[0] = (atomicAdd(reinterpret_cast<unsigned *>(p.hdr + 1), 1u) == gridDim.x - 1u);
	__syncthreads();
	SCAN1_T(4);
	if (!ctl[0]) return;

	// ---- last block of the grid: the result
	__threadfence();
	const int G = (int)gridDim.x;
	if (kreg) {
		// every warp folds its share of the blocks' sorted lists (loaded up front, independent L2 reads), then the same
		// tree over the warps
		u64 mine[10];
		int nm = 0;
#pragma unroll
		for (int u = 0; u < 10; u++) {
			const int b = warp + u * SCAN1_WARPS;
			if (b < G) mine[u] = __ldcg(p.slots + (size_t)b * 32 + lane), nm = u + 1;
			else mine[u] = KEY_INVALID;
		}
		u64 best = mine[0];
#pragma unroll
		for (int u = 1; u < 10; u++)
			if (u < nm) best = warp_merge32(best, mine[u]);
		for (int b = warp + 10 * SCAN1_WARPS; b < G; b += SCAN1_WARPS) best = warp_merge32(best, __ldcg(p.slots + (size_t)b * 32 + lane));
		SCAN1_T(5);
		__syncthreads();
#pragma unroll
		for (int step = 1; step < SCAN1_WARPS; step <<= 1) {
			if ((warp & (2 * step - 1)) == step) xch[(size_t)warp * 32 + lane] = best;
			__syncthreads();
			if ((warp & (2 * step - 1)) == 0) best = warp_merge32(best, xch[(size_t)(warp + step) * 32 + lane]);
			__syncthreads();
		}
		if (warp == 0) {
			const bool ok = best != KEY_INVALID && (int)lane < k;
			const int c = __popc(__ballot_sync(0xFFFFFFFFu, ok));
			if ((int)lane < k) {
				if (p.out_labels) p.out_labels[lane] = ok ? (int64_t)(p.label_base + (long long)key_row(best)) : -1;
				if (p.out_dist) p.out_dist[lane] = ok ? key_dist(best) : __int_as_float(0x7F800000);
				if (p.out_keys) p.out_keys[lane] = ok ? best : KEY_INVALID;
			}
			if (lane == 0 && p.out_counts) p.out_counts[0] = c;
		}
	} else {
		// all slots -> shared memory (filtered by the bound) -> radix select -> sort
		const u64 gt = *reinterpret_cast<volatile u64 *>(g_tau);
		const int total = G * k;
		u64 *pool = xch + 4096; // behind the select's output area
		const int room = (int)(((size_t)SCAN1_WARPS * L.ring_bytes) / 8) - 4096;
		if (threadIdx.x == 0) ctl[1] = 0;
		__syncthreads();
		for (int i0 = 0; i0 < total; i0 += 8 * NT) {
			u64 key[8];
#pragma unroll
			for (int u = 0; u < 8; u++) {
				const int i = i0 + u * NT + (int)threadIdx.x;
				key[u] = i < total ? __ldcg(p.slots + i) : KEY_INVALID;
			}
#pragma unroll
			for (int u = 0; u < 8; u++)
				if (key[u] <= gt && key[u] != KEY_INVALID) {
					const int pos = atomicAdd(&ctl[1], 1);
					if (pos < room) pool[pos] = key[u];
				}
		}
		__syncthreads();
		const int m = ctl[1];
		SCAN1_T(5);
		if (m <= room && k <= 2048) {
			u64 *fin = xch; // [k + 32]
			int cnt_out;
			if (m > k) {
				block_radix_select([=](int i) { return pool[i]; }, m, k, fin);
				cnt_out = k;
			} else {
				for (int i = threadIdx.x; i < k; i += NT) fin[i] = i < m ? pool[i] : KEY_INVALID;
				cnt_out = m;
				__syncthreads();
			}
			int nn = 64;
			while (nn < k) nn <<= 1;
			for (int i = k + threadIdx.x; i < nn; i += NT) fin[i] = KEY_INVALID;
			block_bitonic_sort(fin, nn, IdentityMap());
			for (int i = threadIdx.x; i < k; i += NT) {
				const u64 v = fin[i];
				const bool ok = v != KEY_INVALID;
				if (p.out_labels) p.out_labels[i] = ok ? (int64_t)(p.label_base + (long long)key_row(v)) : -1;
				if (p.out_dist) p.out_dist[i] = ok ? key_dist(v) : __int_as_float(0x7F800000);
				if (p.out_keys) p.out_keys[i] = v;
			}
			if (threadIdx.x == 0 && p.out_counts) p.out_counts[0] = cnt_out;
		} else {
			// more surviving keys than the stage memory holds (huge k x grid): general select straight from the slots
			block_final_select(p.slots, (long long)total, k, p.cap, bufs, ctl, p.label_base, (size_t)0, p.out_labels, p.out_dist,
			                   p.out_counts, p.out_keys);
		}
	}
	SCAN1_T(6);
	// hand the header back clean, then publish completion
	if (threadIdx.x == 0) {
		p.hdr[0] = ~0ull;
		p.hdr[1] = 0ull;
	}
	if (p.done_flag) {
		__threadfence_system();
		__syncthreads();
		if (threadIdx.x == 0) {
			*p.done_flag = p.done_value;
			__threadfence_system();
		}
	}
}

} // namespace lb
