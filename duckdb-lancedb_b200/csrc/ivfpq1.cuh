// ivfpq1.cuh — one IVF_PQ query in ONE cooperative launch (included by ivfpq.cu, inside namespace lb).
//
// The multi-kernel pipeline of ivfpq_host.cuh (prep -> probe -> select -> table -> scan) is right for batches; for the
// single query the SQL surface issues (src/lance_index.cpp:442-465 -> lance_detached_search) it is five launches of
// latency for ~30 MB of traffic: measured 78 us on B200, 0.06 of the HBM roofline.  Here the same arithmetic runs as
// one persistent grid (one 16-warp CTA per SM) with a single grid-wide barrier:
//   phase A (all CTAs): centroid distances (a12; one warp per centroid row, 128-bit loads) -> (distance, partition)
//                       keys, and the query's lookup table T1 (a13; one table entry per thread, same loop as
//                       lut_kernel) — both depend on the query only, so they run side by side;
//   grid barrier;
//   phase B (all CTAs): while the table streams into shared memory (one bulk copy), every CTA selects the nprobes
//                       nearest partitions from the P keys (thread minima -> register sort/merge networks: the
//                       nprobes-th smallest thread minimum bounds the answer, the few keys under it are sorted),
//                       splits the probed lists' 32-row code blocks evenly over the grid and scans its share (ADC,
//                       lane = row, rotated conflict-free lookups, next block prefetched);
//   tail:               register top-32 per warp -> tree over the CTA's warps -> one slot + one ticket per CTA ->
//                       the last CTA merges the slots, re-ranks the k*rf candidates exactly (a14) and writes the
//                       sorted top-k (and the completion word of a polling host call).
// Same probe formula as probe_rows_kernel, same table as lut_kernel, same scan arithmetic as adc_scan_kernel<32, NB>:
// the candidates and distances are those of the batched path.
#pragma once
// (ivfpq.cu includes <cooperative_groups.h> at global scope)
namespace cg = ::cooperative_groups;

static constexpr int IVF1_WARPS = 16;
static constexpr int IVF1_NT = IVF1_WARPS * 32;
static constexpr int IVF1_MAXKEYS = 16; // centroid keys per thread in the probe select: P <= 8192
static constexpr int IVF1_HDR_WORDS = 4; // u64 header: [1] ticket

struct Ivf1Params {
	// query
	const float *q_dev; // or null: inline
	int d, metric;
	// index
	const float *centroids;
	int P, M, dsub;
	const float *codebook;
	const long long *pad_offsets;
	const int32_t *slot_rows;
	const uint8_t *codes_il;
	const float *row_term;
	// filters
	const uint32_t *allow;
	long long allow_bits;
	const uint32_t *deleted;
	// search
	int nprobes, k, kc, refine;
	// scratch (global)
	u64 *probe_keys;   // [P] phase A output
	float *lut_g;      // [256][M] transposed table
	u64 *probe_sorted; // [32] the selected partitions, ascending (read by the debug entry)
	u64 *slots;        // [grid][32]
	u64 *hdr;          // IVF1_HDR_WORDS, zero between launches
	u64 *cand;         // [kc] pre-refine ADC candidates (debug entry)
	// raw rows for the refine step
	const float *X;
	long long label_base;
	int64_t *out_labels;
	float *out_dist;
	int32_t *out_counts;
	volatile uint32_t *done_flag;
	uint32_t done_value;
	float4 qinline[SCAN1_QINLINE_MAX / 4];
};

struct Ivf1Smem {
	size_t off_lut, off_q, off_qraw, off_ctl, off_xch, off_pend, off_cand, total;
};
__host__ __device__ inline Ivf1Smem ivf1_smem(int d, int M) {
	Ivf1Smem s;
	const int dpad = (d + 3) & ~3;
	s.off_lut = 0;
	s.off_q = (size_t)M * 1024;
	s.off_qraw = s.off_q + (size_t)dpad * 4;                  // the caller's query (q' above is normalised for cosine)
	s.off_ctl = s.off_qraw + (size_t)dpad * 4;
	s.off_ctl = (s.off_ctl + 15) & ~(size_t)15;
	s.off_xch = s.off_ctl + 1024;                             // ctl words, mbarrier, probe prefix, probe list
	s.off_pend = s.off_xch + (size_t)IVF1_WARPS * 32 * 8;     // tree exchange [16][32] keys
	s.off_cand = s.off_pend + (size_t)IVF1_WARPS * 32 * 8;    // pending buffers of WarpSelect32
	s.total = s.off_cand + 512 * 8;                           // probe candidates under the bound
	return s;
}

template <int NB> // M / 32
__global__ void __launch_bounds__(IVF1_NT, 1) ivfpq1_kernel(const __grid_constant__ Ivf1Params p) {
	extern __shared__ __align__(128) unsigned char smem_raw[];
	cg::grid_group grid = cg::this_grid();
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int d = p.d, M = p.M, metric = p.metric;
	const int dpad = (d + 3) & ~3;
	const Ivf1Smem L = ivf1_smem(d, M);
	float *lut = reinterpret_cast<float *>(smem_raw + L.off_lut);
	float *qs = reinterpret_cast<float *>(smem_raw + L.off_q);
	float *qraw = reinterpret_cast<float *>(smem_raw + L.off_qraw);
	int *ctl = reinterpret_cast<int *>(smem_raw + L.off_ctl);                 // [0] last flag, [1] counter
	const uint32_t bar = tc::smem_u32(smem_raw + L.off_ctl + 96);
	int *pfx = reinterpret_cast<int *>(smem_raw + L.off_ctl + 128);           // [33] block prefix of the probed lists
	u64 *plist = reinterpret_cast<u64 *>(smem_raw + L.off_ctl + 512);         // [32] selected probe keys
	u64 *xch = reinterpret_cast<u64 *>(smem_raw + L.off_xch);
	u64 *pend = reinterpret_cast<u64 *>(smem_raw + L.off_pend);
	u64 *pcand = reinterpret_cast<u64 *>(smem_raw + L.off_cand);
	const bool is_dot = metric == METRIC_DOT;

	// ---------------- the query: q' = q (l2, dot) or q / |q| (cosine); qn2 = |q'|^2
	if (p.q_dev) {
		for (int i = threadIdx.x; i < dpad; i += IVF1_NT) qs[i] = qraw[i] = i < d ? __ldg(p.q_dev + i) : 0.f;
	} else {
		const int n4 = dpad >> 2;
		for (int i = warp; i < n4; i += IVF1_WARPS) {
			const float4 v = p.qinline[i];
			if (lane == 0) reinterpret_cast<float4 *>(qs)[i] = v, reinterpret_cast<float4 *>(qraw)[i] = v;
		}
	}
	if (threadIdx.x == 0) {
		tc::mbar_init(bar, 1);
		tc::fence_barrier_init();
	}
	__syncthreads();
	float qn2;
	{
		// same reduction as prep_queries_kernel: lane-strided fma, xor butterfly
		float s = 0.f;
		for (int i = lane; i < d; i += 32) s = fmaf(qs[i], qs[i], s);
		for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
		float r = 1.f;
		if (metric == METRIC_COSINE) r = s > 0.f ? rsqrtf(s) : 1.f;
		__syncthreads();
		if (metric == METRIC_COSINE)
			for (int i = threadIdx.x; i < d; i += IVF1_NT) qs[i] *= r;
		__syncthreads();
		float s2 = 0.f;
		for (int i = lane; i < d; i += 32) s2 = fmaf(qs[i], qs[i], s2);
		for (int o = 16; o; o >>= 1) s2 += __shfl_xor_sync(0xFFFFFFFFu, s2, o);
		qn2 = s2;
	}

	// ---------------- phase A1: centroid keys, one warp per centroid row (two rows in flight per warp)
	const int gw = blockIdx.x * IVF1_WARPS + warp, GW = gridDim.x * IVF1_WARPS;
	const bool vec4 = (d & 3) == 0 && ((reinterpret_cast<uintptr_t>(p.centroids) & 15) == 0);
	for (int c0 = gw; c0 < p.P; c0 += 2 * GW) {
		const int c1 = c0 + GW;
		const bool two = c1 < p.P;
		float a0 = 0.f, a1 = 0.f;
		const float *x0 = p.centroids + (size_t)c0 * d, *x1 = p.centroids + (size_t)(two ? c1 : c0) * d;
		if (vec4) {
			const float4 *q4 = reinterpret_cast<const float4 *>(qs);
			const float4 *r0 = reinterpret_cast<const float4 *>(x0), *r1 = reinterpret_cast<const float4 *>(x1);
			const int n4 = d >> 2;
#pragma unroll 4
			for (int c = lane; c < n4; c += 32) {
				const float4 b0 = __ldg(r0 + c), b1 = __ldg(r1 + c);
				const float4 a = q4[c];
				if (is_dot) {
					a0 = fmaf(a.x, b0.x, a0), a0 = fmaf(a.y, b0.y, a0), a0 = fmaf(a.z, b0.z, a0), a0 = fmaf(a.w, b0.w, a0);
					a1 = fmaf(a.x, b1.x, a1), a1 = fmaf(a.y, b1.y, a1), a1 = fmaf(a.z, b1.z, a1), a1 = fmaf(a.w, b1.w, a1);
				} else {
					float e0 = a.x - b0.x, e1 = a.y - b0.y, e2 = a.z - b0.z, e3 = a.w - b0.w;
					a0 = fmaf(e0, e0, a0), a0 = fmaf(e1, e1, a0), a0 = fmaf(e2, e2, a0), a0 = fmaf(e3, e3, a0);
					e0 = a.x - b1.x, e1 = a.y - b1.y, e2 = a.z - b1.z, e3 = a.w - b1.w;
					a1 = fmaf(e0, e0, a1), a1 = fmaf(e1, e1, a1), a1 = fmaf(e2, e2, a1), a1 = fmaf(e3, e3, a1);
				}
			}
		} else {
			for (int c = lane; c < d; c += 32) {
				const float a = qs[c], b0 = __ldg(x0 + c), b1 = __ldg(x1 + c);
				if (is_dot)
					a0 = fmaf(a, b0, a0), a1 = fmaf(a, b1, a1);
				else {
					const float e0 = a - b0, e1 = a - b1;
					a0 = fmaf(e0, e0, a0), a1 = fmaf(e1, e1, a1);
				}
			}
		}
		for (int o = 16; o; o >>= 1) a0 += __shfl_xor_sync(0xFFFFFFFFu, a0, o), a1 += __shfl_xor_sync(0xFFFFFFFFu, a1, o);
		if (lane == 0) {
			p.probe_keys[c0] = make_key(is_dot ? 1.0f - a0 : a0, (uint32_t)c0);
			if (two) p.probe_keys[c1] = make_key(is_dot ? 1.0f - a1 : a1, (uint32_t)c1);
		}
	}
	// ---------------- phase A2: lookup table T1, one entry per thread (transposed [256][M]: the rotated scan's layout)
	{
		const int dsub = p.dsub;
		const int E = M * 256;
		for (int e = blockIdx.x * IVF1_NT + threadIdx.x; e < E; e += gridDim.x * IVF1_NT) {
			const int m = e >> 8, c = e & 255;
			const float *cb = p.codebook + (size_t)e * dsub;
			const float *qm = qs + m * dsub;
			float acc = 0.f;
			if (is_dot) {
				for (int i = 0; i < dsub; i++) acc = fmaf(-qm[i], __ldg(cb + i), acc);
			} else {
				for (int i = 0; i < dsub; i++) {
					const float e2 = qm[i] - __ldg(cb + i);
					acc = fmaf(e2, e2, acc);
				}
			}
			p.lut_g[(size_t)c * M + m] = acc;
		}
	}
	__threadfence();
	grid.sync();

	// ---------------- phase B0: the table streams into shared memory while the partitions are selected
	if (threadIdx.x == 0) {
		asm volatile("fence.proxy.async;" ::: "memory"); // the table was written with ordinary stores by other SMs
		tc::mbar_expect_tx(bar, (uint32_t)M * 1024u);
		const unsigned char *src = reinterpret_cast<const unsigned char *>(p.lut_g);
		for (uint32_t off = 0; off < (uint32_t)M * 1024u; off += 32768u) {
			const uint32_t nb = min(32768u, (uint32_t)M * 1024u - off);
			tc::bulk_load(tc::smem_u32(lut) + off, src + off, nb, bar);
		}
	}
	// ---------------- phase B1: the nprobes nearest partitions (every CTA selects them: no second grid barrier)
	const int np = p.nprobes;
	{
		u64 key[IVF1_MAXKEYS];
		u64 mn = KEY_INVALID;
#pragma unroll
		for (int u = 0; u < IVF1_MAXKEYS; u++) {
			const int i = u * IVF1_NT + (int)threadIdx.x;
			key[u] = i < p.P ? __ldcg(p.probe_keys + i) : KEY_INVALID;
			mn = key[u] < mn ? key[u] : mn;
		}
		// 32nd (np-th) smallest of the 512 thread minima: an upper bound of the np-th smallest key, and a tight one
		u64 best = warp_sort32(mn);
#pragma unroll
		for (int step = 1; step < IVF1_WARPS; step <<= 1) {
			if ((warp & (2 * step - 1)) == step) xch[(size_t)warp * 32 + lane] = best;
			__syncthreads();
			if ((warp & (2 * step - 1)) == 0) best = warp_merge32(best, xch[(size_t)(warp + step) * 32 + lane]);
			__syncthreads();
		}
		if (warp == 0) {
			const u64 t0 = __shfl_sync(0xFFFFFFFFu, best, np - 1);
			if (lane == 0) {
				xch[0] = t0;
				ctl[1] = 0;
			}
		}
		__syncthreads();
		const u64 T0 = xch[0];
		__syncthreads();
#pragma unroll
		for (int u = 0; u < IVF1_MAXKEYS; u++)
			if (key[u] <= T0 && key[u] != KEY_INVALID) {
				const int pos = atomicAdd(&ctl[1], 1);
				if (pos < 512) pcand[pos] = key[u];
			}
		__syncthreads();
		const int nc = min(ctl[1], 512);
		if (warp == 0) {
			// sort the (few) keys under the bound: 32 at a time, merged in registers
			u64 b = KEY_INVALID;
			for (int i0 = 0; i0 < nc; i0 += 32) {
				const u64 v = i0 + (int)lane < nc ? pcand[i0 + lane] : KEY_INVALID;
				b = warp_merge32(b, warp_sort32(v));
			}
			plist[lane] = (int)lane < np ? b : KEY_INVALID;
			// block prefix of the probed lists (32-row blocks)
			int nb = 0;
			if ((int)lane < np && b != KEY_INVALID) {
				const int part = (int)key_row(b);
				nb = (int)((p.pad_offsets[part + 1] - p.pad_offsets[part]) >> 5);
			}
			int incl = nb;
			for (int o = 1; o < 32; o <<= 1) {
				const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
				if ((int)lane >= o) incl += v;
			}
			pfx[lane + 1] = incl;
			if (lane == 0) pfx[0] = 0;
			if (blockIdx.x == 0) p.probe_sorted[lane] = (int)lane < np ? b : KEY_INVALID;
		}
		__syncthreads();
	}

	// ---------------- phase B2: this CTA's share of the probed lists' blocks
	const int TB = pfx[32];
	const int g_lo = (int)((long long)TB * blockIdx.x / gridDim.x), g_hi = (int)((long long)TB * (blockIdx.x + 1) / gridDim.x);
	const int my_pfx = pfx[lane]; // lane i: first block of list i (lists beyond np start at TB)
	const u64 my_pk = plist[lane];
	WarpSelect32 sel;
	sel.init(pend + (size_t)warp * 32, p.kc);
	constexpr int CH = NB * 2;
	const uint32_t lane4 = lane << 2;
	uint4 wn[CH];
	int row_n = -1;
	float rterm_n = 0.f, cst_n = 0.f;
	auto fetch = [&](int g) {
		// list of block g: the last list whose first block is <= g
		const unsigned le = __ballot_sync(0xFFFFFFFFu, my_pfx <= g && (int)lane < np);
		const int li = 31 - __clz(le);
		const u64 pk = __shfl_sync(0xFFFFFFFFu, my_pk, li);
		const int first = __shfl_sync(0xFFFFFFFFu, my_pfx, li);
		const int part = (int)key_row(pk);
		cst_n = is_dot ? 1.0f : key_dist(pk) - qn2;
		const long long blk = (p.pad_offsets[part] >> 5) + (g - first);
		const long long slot = (blk << 5) + lane;
		row_n = __ldg(p.slot_rows + slot);
		rterm_n = is_dot ? 0.f : __ldg(p.row_term + slot);
		const uint4 *cp = reinterpret_cast<const uint4 *>(p.codes_il) + (size_t)blk * CH * 32 + lane;
#pragma unroll
		for (int u = 0; u < CH; u++) wn[u] = __ldg(cp + (size_t)u * 32);
	};
	int g = g_lo + warp;
	if (g < g_hi) fetch(g);
	tc::mbar_wait(bar, 0);
	while (g < g_hi) {
		uint4 w[CH];
#pragma unroll
		for (int u = 0; u < CH; u++) w[u] = wn[u];
		const int row = row_n;
		const float rterm = rterm_n, cst = cst_n;
		g += IVF1_WARPS;
		if (g < g_hi) fetch(g);
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
		if (is_dot)
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

	// ---------------- tail: CTA result -> slot + ticket; the last CTA merges, refines, writes
	sel.flush();
	u64 best = sel.best;
	__syncthreads();
#pragma unroll
	for (int step = 1; step < IVF1_WARPS; step <<= 1) {
		if ((warp & (2 * step - 1)) == step) xch[(size_t)warp * 32 + lane] = best;
		__syncthreads();
		if ((warp & (2 * step - 1)) == 0) best = warp_merge32(best, xch[(size_t)(warp + step) * 32 + lane]);
		__syncthreads();
	}
	if (warp == 0) p.slots[(size_t)blockIdx.x * 32 + lane] = best;
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) ctl[0] = (atomicAdd(reinterpret_cast<unsigned *>(p.hdr + 1), 1u) == gridDim.x - 1u);
	__syncthreads();
	if (!ctl[0]) return;
	__threadfence();
	const int G = (int)gridDim.x;
	{
		u64 mine[10];
		int nm = 0;
#pragma unroll
		for (int u = 0; u < 10; u++) {
			const int b = warp + u * IVF1_WARPS;
			if (b < G) mine[u] = __ldcg(p.slots + (size_t)b * 32 + lane), nm = u + 1;
			else mine[u] = KEY_INVALID;
		}
		best = mine[0];
#pragma unroll
		for (int u = 1; u < 10; u++)
			if (u < nm) best = warp_merge32(best, mine[u]);
		for (int b = warp + 10 * IVF1_WARPS; b < G; b += IVF1_WARPS) best = warp_merge32(best, __ldcg(p.slots + (size_t)b * 32 + lane));
		__syncthreads();
#pragma unroll
		for (int step = 1; step < IVF1_WARPS; step <<= 1) {
			if ((warp & (2 * step - 1)) == step) xch[(size_t)warp * 32 + lane] = best;
			__syncthreads();
			if ((warp & (2 * step - 1)) == 0) best = warp_merge32(best, xch[(size_t)(warp + step) * 32 + lane]);
			__syncthreads();
		}
	}
	// warp 0 holds the sorted ADC candidates; the first kc of them are the candidate list
	if (warp == 0) {
		const u64 c = (int)lane < p.kc ? best : KEY_INVALID;
		xch[lane] = c;
		if ((int)lane < p.kc && p.cand) p.cand[lane] = c;
	}
	__syncthreads();
	// refine (a14): exact distance of every candidate, one warp each (the caller's query, not the normalised one)
	u64 *fin = xch + 64;
	if (warp == 0) fin[lane] = KEY_INVALID;
	__syncthreads();
	const bool v4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.X) & 15) == 0);
	for (int i = warp; i < p.kc; i += IVF1_WARPS) {
		const u64 key = xch[i];
		u64 res = key;
		if (p.refine && key != KEY_INVALID) {
			const uint32_t row = key_row(key);
			res = make_key(warp_exact_distance<false>(metric, qraw, p.X + (size_t)row * d, d, v4), row);
		}
		if (lane == 0) fin[i] = res;
	}
	__syncthreads();
	if (warp == 0) {
		const u64 v = warp_sort32(fin[lane]);
		const bool ok = v != KEY_INVALID && (int)lane < p.k;
		const int cnt = __popc(__ballot_sync(0xFFFFFFFFu, ok));
		if ((int)lane < p.k) {
			p.out_labels[lane] = ok ? (int64_t)(p.label_base + (long long)key_row(v)) : -1;
			p.out_dist[lane] = ok ? key_dist(v) : __int_as_float(0x7F800000);
		}
		if (lane == 0) {
			p.out_counts[0] = cnt;
			p.hdr[1] = 0ull; // ticket back to zero for the next launch
		}
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
