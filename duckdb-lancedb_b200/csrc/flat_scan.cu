// flat_scan.cu — exact fp32 flat kNN scan for small query batches (1..4 queries per pass).
//
// Replaces the CPU hot loop the reference enters at rust_lib/src/lance_manager.rs:411-419 when the
// table has no ANN index: for every live row d(q,x), keep the k smallest.
//   L2 = sum (q-x)^2 (squared, test/sql/lance_basic.test:40-42), cosine = 1 - q.x/(|q||x|),
//   dot = 1 - q.x.
//
// Roofline: HBM.  One pass over X (n*d*4 bytes) per batch of <= 4 queries; arithmetic intensity is
// <= 2 flop/B, far under the fp32 ridge, so no tensor cores here (the batched tcgen05 path lives in
// flat_gemm.cu).  Layout/mapping:
//   * X is row-major fp32 [n][d]; a warp takes 8 consecutive rows (8*d*4 contiguous bytes) and reads
//     them with 128-bit loads, lane l covering float4 chunks l, l+32, ... of every row -> each load
//     instruction of the warp covers 512 contiguous bytes, 8 independent loads in flight per lane;
//   * queries sit in shared memory (one conflict-free LDS.128 per chunk, reused for the 8 rows);
//   * the 8 per-row partial sums are combined with a transposed butterfly (9 shuffles per 8 rows
//     instead of 40), after which lane l holds the total of row (l>>2)&7;
//   * selection is fused (topk.cuh): nothing but the k best keys per block leaves the SM;
//   * blocks share their k-th best key through one global word per query (atomicMin) so late
//     blocks reject almost everything with a single compare.
#include <algorithm>
#include <cstring>
#include <mutex>

#include "kernels.h"
#include "scan_common.cuh"
#include "flat_scan1.cuh"
#include "topk.cuh"

namespace lb {

static constexpr int SCAN_MAXG = 32;       // groups of the two-level final select
static constexpr int SCAN_HDR_WORDS = 128; // u64 words of control state in front of the lists

template <int METRIC, int QB, int VEC>
__global__ void __launch_bounds__(256, (QB == 1 ? 3 : (QB == 2 ? 2 : 1))) flat_scan_kernel(const float *__restrict__ X, long long row_begin, long long n,
                                                         int d, const float *__restrict__ Q, int nq_valid, int k, int cap,
                                                         const uint32_t *__restrict__ allow, long long allow_bits,
                                                         const uint32_t *__restrict__ deleted,
                                                         u64 *__restrict__ partials, u64 *g_tau, int G, long long label_base,
                                                         int64_t *__restrict__ out_labels, float *__restrict__ out_dist,
                                                         int32_t *__restrict__ out_counts, u64 *__restrict__ out_keys) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const int nwarps = blockDim.x >> 5;
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	const int dpad = (d + 3) & ~3;
	float *qs = reinterpret_cast<float *>(smem_raw);                  // [QB][dpad]
	float *qn = qs + QB * dpad;                                       // [QB] (padded to 4)
	int *cnts = reinterpret_cast<int *>(qn + 4);                      // [nwarps][QB]
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw + (((size_t)(QB * dpad + 4) * 4 + (size_t)nwarps * QB * 4 + 15) & ~(size_t)15));

	for (int i = threadIdx.x; i < QB * dpad; i += blockDim.x) {
		int j = i / dpad, c = i - j * dpad;
		qs[i] = (j < nq_valid && c < d) ? Q[(size_t)j * d + c] : 0.f;
	}
	__syncthreads();
	if (warp == 0) {
		for (int j = 0; j < QB; j++) {
			float s = 0.f;
			for (int c = lane; c < d; c += 32) s = fmaf(qs[j * dpad + c], qs[j * dpad + c], s);
			for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
			if (lane == 0) qn[j] = s;
		}
	}
	__syncthreads();

	// control words (all-ones at launch): list length - 1 per (query, group) | arrival ticket - 1 per group | final ticket - 1
	unsigned *ctl = reinterpret_cast<unsigned *>(g_tau + 8);
	const int grp = blockIdx.x % G;
	const int gblocks = ((int)gridDim.x - grp + G - 1) / G;              // blocks in this group
	const size_t gcap = (size_t)((gridDim.x + G - 1) / G) * k;           // key slots of a group list
	WarpSelect sel[QB];
#pragma unroll
	for (int j = 0; j < QB; j++) sel[j].init(bufs + ((size_t)warp * QB + j) * cap, cap, k);

	const int nvec = d / VEC;
	const long long ngroups = (n + 7) >> 3;
	const long long gstride = (long long)gridDim.x * nwarps;
	int iter = 0;
	// The warp's next 32 row groups (g, g + gstride, ..) are looked at together: lane l fetches the liveness of the
	// l-th of them (prefilter / tombstone bits), a ballot keeps the groups with a live row and only those are
	// visited, in the same order as before.  With a selective prefilter (C5's 1 % case) the dead groups used to cost
	// a dependent bitmap load each; now they cost nothing.
	for (long long gb = (row_begin >> 3) + (long long)blockIdx.x * nwarps + warp; gb < ngroups; gb += 32 * gstride) {
		const long long gl = gb + (long long)lane * gstride;
		const unsigned live_l = gl < ngroups ? live_mask8(gl << 3, row_begin, n, allow, allow_bits, deleted) : 0u;
		unsigned todo = __ballot_sync(0xFFFFFFFFu, live_l != 0u);
	while (todo) {
		const int src = __ffs(todo) - 1;
		todo &= todo - 1;
		const unsigned live = __shfl_sync(0xFFFFFFFFu, live_l, src);
		const long long row0 = (gb + (long long)src * gstride) << 3;
		float acc[QB][8];
		float nx[8];
#pragma unroll
		for (int r = 0; r < 8; r++) {
			nx[r] = 0.f;
#pragma unroll
			for (int j = 0; j < QB; j++) acc[j][r] = 0.f;
		}
		// Rows that are filtered out (or beyond n) are redirected to the group's first live row: the load
		// stays unconditional (so the 8 loads issue back to back) but touches no extra cache line.
		const int first_live = __ffs(live) - 1;
		const float *rp[8];
#pragma unroll
		for (int r = 0; r < 8; r++)
			rp[r] = X + ((size_t)row0 + (((live >> r) & 1u) ? r : first_live)) * d + (size_t)lane * VEC;
		for (int c = lane; c < nvec; c += 32) {
			XVec<VEC> xv[8];
#pragma unroll
			for (int r = 0; r < 8; r++) {
				xv[r].load(rp[r]);
				rp[r] += 32 * VEC;
			}
			XVec<VEC> qv[QB];
#pragma unroll
			for (int j = 0; j < QB; j++) qv[j].load_shared(qs + j * dpad + c * VEC);
#pragma unroll
			for (int r = 0; r < 8; r++) {
#pragma unroll
				for (int j = 0; j < QB; j++) accum<METRIC>(acc[j][r], qv[j], xv[r]);
				if (METRIC == METRIC_COSINE) accum_norm(nx[r], xv[r]);
			}
		}
		float xn = 0.f;
		if (METRIC == METRIC_COSINE) xn = reduce8(nx);
		const int r = (lane >> 2) & 7;
		const bool row_live = (live >> r) & 1u;
		const uint32_t row = (uint32_t)(row0 + r);
#pragma unroll
		for (int j = 0; j < QB; j++) {
			float s = reduce8(acc[j]);
			float dist = finish_distance<METRIC>(s, qn[j], xn);
			bool mine = row_live && ((lane & 3u) == (unsigned)(j & 3)) && (j < nq_valid);
			sel[j].push(mine, make_key(dist, row));
		}
		if ((iter & 15) == 15) {
#pragma unroll
			for (int j = 0; j < QB; j++) {
				u64 gt = *reinterpret_cast<volatile u64 *>(g_tau + j);
				if (sel[j].tau < gt) {
					if (lane == 0) atomicMin(g_tau + j, sel[j].tau);
				} else {
					sel[j].tighten(gt);
				}
			}
		}
		iter++;
	}
	}

#pragma unroll
	for (int j = 0; j < QB; j++) sel[j].pad();
	// block merge: one block-wide sort per query over the warps' buffers (query j's buffers are QB*cap apart);
	// afterwards the block's k best keys of query j are the first k entries of warp 0's buffer j
#pragma unroll
	for (int j = 0; j < QB; j++) {
		const int lc = 31 - __clz(cap);
		block_bitonic_sort(bufs, nwarps * cap, [=](int i) { return (((i >> lc) * QB + j) << lc) + (i & (cap - 1)); });
	}
#pragma unroll
	for (int j = 0; j < QB; j++) {
		if (warp != (j % nwarps)) continue;
		const u64 *mine = bufs + (size_t)j * cap;
		// publish this block's k-th best, then write only what can still make the global top-k: every
		// block's k-th best bounds the global k-th best from above, so keys beyond g_tau are dead already
		if (mine[k - 1] != KEY_INVALID) {
			if (lane == 0) atomicMin(g_tau + j, mine[k - 1]);
			__syncwarp();
		}
		// append only what can still make the global top-k (a sorted prefix) to query j's shared list
		const u64 gt = *reinterpret_cast<volatile u64 *>(g_tau + j);
		int npass = 0;
		for (int i0 = 0; i0 < k; i0 += 32) {
			int i = i0 + lane;
			npass += __popc(__ballot_sync(0xFFFFFFFFu, i < k && mine[i] != KEY_INVALID && mine[i] <= gt));
		}
		unsigned base = 0;
		if (lane == 0 && npass > 0) base = atomicAdd(ctl + j * SCAN_MAXG + grp, (unsigned)npass) + 1u; // counters start at 0xFFFFFFFF
		base = __shfl_sync(0xFFFFFFFFu, base, 0);
		u64 *dst = partials + ((size_t)j * G + grp) * gcap + base;
		for (int i = lane; i < npass; i += 32) dst[i] = mine[i];
	}
	// Two-level final select without a second launch: the last block of each group reduces the group's list to k
	// keys (the groups finish together, so these run in parallel on G SMs); the last group then selects the result.
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) cnts[0] = (atomicAdd(ctl + 4 * SCAN_MAXG + grp, 1u) + 1u == (unsigned)gblocks - 1u);
	__syncthreads();
	bool last = cnts[0] != 0;
	__syncthreads();
	if (!last) return;
	__threadfence();
	u64 *level2 = partials + (size_t)QB * G * gcap;
	for (int j = 0; j < QB && j < nq_valid; j++) {
		const unsigned m = *reinterpret_cast<volatile unsigned *>(ctl + j * SCAN_MAXG + grp) + 1u;
		block_final_select(partials + ((size_t)j * G + grp) * gcap, (long long)m, k, cap, bufs, cnts, 0, 0, nullptr, nullptr,
		                   nullptr, level2 + ((size_t)j * G + grp) * k);
	}
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) cnts[0] = (atomicAdd(ctl + 5 * SCAN_MAXG, 1u) + 1u == (unsigned)G - 1u);
	__syncthreads();
	last = cnts[0] != 0;
	__syncthreads();
	if (!last) return;
	__threadfence();
	for (int j = 0; j < QB && j < nq_valid; j++)
		block_final_select(level2 + (size_t)j * G * k, (long long)G * k, k, cap, bufs, cnts, label_base, (size_t)j, out_labels,
		                   out_dist, out_counts, out_keys);
	// Every other block took its tickets after its last access to the header, so the header can be handed back
	// all-ones: the next pass over a persistent header needs no memset launch.
	for (int i = threadIdx.x; i < SCAN_HDR_WORDS; i += blockDim.x) g_tau[i] = ~0ull;
}

// Final select: keys[q][m] (KEY_INVALID entries ignored) -> sorted (label, distance) lists.
// One block per query; also used for the multi-GPU merge (keys rebuilt from gathered lists).
__global__ void __launch_bounds__(256) topk_merge_kernel(const u64 *__restrict__ keys, long long m, int k, int cap,
                                                          long long label_base, int64_t *__restrict__ out_labels,
                                                          float *__restrict__ out_dist, int32_t *__restrict__ out_counts,
                                                          u64 *__restrict__ out_keys) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	int *cnts = reinterpret_cast<int *>(smem_raw);
	u64 *bufs = reinterpret_cast<u64 *>(smem_raw + 64);
	const int q = blockIdx.x;
	block_final_select<true>(keys + (size_t)q * m, m, k, cap, bufs, cnts, label_base, (size_t)q, out_labels, out_dist,
	                         out_counts, out_keys);
}

__global__ void fill_u64_kernel(u64 *p, long long n, u64 v) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n) p[i] = v;
}

// ------------------------------------------------------------------------------------------------ host

typedef void (*scan_kernel_t)(const float *, long long, long long, int, const float *, int, int, int, const uint32_t *, long long,
                              const uint32_t *, u64 *, u64 *, int, long long, int64_t *, float *, int32_t *, u64 *);

template <int METRIC, int QB>
static scan_kernel_t scan_kernel_vec(bool vec4) {
	return vec4 ? (scan_kernel_t)flat_scan_kernel<METRIC, QB, 4> : (scan_kernel_t)flat_scan_kernel<METRIC, QB, 1>;
}
template <int QB>
static scan_kernel_t scan_kernel_metric(int metric, bool vec4) {
	switch (metric) {
	case METRIC_COSINE: return scan_kernel_vec<METRIC_COSINE, QB>(vec4);
	case METRIC_DOT: return scan_kernel_vec<METRIC_DOT, QB>(vec4);
	default: return scan_kernel_vec<METRIC_L2, QB>(vec4);
	}
}
static scan_kernel_t scan_kernel(int metric, int QB, bool vec4) {
	switch (QB) {
	case 4: return scan_kernel_metric<4>(metric, vec4);
	case 2: return scan_kernel_metric<2>(metric, vec4);
	default: return scan_kernel_metric<1>(metric, vec4);
	}
}

static size_t scan_smem_bytes(int QB, int d, int nwarps, int cap) {
	int dpad = (d + 3) & ~3;
	size_t head = (((size_t)(QB * dpad + 4) * 4 + (size_t)nwarps * QB * 4 + 15) & ~(size_t)15);
	return head + (size_t)nwarps * QB * cap * sizeof(u64);
}

static const size_t SMEM_BUDGET = 200 * 1024;

FlatScanPlan plan_flat_scan(const float *X, int metric, int d, int k, int nq, long long n_scan, int num_sms) {
	const long long n = n_scan;
	FlatScanPlan p;
	p.cap = warpsel_cap(k);
	int QB = nq >= 4 ? 4 : (nq >= 2 ? 2 : 1);
	int nwarps = 8;
	while (true) {
		while (nwarps > 1 && scan_smem_bytes(QB, d, nwarps, p.cap) > SMEM_BUDGET) nwarps >>= 1;
		if (scan_smem_bytes(QB, d, nwarps, p.cap) <= SMEM_BUDGET) break;
		if (QB == 1)
			throw Error("k=" + std::to_string(k) + " with dimension " + std::to_string(d) +
			            " is too large for the device scan");
		QB >>= 1;
		nwarps = 8;
	}
	// small tables (the IVF centroid table, C1-sized shards): narrower blocks, so that every SM gets work
	{
		const long long ngroups_all = (n + 7) / 8;
		while (nwarps > 2 && ngroups_all / nwarps < 2ll * num_sms) nwarps >>= 1;
	}
	p.QB = QB;
	p.nwarps = nwarps;
	p.smem = scan_smem_bytes(QB, d, nwarps, p.cap);
	p.vec4 = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
	p.kernel = (const void *)scan_kernel(metric, QB, p.vec4);
	if (p.smem > 48 * 1024)
		LB_CUDA(cudaFuncSetAttribute(p.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
	int per_sm = 1;
	LB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, p.kernel, nwarps * 32, p.smem));
	if (per_sm < 1) per_sm = 1;
	long long ngroups = (n + 7) / 8;
	long long want = (ngroups + nwarps - 1) / nwarps;
	long long full = (long long)num_sms * per_sm;
	p.grid = (int)(want < full ? (want < 1 ? 1 : want) : full);
	return p;
}

// groups of the two-level final select: about sqrt(grid), so both levels see equally short lists
static int scan_groups(int grid, int k) {
	if ((long long)grid * k <= 4096) return 1; // short lists: one level
	int g = 1;
	while (g < SCAN_MAXG && g * g < grid) g++;
	return g;
}

size_t flat_scan_scratch_bytes(const FlatScanPlan &p, int k) {
	const int G = scan_groups(p.grid, k);
	return ((size_t)p.QB * G * ((p.grid + G - 1) / G) * k + (size_t)p.QB * G * k + SCAN_HDR_WORDS) * sizeof(u64);
}

static void merge_launch_params(long long m, int k, int &nwarps, int &cap, size_t &smem) {
	cap = warpsel_cap(k);
	nwarps = 8;
	while (nwarps > 1 && (long long)(nwarps / 2) * 32 >= m) nwarps >>= 1;
	while (nwarps > 1 && 64 + (size_t)nwarps * cap * 8 > SMEM_BUDGET) nwarps >>= 1;
	smem = 64 + (size_t)nwarps * cap * 8;
	if (smem > SMEM_BUDGET) throw Error("k too large for the device merge");
}

void launch_topk_merge(const u64 *keys, int nq, long long m, int k, long long label_base, int64_t *out_labels,
                       float *out_dist, int32_t *out_counts, u64 *out_keys, cudaStream_t s, int64_t *launches) {
	if (nq <= 0) return;
	int nwarps, cap;
	size_t smem;
	merge_launch_params(m, k, nwarps, cap, smem);
	if (smem > 48 * 1024)
		LB_CUDA(cudaFuncSetAttribute(topk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	topk_merge_kernel<<<nq, nwarps * 32, smem, s>>>(keys, m, k, cap, label_base, out_labels, out_dist, out_counts,
	                                                  out_keys);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

void fill_u64(u64 *p, long long n, u64 v, cudaStream_t s, int64_t *launches) {
	if (n <= 0) return;
	fill_u64_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p, n, v);
	LB_CUDA(cudaGetLastError());
	if (launches) (*launches)++;
}

// ------------------------------------------------------------------ single-query latency path (flat_scan1.cuh)
static constexpr size_t SCAN1_SMEM_MAX = 232448 - 4096; // dynamic part: the kernel has ~3 KB of static shared memory

bool scan1_eligible(const float *X, int d, int k, int nq) {
	if (nq != 1 || d < 4 || (d & 3) != 0 || k < 1) return false;
	if ((reinterpret_cast<uintptr_t>(X) & 15) != 0) return false; // bulk copies need 16-byte aligned rows
	static const bool off = [] { // A/B hook
		const char *e = getenv("LANCE_B200_SCAN1");
		return e && atoi(e) == 0;
	}();
	if (off) return false;
	return scan1_smem(d, k).total <= SCAN1_SMEM_MAX;
}

size_t scan1_scratch_bytes(int k, int num_sms) { return (size_t)num_sms * scan1_slot_keys(k) * sizeof(u64); }

void init_scan1_header(u64 *hdr, cudaStream_t s) {
	static const u64 init[SCAN1_HDR_WORDS] = {~0ull, 0ull, 0ull, 0ull};
	LB_CUDA(cudaMemcpyAsync(hdr, init, sizeof(init), cudaMemcpyHostToDevice, s));
}

template <int METRIC>
static void launch_scan1(const Scan1Params &p, int grid, size_t smem, cudaStream_t s) {
	// the kernel also has a few KB of static shared memory (radix-select histograms): raise the dynamic limit to what
	// this launch needs, once per size
	static std::mutex mu;
	static size_t granted_by_device[64] = {0}; // the attribute is per device (in-process sharded handles)
	{
		std::lock_guard<std::mutex> g(mu);
		int dev = 0;
		LB_CUDA(cudaGetDevice(&dev));
		size_t &granted = granted_by_device[dev & 63];
		if (granted == 0) granted = 48 * 1024;
		if (smem > granted) {
			LB_CUDA(cudaFuncSetAttribute(flat_scan1_kernel<METRIC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
			granted = smem;
		}
	}
	flat_scan1_kernel<METRIC><<<grid, SCAN1_WARPS * 32, smem, s>>>(p);
	LB_CUDA(cudaGetLastError());
}

void run_flat_scan1(const FlatScanArgs &a, void *scratch, int num_sms, cudaStream_t s, int64_t *launches) {
	Scan1Params p;
	p.X = a.X, p.n = a.n, p.row_begin = a.row_begin, p.d = a.d, p.k = a.k, p.cap = warpsel_cap(a.k);
	p.allow = a.allow, p.allow_bits = a.allow_bits, p.deleted = a.deleted;
	p.hdr = a.scan1_hdr, p.slots = reinterpret_cast<u64 *>(scratch), p.label_base = a.label_base;
	p.out_labels = a.out_labels, p.out_dist = a.out_dist, p.out_counts = a.out_counts, p.out_keys = a.out_keys;
	p.done_flag = a.done_flag, p.done_value = a.done_value;
	p.trace = nullptr, p.trace2 = nullptr;
	if (a.hQ && a.d <= SCAN1_QINLINE_MAX) {
		p.q_dev = nullptr;
		memcpy(p.qinline, a.hQ, (size_t)a.d * sizeof(float));
	} else {
		p.q_dev = a.dQ;
	}
	const long long units = ((a.n + 7) >> 3) - (a.row_begin >> 3);
	int grid = (int)std::min<long long>(std::max<long long>((units + SCAN1_WARPS - 1) / SCAN1_WARPS, 1), num_sms);
	p.unit_shift = scan1_unit_shift(a.n - a.row_begin, grid);
	const size_t smem = scan1_smem(a.d, a.k).total;
	int tix = a.timer ? a.timer->begin(KC_FLAT_SCAN, s) : -1;
	switch (a.metric) {
	case METRIC_COSINE: launch_scan1<METRIC_COSINE>(p, grid, smem, s); break;
	case METRIC_DOT: launch_scan1<METRIC_DOT>(p, grid, smem, s); break;
	default: launch_scan1<METRIC_L2>(p, grid, smem, s);
	}
	if (a.timer) a.timer->end(tix, s);
	if (launches) (*launches)++;
}

// Runs the scan for a.nq queries (any count) in passes of plan.QB; results land in the out_* device buffers.
void run_flat_scan(const FlatScanArgs &a, const FlatScanPlan &p, void *scratch, cudaStream_t s, int64_t *launches) {
	// scratch header (all-ones at the start of a pass): g_tau[8] | control words of the final select
	static_assert(SCAN_HDR_BYTES == SCAN_HDR_WORDS * sizeof(u64), "header size");
	const bool persistent = a.hdr && a.hdr_clean;
	u64 *g_tau = persistent ? a.hdr : reinterpret_cast<u64 *>(scratch);
	u64 *partials = reinterpret_cast<u64 *>(scratch) + SCAN_HDR_WORDS;
	int G = scan_groups(p.grid, a.k);
	for (int q0 = 0; q0 < a.nq; q0 += p.QB) {
		int nv = a.nq - q0 < p.QB ? a.nq - q0 : p.QB;
		if (!persistent || !*a.hdr_clean) LB_CUDA(cudaMemsetAsync(g_tau, 0xFF, SCAN_HDR_WORDS * sizeof(u64), s));
		if (persistent) *a.hdr_clean = false; // until the launch below is in the stream
		const float *dQ = a.dQ + (size_t)q0 * a.d;
		const float *X = a.X;
		long long n = a.n, allow_bits = a.allow_bits, row_begin = a.row_begin, label_base = a.label_base;
		int d = a.d, k = a.k, cap = p.cap;
		const uint32_t *allow = a.allow, *deleted = a.deleted;
		int64_t *out_labels = a.out_labels ? a.out_labels + (size_t)q0 * a.k : nullptr;
		float *out_dist = a.out_dist ? a.out_dist + (size_t)q0 * a.k : nullptr;
		int32_t *out_counts = a.out_counts ? a.out_counts + q0 : nullptr;
		u64 *out_keys = a.out_keys ? a.out_keys + (size_t)q0 * a.k : nullptr;
		void *args[] = {&X,     &row_begin, &n,        &d,     &dQ,         &nv,         &k,        &cap,        &allow,
		                &allow_bits, &deleted, &partials, &g_tau, &G, &label_base, &out_labels, &out_dist, &out_counts, &out_keys};
		int tix = a.timer ? a.timer->begin(KC_FLAT_SCAN, s) : -1;
		LB_CUDA(cudaLaunchKernel(p.kernel, dim3(p.grid), dim3(p.nwarps * 32), args, p.smem, s));
		if (a.timer) a.timer->end(tix, s);
		if (launches) (*launches)++;
		if (persistent) *a.hdr_clean = true; // the pass's last block restores the header
	}
}

} // namespace lb
