// flat_gemm.cu — batched exact flat search on the tensor cores (sm_100a tcgen05 / TMEM / TMA).
//
// Replaces, for query batches, the CPU loop the reference enters at rust_lib/src/lance_manager.rs:411-419
// (SURVEY.md 8a row a11).  The contraction q.x is GEMM-shaped, so it runs as bf16 UMMA over a bf16 shadow of
// the table; exactness is recovered with an error bound and an fp32 re-rank:
//
//   1. shadow rows (once per appended row): bf16 of x (dot), x/|x| (cosine) or [x, -|x|^2/2] (l2), so that
//      every metric ranks by ONE inner product  score = <Qs[q], Xs[row]>  (larger = nearer);
//   2. |score - exact| <= err(q) = ERR_C |q| max|x|  (flat_gemm_aux.cuh).  If A_k is the k-th best approximate
//      score over any set of live rows, every true top-k row has score >= A_k - 2 err(q): that is the
//      candidate threshold.  Rows are processed in geometrically growing segments; after each one the
//      threshold of every query is refreshed from its candidate list and the list is compacted, so the
//      large final segments emit a few dozen candidates per query;
//   3. flat_gemm_kernel (flat_gemm_kernel.cuh) streams a segment: TMA -> smem -> tcgen05.mma -> TMEM ->
//      fused threshold filter.  No distance matrix reaches HBM;
//   4. rerank_kernel computes the exact fp32 distance of the surviving candidates and selects the top k
//      under the library's (distance, label) order — the result is the exact flat search, not an approximation;
//   5. a query whose candidate list overflowed (degenerate data: masses of ties) is re-run by the fp32 scan.
//
// Roofline: tensor pipe for large batches (2*B*n*d flops), HBM for small ones (n*Kpad*2 shadow bytes).
#include "flat_gemm.h"

#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

#include "flat_gemm_aux.cuh"
#include "flat_gemm_kernel.cuh"
#include "flat_gemm_ts_kernel.cuh"
#include "flat_gemm_tsk_kernel.cuh"
#include <cstdlib>

namespace lb {

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode_fn() {
	static encode_tiled_fn fn = nullptr;
	if (fn) return fn;
	void *p = nullptr;
	cudaDriverEntryPointQueryResult qres;
	LB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
	if (!p || qres != cudaDriverEntryPointSuccess) throw Error("cuTensorMapEncodeTiled is not available in this driver");
	fn = reinterpret_cast<encode_tiled_fn>(p);
	return fn;
}

// bf16 row-major [rows][Kpad] -> tiled map with a {64, 128} box and the 128-byte swizzle the UMMA descriptors expect
static void encode_rows_map(CUtensorMap *m, const void *base, long long rows, int Kpad, int box_rows = 128) {
	cuuint64_t dims[2] = {(cuuint64_t)Kpad, (cuuint64_t)rows};
	cuuint64_t strides[1] = {(cuuint64_t)Kpad * 2};
	cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
	cuuint32_t estr[2] = {1, 1};
	CUresult r = get_encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
	                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
	                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
	if (r != CUDA_SUCCESS) throw Error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
}

static constexpr int TC_CAP = 16384;      // candidate slots per query
static constexpr int TC_QCHUNK = 4096;    // queries per pipeline pass (candidate buffer = 512 MB)
static constexpr int TC_SEED_TILES = 16;  // first segment: 2048 rows, everything is a candidate
static constexpr size_t TC_SMEM_MAX = 232448;

struct FlatGemm::Impl {
	int device, num_sms, dim, metric;
	int Kreal16, Kpad, nkb, ksteps_last;
	bool bias = false; // l2 on the K-streamed kernel: -|x|^2/2 is added in the epilogue from bias_buf (no extra K columns)
	DevBuf bias_buf;   // fp32 [cap_rows + 256]
	DevBuf shadow;   // bf16 [cap_rows][Kpad]
	DevBuf max_norm; // one float (as bits)
	long long rows_done = 0, cap_rows = 0;
	CUtensorMap tmX, tmX64, tmX32; // 128-, 64- and 32-row boxes over the shadow
	const void *tmX_base = nullptr;
	long long tmX_rows = -1;
	std::once_flag attr_once;
	// the shadow / max_norm work of sync_rows is enqueued on the first caller's stream: every search waits for this
	// event on ITS stream before it reads the shadow (two concurrent callers use two streams)
	cudaEvent_t shadow_ev = nullptr;
	bool shadow_ev_set = false;
	// deferred overflow check of a search (FlatGemm::search with defer = true), one per stream
	struct Pending {
		bool active = false;
		cudaStream_t s = nullptr;
		const float *X = nullptr, *Q = nullptr;
		long long n = 0, allow_bits = 0, label_base = 0;
		const uint32_t *deleted = nullptr, *allow = nullptr;
		const uint32_t *flag_out = nullptr, *total = nullptr;
		int nq = 0, k = 0;
		int64_t *out_labels = nullptr;
		float *out_dist = nullptr;
		int32_t *out_counts = nullptr;
		uint32_t *h_total = nullptr; // pinned, owned by this slot
	};
	std::vector<Pending> pend; // one slot per stream that ever deferred (a handful)
	std::mutex pend_mu;
	Pending *slot_for(cudaStream_t s, bool create) { // pend_mu held
		for (auto &p : pend)
			if (p.s == s) return &p;
		if (!create) return nullptr;
		pend.emplace_back();
		pend.back().s = s;
		LB_CUDA(cudaMallocHost(&pend.back().h_total, 16));
		return &pend.back();
	}
	~Impl() {
		for (auto &p : pend)
			if (p.h_total) cudaFreeHost(p.h_total);
		if (shadow_ev) cudaEventDestroy(shadow_ev);
	}
};

static int gemm_variant_override();
// l2 whose plain dimension lands on the K-streamed kernel (5..12 k-blocks) carries its -|x|^2/2 term in the epilogue
static bool use_bias_epilogue(int dim, int metric) {
	const int nkb_plain = (((dim + 15) & ~15) + 63) / 64;
	return metric == METRIC_L2 && nkb_plain >= 5 && nkb_plain <= 12 && gemm_variant_override() != 1;
}

static void shape_for(int dim, int metric, int &Kreal16, int &Kpad, int &nkb, int &ksteps_last) {
	int kr = (metric == METRIC_L2 && !use_bias_epilogue(dim, metric)) ? dim + 3 : dim;
	Kreal16 = (kr + 15) & ~15;
	Kpad = (Kreal16 + 63) & ~63;
	nkb = Kpad / 64;
	ksteps_last = (Kreal16 - (nkb - 1) * 64) / 16;
}

FlatGemm::FlatGemm(int device, int num_sms, int dim, int metric) : impl_(new Impl()) {
	impl_->device = device;
	impl_->num_sms = num_sms;
	impl_->dim = dim;
	impl_->metric = metric;
	shape_for(dim, metric, impl_->Kreal16, impl_->Kpad, impl_->nkb, impl_->ksteps_last);
	impl_->bias = use_bias_epilogue(dim, metric);
	impl_->max_norm.alloc(16); // [0] max |x| (float bits), [1] "a row with a non-finite norm exists"
	LB_CUDA(cudaMemset(impl_->max_norm.p, 0, 16));
	LB_CUDA(cudaEventCreateWithFlags(&impl_->shadow_ev, cudaEventDisableTiming));
}
FlatGemm::~FlatGemm() { delete impl_; }

bool FlatGemm::supported(int dim, int metric, int k) {
	int Kreal16, Kpad, nkb, ks;
	shape_for(dim, metric, Kreal16, Kpad, nkb, ks);
	if (dim < 16 || k < 1 || k > 1024) return false;
	// one resident query tile + two single-k-block stages must fit
	return gemm_smem_bytes(1, nkb, 1, 2) <= TC_SMEM_MAX;
}

bool FlatGemm::worthwhile(int nq, long long n, int dim, int k) {
	(void)dim;
	(void)k;
	return nq >= 8 && n >= 65536 && n < (1ll << 31) - 256;
}

void FlatGemm::invalidate() {
	impl_->rows_done = 0;
	impl_->tmX_base = nullptr;
}

void FlatGemm::sync_rows(const float *X, long long n, long long cap_rows, cudaStream_t s, int64_t *launches) {
	Impl &im = *impl_;
	if (im.cap_rows < cap_rows || !im.shadow.p) {
		// the store grew (and called invalidate()): rebuild the shadow at the new capacity
		im.shadow.alloc((size_t)cap_rows * im.Kpad * 2 + 1024);
		if (im.bias) {
			// the epilogue reads whole 128-row tiles: rows beyond the table must hold finite values (zeros)
			im.bias_buf.alloc(((size_t)cap_rows + 256) * sizeof(float));
			LB_CUDA(cudaMemsetAsync(im.bias_buf.p, 0, im.bias_buf.bytes, s));
		}
		im.cap_rows = cap_rows;
		im.rows_done = 0;
		im.tmX_base = nullptr;
		LB_CUDA(cudaMemsetAsync(im.max_norm.p, 0, 16, s));
	}
	if (im.rows_done < n) {
		long long cnt = n - im.rows_done;
		shadow_rows_kernel<<<(unsigned)((cnt + 7) / 8), 256, 0, s>>>(X, im.rows_done, n, im.dim, im.metric, im.Kpad,
		                                                             im.shadow.as<__nv_bfloat16>(),
		                                                             im.max_norm.as<uint32_t>(),
		                                                             im.bias ? im.bias_buf.as<float>() : nullptr);
		LB_CUDA(cudaGetLastError());
		if (launches) (*launches)++;
		im.rows_done = n;
		LB_CUDA(cudaEventRecord(im.shadow_ev, s));
		im.shadow_ev_set = true;
	}
	// called under the store's gemm mutex: whoever comes next (any stream) orders itself behind the shadow work
	if (im.shadow_ev_set) LB_CUDA(cudaStreamWaitEvent(s, im.shadow_ev, 0));
	if (im.tmX_base != im.shadow.p || im.tmX_rows != n) {
		encode_rows_map(&im.tmX, im.shadow.p, n, im.Kpad); // rows beyond n read as zeros (TMA out-of-bounds fill)
		encode_rows_map(&im.tmX64, im.shadow.p, n, im.Kpad, 64);
		encode_rows_map(&im.tmX32, im.shadow.p, n, im.Kpad, 32);
		im.tmX_base = im.shadow.p;
		im.tmX_rows = n;
	}
}

struct GemmPlan {
	int QT, KS, S, NG, R, EB = 0;
	size_t smem;
	bool ts;  // queries resident in tensor memory, whole-K X tiles (flat_gemm_ts_kernel, K <= 256)
	bool tsk; // queries resident in tensor memory, K streamed (flat_gemm_tsk_kernel, K <= 768)
	bool mc = false; // tsk: clusters of two query groups share every X tile through TMA multicast (NG even)
};

// 0 auto, 1 all-shared operands, 2 queries in TMEM (needs Kpad <= 256); LANCE_B200_GEMM=ss|ts
static int gemm_variant_override() {
	const char *e = getenv("LANCE_B200_GEMM");
	if (!e) return 0;
	return e[0] == 's' ? 1 : (e[0] == 't' ? 2 : 0);
}

static GemmPlan plan_gemm(int tilesQ, int nkb, int num_sms) {
	GemmPlan best{};
	const int variant = gemm_variant_override();
	if (nkb <= 4 && variant != 1) {
		// TMEM holds 256 columns of query tiles: K/2 columns each
		int qt_max = std::min(8 / nkb, tilesQ);
		int NG = (tilesQ + qt_max - 1) / qt_max;
		best.ts = true;
		best.NG = NG;
		best.QT = (tilesQ + NG - 1) / NG;
		best.KS = nkb;
		best.EB = 8;
		best.S = 2;
		while (best.S < 4 && gemm_ts_smem_bytes(best.QT, nkb, best.S + 1, best.EB) <= TC_SMEM_MAX) best.S++;
		best.R = std::max(1, num_sms / NG);
		best.smem = gemm_ts_smem_bytes(best.QT, nkb, best.S, best.EB);
		return best;
	}
	if (nkb <= 12 && variant != 1) {
		best.tsk = true;
		best.QT = 1;
		best.NG = tilesQ;
		best.KS = std::min(4, nkb);
		if (const char *e = getenv("LANCE_B200_TSK_KS")) best.KS = std::max(1, std::min(nkb, atoi(e)));
		best.EB = 12; // 8 -> 12 slots per lane: fewer warp-wide flushes (C2: 6.81 -> 6.75 ms), 16 no better
		if (const char *e = getenv("LANCE_B200_TSK_EB")) best.EB = std::max(1, std::min(32, atoi(e)));
		best.S = 2;
		while (best.S < 12 && gemm_tsk_smem_bytes(best.S + 1, best.KS, best.EB) <= TC_SMEM_MAX) best.S++;
		best.R = std::max(1, num_sms / best.NG);
		best.smem = gemm_tsk_smem_bytes(best.S, best.KS, best.EB);
		static const bool mc_on = [] {
			const char *e = getenv("LANCE_B200_TSK_MC");
			return e ? atoi(e) != 0 : true;
		}();
		best.mc = mc_on && best.NG >= 2 && (best.NG % 2) == 0;
		return best;
	}
	if (variant == 2) throw Error("LANCE_B200_GEMM=ts needs K <= 768");
	// largest resident query block that still leaves two whole-tile stages
	int qt_max = 1;
	for (int qt = 2; qt <= 8; qt++)
		if (gemm_smem_bytes(qt, nkb, nkb, 2) <= TC_SMEM_MAX) qt_max = qt;
	if (qt_max > tilesQ) qt_max = tilesQ;
	int NG = (tilesQ + qt_max - 1) / qt_max;
	int QT = (tilesQ + NG - 1) / NG;
	best.QT = QT;
	best.NG = NG;
	if (QT > 1) {
		best.KS = nkb;
		best.S = 2;
		while (best.S < 4 && gemm_smem_bytes(QT, nkb, nkb, best.S + 1) <= TC_SMEM_MAX) best.S++;
	} else {
		// one query tile: stages may be single k-blocks; prefer >= 3 stages
		best.KS = nkb;
		while (best.KS > 1 && (nkb % best.KS != 0 || gemm_smem_bytes(1, nkb, best.KS, 3) > TC_SMEM_MAX)) best.KS--;
		best.S = 2;
		while (best.S < 6 && gemm_smem_bytes(1, nkb, best.KS, best.S + 1) <= TC_SMEM_MAX) best.S++;
		if (gemm_smem_bytes(1, nkb, best.KS, best.S) > TC_SMEM_MAX) throw Error("dimension too large for the tcgen05 flat path");
	}
	best.R = std::max(1, num_sms / NG);
	best.smem = gemm_smem_bytes(best.QT, nkb, best.KS, best.S);
	return best;
}

// the opt-in shared-memory sizes of every kernel of this file, once per object
static void ensure_attrs(FlatGemm::Impl &im) {
	std::call_once(im.attr_once, [] {
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_tsk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_tsk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_tsk_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_tsk_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(flat_gemm_ts_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_MAX));
		LB_CUDA(cudaFuncSetAttribute(refresh_threshold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2048 * 8));
		LB_CUDA(cudaFuncSetAttribute(rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2048 * 8));
		LB_CUDA(cudaFuncSetAttribute(rerank_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 2048 * 8));
	});
}

// query tiles of the next pass: at most TC_QCHUNK queries; when every CTA holds a single query tile (large K), a tile
// count c with c * floor(SMs / c) close to the SM count (32 tiles on 148 SMs would idle 20 of them)
static int pass_query_tiles(int nkb, int num_sms, long long remaining_queries) {
	int rem_tiles = (int)((std::min<long long>(TC_QCHUNK, remaining_queries) + 127) / 128);
	int take = rem_tiles;
	const GemmPlan probe = plan_gemm(std::max(2, rem_tiles & ~1), nkb, num_sms);
	if (plan_gemm(rem_tiles, nkb, num_sms).QT == 1 && rem_tiles > 1) {
		// K-streamed kernel with multicast pairs: only EVEN tile counts pair up (24 + 8 tiles for a batch of 4096:
		// 144 of 148 SMs in both passes, every X tile out of the L2 once per pair)
		const bool want_even = probe.tsk && probe.mc && rem_tiles >= 4;
		int best_used = 0;
		for (int c = want_even ? (rem_tiles & ~1) : rem_tiles; c >= (rem_tiles + 1) / 2; c -= want_even ? 2 : 1) {
			int used = c * std::max(1, num_sms / c);
			if (used > num_sms) used = num_sms;
			if (used * 100 >= num_sms * 95) {
				take = c;
				break;
			}
			if (used > best_used) best_used = used, take = c;
		}
	}
	return take;
}

// one launch of the candidate kernel over tiles [p.tile_lo, p.tile_hi) (p.seed: the seeding variant)
static void launch_candidates(FlatGemm::Impl &im, const GemmPlan &pl, GemmParams &p, const CUtensorMap &tmQ, cudaStream_t s) {
	p.R = (int)std::min<long long>(pl.R, p.tile_hi - p.tile_lo);
	if (pl.tsk) {
		if (pl.mc) {
			cudaLaunchConfig_t cfg = {};
			cfg.gridDim = dim3(pl.NG * p.R), cfg.blockDim = dim3(GEMM_THREADS), cfg.dynamicSmemBytes = pl.smem, cfg.stream = s;
			cudaLaunchAttribute at[1];
			at[0].id = cudaLaunchAttributeClusterDimension;
			at[0].val.clusterDim.x = 2, at[0].val.clusterDim.y = 1, at[0].val.clusterDim.z = 1;
			cfg.attrs = at, cfg.numAttrs = 1;
			if (im.bias) LB_CUDA(cudaLaunchKernelEx(&cfg, flat_gemm_tsk_kernel<true, true>, im.tmX64, p));
			else LB_CUDA(cudaLaunchKernelEx(&cfg, flat_gemm_tsk_kernel<true>, im.tmX64, p));
		} else if (im.bias) {
			flat_gemm_tsk_kernel<false, true><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p);
		} else {
			flat_gemm_tsk_kernel<false><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p);
		}
	} else if (pl.ts) {
		if (p.seed) {
			switch (im.nkb) {
			case 1: flat_gemm_ts_kernel<1, true><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			case 2: flat_gemm_ts_kernel<2, true><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			case 3: flat_gemm_ts_kernel<3, true><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			default: flat_gemm_ts_kernel<4, true><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p);
			}
		} else {
			switch (im.nkb) {
			case 1: flat_gemm_ts_kernel<1, false><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			case 2: flat_gemm_ts_kernel<2, false><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			case 3: flat_gemm_ts_kernel<3, false><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p); break;
			default: flat_gemm_ts_kernel<4, false><<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(im.tmX, p);
			}
		}
	} else {
		flat_gemm_kernel<<<pl.NG * p.R, GEMM_THREADS, pl.smem, s>>>(tmQ, im.tmX, p);
	}
	LB_CUDA(cudaGetLastError());
}

void FlatGemm::search(const float *X, long long n, const uint32_t *deleted, const float *d_queries, int nq_total, int k,
                      const uint32_t *d_allow, long long allow_bits, long long label_base, int64_t *d_out_labels,
                      float *d_out_dist, int32_t *d_out_counts, Workspace &ws, cudaStream_t s, int64_t *launches,
                      int64_t *fallback_queries, KernelTimer *timer, bool defer) {
	Impl &im = *impl_;
	finish(ws, s, launches, fallback_queries); // a previous deferred check on this stream still uses the scratch buffers
	const int d = im.dim, Kpad = im.Kpad;
	ensure_attrs(im);
	const int cap = warpsel_cap(k);
	const size_t sel_smem = (size_t)8 * cap * sizeof(u64);
	const long long ntiles = (n + GEMM_BN - 1) / GEMM_BN;
	// segment growth: candidates per refresh ~ k * growth * (margin inflation); keep that well under TC_CAP
	// total slow-path hits per query ~ (#segments) * growth * k: small k affords few, large segments
	// A single query tile (<= 128 queries) emits little, so its launches + refreshes dominate: grow faster
	// (B200, B=64: C2 0.535 -> 0.49 ms, C5 shape 1.53 -> 1.44 ms per batch); full batches are emission-bound at 4.
	const bool one_tile = nq_total <= 128; // query tile rows (the MMA M dimension)
	// K-streamed kernel (K > 256, C2): with the warp-wide emit flush a hit is cheap there, and x8 beats x4 (round 2, C2
	// B = 4096: 6.81 -> 6.50 ms per batch; x12 6.62, x16 6.95); the TMEM-resident kernel of C4 stays at x4 (x8: +3 %)
	const bool streamed = im.nkb > 4;
	const int growth_hi = one_tile ? (k <= 32 ? 16 : 8) : (streamed ? 8 : 4);
	int growth = std::max(2, std::min(growth_hi, std::min(TC_CAP / (8 * k), (one_tile ? 2048 : (streamed ? 1024 : 512)) / k)));
	long long seed_tiles = TC_SEED_TILES;
	if (const char *e = getenv("LANCE_B200_TC_GROWTH")) growth = std::max(2, atoi(e)); // tuning hooks
	if (const char *e = getenv("LANCE_B200_TC_SEED")) seed_tiles = std::max(1, atoi(e));
	// growth of the second segment (tuning hook): the seed's threshold passes ~k/seed_rows of the rows, so a jump of j
	// costs ~j*k emits per query and saves log4(j) launches + refreshes
	int jump = growth; // measured on B200: larger jumps lose more to emission than they save in launches
	if (const char *e = getenv("LANCE_B200_TC_JUMP")) jump = std::max(2, atoi(e));

	for (int q0 = 0, nq = 0; q0 < nq_total; q0 += nq) {
		nq = std::min(pass_query_tiles(im.nkb, im.num_sms, nq_total - q0) * 128, nq_total - q0);
		const int tilesQ = (nq + 127) / 128;
		GemmPlan pl = plan_gemm(tilesQ, im.nkb, im.num_sms);
		const int nslots = pl.NG * pl.QT * 128;

		ws.work2.ensure((size_t)nslots * Kpad * 2 + 1024);
		ws.work3.ensure((size_t)nslots * 5 * 4 + 64);
		ws.work.ensure((size_t)nslots * TC_CAP * sizeof(uint2));
		__nv_bfloat16 *Qs = ws.work2.as<__nv_bfloat16>();
		float *margin = ws.work3.as<float>();
		float *thr = margin + nslots;
		uint32_t *cnt = reinterpret_cast<uint32_t *>(thr + nslots);
		uint32_t *flag = cnt + nslots;
		uint32_t *flag_out = flag + nslots; // + one total counter behind it
		uint32_t *total = flag_out + nslots;
		uint2 *cand = ws.work.as<uint2>();
		const float *Qchunk = d_queries + (size_t)q0 * d;

		prep_query_tiles_kernel<<<(nslots + 7) / 8, 256, 0, s>>>(Qchunk, nq, nslots, d, im.metric, Kpad,
		                                                         im.max_norm.as<uint32_t>(), Qs, margin, thr, cnt, flag,
		                                                         im.bias ? 1 : 0);
		LB_CUDA(cudaGetLastError());
		(*launches)++;
		CUtensorMap tmQ;
		encode_rows_map(&tmQ, Qs, nslots, Kpad);

		GemmParams p;
		p.nkb = im.nkb, p.ksteps_last = im.ksteps_last, p.QT = pl.QT, p.KS = pl.KS, p.S = pl.S, p.NG = pl.NG;
		p.n_rows = n, p.thr = thr, p.cnt = cnt, p.cand = cand, p.CAP = TC_CAP;
		p.deleted = deleted, p.allow = d_allow, p.allow_bits = allow_bits;
		p.Qs = Qs, p.Kpad = Kpad, p.EB = pl.EB;
		p.bias = im.bias ? im.bias_buf.as<float>() : nullptr;
		if (im.bias && !pl.tsk) throw Error("internal: the bias epilogue needs the K-streamed kernel");
		static const int tsk_dbg = [] {
			const char *e = getenv("LANCE_B200_TSK_DBG");
			return e ? atoi(e) : 0;
		}();
		p.dbg = tsk_dbg;
		auto launch_gemm = [&]() { launch_candidates(im, pl, p, tmQ, s); };
		long long t0 = 0, seg = seed_tiles;
		// Seeding (both tensor-memory kernels): one pass over the first `sm_tiles` tiles that emits nothing and keeps the
		// best score of every tile (K-streamed kernel: of every half tile); the k-th largest of those is the threshold the
		// geometric segments start from.  It replaces the segments below sm_tiles, whose loose thresholds let through
		// hundreds of hits per query (on a 12.5 M-row shard they cost 0.3 ms of a 3.4 ms batch), by a pass that costs its
		// MMAs only.  Full batches on the TMEM-resident kernel seed with 1 024 tiles; the K-streamed kernel (large K,
		// tables of a million rows) with 128, or k rounded up to 64 (measured on C2: B=64 0.490 -> 0.44 ms, B=4096
		// 6.71 -> 6.49 ms with 128 tiles; 256 tiles 0.438 / 6.60 ms).
		long long sm_tiles = pl.tsk ? std::min<long long>(1024, std::max<long long>(128, (k + 63) / 64 * 64)) : (one_tile ? 0 : 1024);
		if (const char *e = getenv("LANCE_B200_TC_SEEDMAX")) sm_tiles = std::max(0, atoi(e));
		const long long sm_entries = sm_tiles * (pl.tsk ? 2 : 1);
		p.seed = 0;
		if ((pl.ts || pl.tsk) && sm_entries >= 2 * (long long)k && sm_entries <= TC_CAP && ntiles >= 2 * growth * sm_tiles) {
			p.tile_lo = 0, p.tile_hi = sm_tiles, p.seed = 1;
			int tix = timer ? timer->begin(KC_FLAT_GEMM, s) : -1;
			launch_gemm();
			if (timer) timer->end(tix, s);
			p.seed = 0;
			// the maxima stand in as a candidate list for one refresh (threshold only), then the list is emptied again
			fill_u32_kernel<<<(nslots + 255) / 256, 256, 0, s>>>(cnt, nslots, (uint32_t)sm_entries);
			refresh_threshold_kernel<<<nq, 256, sel_smem, s>>>(cand, TC_CAP, cnt, thr, margin, flag, k, cap);
			LB_CUDA(cudaGetLastError());
			LB_CUDA(cudaMemsetAsync(cnt, 0, (size_t)nslots * 4, s));
			(*launches) += 4;
			seg = sm_tiles * growth; // the first emitting segment covers the seeded rows again
		}
		while (t0 < ntiles) {
			long long t1 = std::min(ntiles, t0 + seg);
			if (ntiles - t1 < seg) t1 = ntiles; // do not leave a short tail segment
			p.tile_lo = t0, p.tile_hi = t1;
			int tix = timer ? timer->begin(KC_FLAT_GEMM, s) : -1;
			launch_gemm();
			if (timer) timer->end(tix, s);
			refresh_threshold_kernel<<<nq, 256, sel_smem, s>>>(cand, TC_CAP, cnt, thr, margin, flag, k, cap);
			LB_CUDA(cudaGetLastError());
			(*launches) += 2;
			t0 = t1;
			seg *= (t0 == seed_tiles ? jump : growth);
		}
		int tix = timer ? timer->begin(KC_RERANK, s) : -1;
		if (nq >= 2 * im.num_sms) {
			rerank_kernel<<<nq, 256, sel_smem, s>>>(X, d, im.metric, Qchunk, cand, TC_CAP, cnt, flag, k, cap, label_base,
			                                        d_out_labels + (size_t)q0 * k, d_out_dist + (size_t)q0 * k,
			                                        d_out_counts + q0);
		} else {
			const int T = std::min(32, (2 * im.num_sms + nq - 1) / nq);
			rerank_dist_kernel<<<dim3(nq, T), 256, 0, s>>>(X, d, im.metric, Qchunk, cand, TC_CAP, cnt, flag);
			LB_CUDA(cudaGetLastError());
			rerank_select_kernel<<<nq, 256, sel_smem, s>>>(cand, TC_CAP, cnt, flag, k, cap, label_base,
			                                               d_out_labels + (size_t)q0 * k, d_out_dist + (size_t)q0 * k,
			                                               d_out_counts + q0);
			(*launches)++;
		}
		LB_CUDA(cudaGetLastError());
		if (timer) timer->end(tix, s);
		LB_CUDA(cudaMemsetAsync(total, 0, 4, s));
		count_flags_kernel<<<(nq + 255) / 256, 256, 0, s>>>(flag, cnt, TC_CAP, nq, d_out_counts + q0, k,
		                                                    im.max_norm.as<uint32_t>() + 1, flag_out, total);
		LB_CUDA(cudaGetLastError());
		(*launches) += 2;

		// overflowed lists (degenerate data) -> exact fp32 scan for those queries.  One small read-back per pass,
		// which the caller may defer to finish() when the whole batch is a single pass.
		const bool last_pass = q0 + nq >= nq_total;
		if (defer && q0 == 0 && last_pass) {
			std::lock_guard<std::mutex> g(im.pend_mu);
			Impl::Pending &pd = *im.slot_for(s, true);
			*pd.h_total = 0;
			LB_CUDA(cudaMemcpyAsync(pd.h_total, total, 4, cudaMemcpyDeviceToHost, s));
			pd.active = true, pd.X = X, pd.Q = Qchunk, pd.n = n, pd.allow_bits = allow_bits, pd.label_base = label_base;
			pd.deleted = deleted, pd.allow = d_allow, pd.flag_out = flag_out, pd.total = total, pd.nq = nq, pd.k = k;
			pd.out_labels = d_out_labels, pd.out_dist = d_out_dist, pd.out_counts = d_out_counts;
			return;
		}
		uint32_t h_total = 0;
		LB_CUDA(cudaMemcpyAsync(&h_total, total, 4, cudaMemcpyDeviceToHost, s));
		LB_CUDA(cudaStreamSynchronize(s));
		if (h_total > 0)
			rerun_flagged(X, n, deleted, Qchunk, nq, k, d_allow, allow_bits, label_base, d_out_labels + (size_t)q0 * k,
			              d_out_dist + (size_t)q0 * k, d_out_counts + q0, flag_out, ws, s, launches, fallback_queries);
	}
}

bool FlatGemm::nearest_supported(int dim, int metric, long long table_rows) {
	if (metric == METRIC_COSINE || dim < 16 || gemm_variant_override() == 1) return false;
	int Kreal16, Kpad, nkb, ks;
	shape_for(dim, metric, Kreal16, Kpad, nkb, ks);
	if (nkb > 12) return false; // neither the TMEM-resident (<= 4 k-blocks) nor the K-streamed kernel
	const long long entries = (table_rows + GEMM_BN - 1) / GEMM_BN * (nkb > 4 ? 2 : 1);
	return table_rows >= 256 && entries <= 4096;
}

bool FlatGemm::nearest(const float *table, long long table_rows, const float *d_points, long long npts, int32_t *d_out_idx,
                       Workspace &ws, cudaStream_t s, int64_t *launches) {
	Impl &im = *impl_;
	if (!nearest_supported(im.dim, im.metric, table_rows)) return false;
	if (!plan_gemm(1, im.nkb, im.num_sms).ts && !plan_gemm(1, im.nkb, im.num_sms).tsk) return false;
	ensure_attrs(im);
	sync_rows(table, table_rows, table_rows, s, launches);
	const int d = im.dim, Kpad = im.Kpad;
	const long long ntiles = (table_rows + GEMM_BN - 1) / GEMM_BN;
	CUtensorMap tmQ_unused = {};
	for (long long q0 = 0; q0 < npts;) {
		const int nq = (int)std::min<long long>((long long)pass_query_tiles(im.nkb, im.num_sms, npts - q0) * 128, npts - q0);
		const GemmPlan pl = plan_gemm((nq + 127) / 128, im.nkb, im.num_sms);
		if (!pl.ts && !pl.tsk) throw Error("internal: nearest() planned the all-shared kernel");
		const int nslots = pl.NG * pl.QT * 128;
		const int E = (int)(ntiles * (pl.tsk ? 2 : 1));
		ws.work2.ensure((size_t)nslots * Kpad * 2 + 1024);
		ws.work3.ensure((size_t)nslots * 5 * 4 + 64);
		ws.work.ensure((size_t)nslots * E * sizeof(uint2));
		__nv_bfloat16 *Qs = ws.work2.as<__nv_bfloat16>();
		float *margin = ws.work3.as<float>();
		float *thr = margin + nslots;
		uint32_t *cnt = reinterpret_cast<uint32_t *>(thr + nslots);
		uint32_t *flag = cnt + nslots;
		const float *Qchunk = d_points + (size_t)q0 * d;
		prep_query_tiles_kernel<<<(nslots + 7) / 8, 256, 0, s>>>(Qchunk, nq, nslots, d, im.metric, Kpad,
		                                                         im.max_norm.as<uint32_t>(), Qs, margin, thr, cnt, flag,
		                                                         im.bias ? 1 : 0);
		LB_CUDA(cudaGetLastError());
		GemmParams p;
		p.nkb = im.nkb, p.ksteps_last = im.ksteps_last, p.QT = pl.QT, p.KS = pl.KS, p.S = pl.S, p.NG = pl.NG;
		p.n_rows = table_rows, p.thr = thr, p.cnt = cnt, p.cand = ws.work.as<uint2>(), p.CAP = E;
		p.deleted = nullptr, p.allow = nullptr, p.allow_bits = 0;
		p.Qs = Qs, p.Kpad = Kpad, p.EB = pl.EB;
		p.bias = im.bias ? im.bias_buf.as<float>() : nullptr;
		p.dbg = 0, p.seed = 1;
		p.tile_lo = 0, p.tile_hi = ntiles;
		launch_candidates(im, pl, p, tmQ_unused, s);
		nearest_from_maxima_kernel<<<(nq + 7) / 8, 256, 0, s>>>(ws.work.as<uint2>(), E, E, table, d, im.metric, Qchunk, nq,
		                                                         d_out_idx + q0);
		LB_CUDA(cudaGetLastError());
		if (launches) *launches += 3;
		q0 += nq;
	}
	return true;
}

void FlatGemm::rerun_flagged(const float *X, long long n, const uint32_t *deleted, const float *Qchunk, int nq, int k,
                             const uint32_t *d_allow, long long allow_bits, long long label_base, int64_t *d_out_labels,
                             float *d_out_dist, int32_t *d_out_counts, const uint32_t *flag_out, Workspace &ws,
                             cudaStream_t s, int64_t *launches, int64_t *fallback_queries) {
	Impl &im = *impl_;
	const int d = im.dim;
	std::vector<uint32_t> hf(nq);
	LB_CUDA(cudaMemcpyAsync(hf.data(), flag_out, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
	LB_CUDA(cudaStreamSynchronize(s));
	std::vector<int> flagged;
	for (int q = 0; q < nq; q++)
		if (hf[q]) flagged.push_back(q);
	if (flagged.empty()) return;
	// The flagged queries are gathered and scanned TOGETHER (the exact scan takes four queries per pass over the table:
	// a flagged batch costs ceil(m / 4) passes instead of m), all enqueued at once; results are scattered back.
	const int m = (int)flagged.size();
	const size_t qb = (size_t)m * d * sizeof(float), lb = (size_t)m * k * sizeof(int64_t), db = (size_t)m * k * sizeof(float);
	const size_t off_l = (qb + 255) & ~(size_t)255, off_d = (off_l + lb + 255) & ~(size_t)255, off_c = (off_d + db + 255) & ~(size_t)255;
	ws.rerun.ensure(off_c + (size_t)m * sizeof(int32_t));
	char *base = static_cast<char *>(ws.rerun.p);
	float *gq = reinterpret_cast<float *>(base);
	int64_t *gl = reinterpret_cast<int64_t *>(base + off_l);
	float *gd = reinterpret_cast<float *>(base + off_d);
	int32_t *gc = reinterpret_cast<int32_t *>(base + off_c);
	for (int i = 0; i < m; i++)
		LB_CUDA(cudaMemcpyAsync(gq + (size_t)i * d, Qchunk + (size_t)flagged[(size_t)i] * d, (size_t)d * sizeof(float),
		                        cudaMemcpyDeviceToDevice, s));
	FlatScanArgs a;
	a.X = X, a.n = n, a.d = d, a.metric = im.metric;
	a.dQ = gq;
	a.nq = m, a.k = k;
	a.allow = d_allow, a.allow_bits = allow_bits, a.deleted = deleted, a.label_base = label_base;
	a.out_labels = gl, a.out_dist = gd, a.out_counts = gc;
	FlatScanPlan fp = plan_flat_scan(X, im.metric, d, k, m, n, im.num_sms);
	ws.work4.ensure(flat_scan_scratch_bytes(fp, k));
	run_flat_scan(a, fp, ws.work4.p, s, launches);
	for (int i = 0; i < m; i++) {
		const int q = flagged[(size_t)i];
		LB_CUDA(cudaMemcpyAsync(d_out_labels + (size_t)q * k, gl + (size_t)i * k, (size_t)k * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
		LB_CUDA(cudaMemcpyAsync(d_out_dist + (size_t)q * k, gd + (size_t)i * k, (size_t)k * sizeof(float), cudaMemcpyDeviceToDevice, s));
		LB_CUDA(cudaMemcpyAsync(d_out_counts + q, gc + i, sizeof(int32_t), cudaMemcpyDeviceToDevice, s));
	}
	*fallback_queries += m;
}

// Completes the deferred overflow check of stream s: waits for the stream, and if some candidate list overflowed
// re-runs those queries with the exact scan (and waits again).  Returns the number of re-run queries; 0 and no wait
// when nothing is pending on that stream.  Other streams' pending searches are not touched.
int FlatGemm::finish(Workspace &ws, cudaStream_t s, int64_t *launches, int64_t *fallback_queries) {
	Impl &im = *impl_;
	Impl::Pending pd;
	{
		std::lock_guard<std::mutex> g(im.pend_mu);
		Impl::Pending *slot = im.slot_for(s, false);
		if (!slot || !slot->active) return 0;
		pd = *slot;
		slot->active = false;
	}
	LB_CUDA(cudaStreamSynchronize(pd.s));
	if (*pd.h_total == 0) return 0;
	int64_t before = *fallback_queries;
	rerun_flagged(pd.X, pd.n, pd.deleted, pd.Q, pd.nq, pd.k, pd.allow, pd.allow_bits, pd.label_base, pd.out_labels, pd.out_dist,
	              pd.out_counts, pd.flag_out, ws, pd.s, launches, fallback_queries);
	LB_CUDA(cudaStreamSynchronize(pd.s));
	return (int)(*fallback_queries - before);
}

const uint32_t *FlatGemm::pending_total(cudaStream_t s) const {
	std::lock_guard<std::mutex> g(impl_->pend_mu);
	Impl::Pending *slot = impl_->slot_for(s, false);
	return slot && slot->active ? slot->total : nullptr;
}

bool FlatGemm::has_pending(cudaStream_t s) const {
	std::lock_guard<std::mutex> g(impl_->pend_mu);
	Impl::Pending *slot = impl_->slot_for(s, false);
	return slot && slot->active;
}

} // namespace lb
