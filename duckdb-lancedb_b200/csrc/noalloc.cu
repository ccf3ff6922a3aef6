// noalloc.cu — lance_b200_flat_search_noalloc: an exact flat top-k of ONE query over a caller-owned device table that
// allocates nothing (no cudaMalloc / cudaFree / allocating library call), for callers that cannot allocate: code under
// CUDA graph capture, tight serving loops, harnesses that forbid it.  No counterpart in the reference.
// It is a variant of the streaming flat scan (flat_scan.cu): that kernel keeps its
// per-block partial top-k lists in a per-stream Workspace grown with cudaMalloc; here the partial lists live in ONE
// file-scope device array of fixed size, which every GPU that loads the library holds until the process exits.
//
// Limits: k <= NOALLOC_MAX_K (128), rows < 2^32; scratch = NOALLOC_BLOCKS x NOALLOC_WARPS x NOALLOC_MAX_K keys of 8 bytes
// = 2 424 832 bytes per device.  Every call on a device shares the array and it keeps the last call's contents: the
// scan kernel rewrites every slot the select kernel reads (padding with KEY_INVALID), the arrival counter is cleared
// by the call itself, and two calls that overlap on the same GPU (different streams) RACE — serialise them.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>

#include "exact.cuh"
#include "../../include/lance_b200.h"

namespace lb {

static constexpr int NOALLOC_MAX_K = 128;
static constexpr int NOALLOC_BLOCKS = 296; // two resident CTAs per SM on a 148-SM B200
static constexpr int NOALLOC_WARPS = 8;
static constexpr size_t NOALLOC_KEYS = (size_t)NOALLOC_BLOCKS * NOALLOC_WARPS * NOALLOC_MAX_K;

// In HOST code these names are host-side stand-ins, not device addresses (a kernel handed one faults, or silently
// uses the wrong memory where the GPU can reach host memory): the host gets the real addresses once per device from
// cudaGetSymbolAddress (noalloc_setup) and passes THOSE to the kernels and to cudaMemsetAsync.
static __device__ __align__(256) u64 g_noalloc_keys[NOALLOC_KEYS];
static __device__ __align__(256) unsigned g_noalloc_done[64]; // [0]: blocks that finished the scan of the current call

// Scan: a warp takes 32 consecutive rows per round (lane r keeps the distance of row r), exact fp32 distance by
// the whole warp per row (exact.cuh: the arithmetic of the library's re-rank / refine), warp-level selection in
// shared memory (topk.cuh).  Each warp leaves exactly k keys (sorted, padded with KEY_INVALID) in its slot.
__global__ void __launch_bounds__(NOALLOC_WARPS * 32) noalloc_scan_kernel(const float *__restrict__ X, long long n, int d,
                                                                          int metric, const float *__restrict__ q, int k,
                                                                          const uint32_t *__restrict__ allow,
                                                                          long long allow_bits, u64 *__restrict__ scratch,
                                                                          unsigned *__restrict__ done) {
	__shared__ u64 sel_buf[NOALLOC_WARPS][256]; // warpsel_cap(128) = 256
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	WarpSelect sel;
	sel.init(sel_buf[warp], warpsel_cap(k), k);
	const bool vec4 = (d & 3) == 0 && (reinterpret_cast<size_t>(X) & 15) == 0 && (reinterpret_cast<size_t>(q) & 15) == 0;
	const long long gw = (long long)blockIdx.x * NOALLOC_WARPS + warp, nw = (long long)gridDim.x * NOALLOC_WARPS;
	for (long long base = gw * 32; base < n; base += nw * 32) {
		float mine = 0.f;
		const int rows = (int)min(32ll, n - base);
		for (int r = 0; r < rows; r++) {
			const float dist = warp_exact_distance(metric, q, X + (size_t)(base + r) * d, d, vec4);
			if ((int)lane == r) mine = dist;
		}
		const long long row = base + lane;
		bool ok = (int)lane < rows;
		if (ok && allow) ok = row < allow_bits && bit_test(allow, row);
		sel.push(ok, make_key(mine, (uint32_t)row));
	}
	sel.finalize();
	u64 *slot = scratch + ((size_t)blockIdx.x * NOALLOC_WARPS + warp) * k;
	for (int i = lane; i < k; i += 32) slot[i] = i < sel.cnt ? sel.buf[i] : KEY_INVALID;
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) atomicAdd(done, 1u);
}

// Select: one block reduces the grid's slots to the k best and writes labels / distances / count.
__global__ void __launch_bounds__(NOALLOC_WARPS * 32) noalloc_select_kernel(int slots, int k, long long label_base,
                                                                            int64_t *__restrict__ out_labels,
                                                                            float *__restrict__ out_dist,
                                                                            int32_t *__restrict__ out_count,
                                                                            const u64 *__restrict__ scratch,
                                                                            const unsigned *__restrict__ done) {
	__shared__ u64 sel_buf[NOALLOC_WARPS][256];
	__shared__ int cnts[NOALLOC_WARPS];
	const int warp = threadIdx.x >> 5;
	const unsigned lane = lane_id();
	WarpSelect sel;
	sel.init(sel_buf[warp], warpsel_cap(k), k);
	const long long m = (long long)slots * k;
	for (long long i0 = (long long)warp * 32; i0 < m; i0 += NOALLOC_WARPS * 32) {
		const long long i = i0 + lane;
		const u64 key = i < m ? scratch[i] : KEY_INVALID;
		sel.push(key != KEY_INVALID, key);
	}
	sel.finalize();
	if (lane == 0) cnts[warp] = sel.cnt;
	__syncthreads();
	if (warp != 0) return;
	for (int w = 1; w < NOALLOC_WARPS; w++) {
		const int c = cnts[w];
		for (int i0 = 0; i0 < c; i0 += 32) {
			const int i = i0 + lane;
			sel.push(i < c, i < c ? sel_buf[w][i] : KEY_INVALID);
		}
	}
	sel.finalize();
	for (int i = lane; i < k; i += 32) {
		const bool live = i < sel.cnt;
		out_labels[i] = live ? label_base + (long long)key_row(sel.buf[i]) : -1;
		out_dist[i] = live ? key_dist(sel.buf[i]) : __int_as_float(0x7f800000);
	}
	// every block of this call's scan must have arrived: anything else means another call reset or bumped the shared
	// counter while this one was in flight (overlapping calls race) — reported as count -1 instead of a wrong list
	if (lane == 0) *out_count = done[0] * NOALLOC_WARPS == (unsigned)slots ? sel.cnt : -1;
}

struct NoAllocDevice {
	bool ready = false;
	u64 *keys = nullptr;
	unsigned *done = nullptr;
};
static std::mutex g_na_mu;
static NoAllocDevice g_na[64];

// once per device, with that device current: the device addresses of the file-scope arrays
static NoAllocDevice &noalloc_setup(int device) {
	std::lock_guard<std::mutex> g(g_na_mu);
	NoAllocDevice &nd = g_na[device];
	if (!nd.ready) {
		cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&nd.keys), g_noalloc_keys);
		if (e != cudaSuccess) throw Error(std::string("cudaGetSymbolAddress(g_noalloc_keys): ") + cudaGetErrorString(e));
		e = cudaGetSymbolAddress(reinterpret_cast<void **>(&nd.done), g_noalloc_done);
		if (e != cudaSuccess) throw Error(std::string("cudaGetSymbolAddress(g_noalloc_done): ") + cudaGetErrorString(e));
		nd.ready = true;
	}
	return nd;
}

} // namespace lb

extern "C" int32_t lance_b200_flat_search_noalloc(int32_t device, const float *d_vectors, int64_t rows, int32_t dim,
                                                  int32_t metric, const float *d_query, int32_t k, int64_t label_base,
                                                  const uint32_t *d_allow, int64_t bitmap_bits, int64_t *d_out_labels,
                                                  float *d_out_distances, int32_t *d_out_count, void *stream,
                                                  char *err_buf, int err_buf_len) {
	auto fail = [&](const std::string &msg) {
		if (err_buf && err_buf_len > 0) {
			const size_t n = msg.size() < (size_t)(err_buf_len - 1) ? msg.size() : (size_t)(err_buf_len - 1);
			memcpy(err_buf, msg.data(), n);
			err_buf[n] = '\0';
		}
		return -1;
	};
	try {
		using namespace lb;
		if (device < 0 || device >= 64) return fail("flat_search_noalloc failed: device index out of range");
		if (k < 1 || k > NOALLOC_MAX_K)
			return fail("flat_search_noalloc failed: k = " + std::to_string(k) + " exceeds the static scratch (1 <= k <= " +
			            std::to_string(NOALLOC_MAX_K) + ")");
		if (rows < 0 || rows >= (1ll << 32)) return fail("flat_search_noalloc failed: rows must be in [0, 2^32)");
		if (dim < 1) return fail("flat_search_noalloc failed: dimension must be positive");
		if (metric < METRIC_L2 || metric > METRIC_DOT) return fail("flat_search_noalloc failed: metric must be 0 (l2), 1 (cosine) or 2 (dot)");
		if ((!d_vectors && rows > 0) || !d_query || !d_out_labels || !d_out_distances || !d_out_count)
			return fail("flat_search_noalloc failed: null buffer");
		LB_CUDA(cudaSetDevice(device));
		NoAllocDevice &nd = noalloc_setup(device);
		cudaStream_t s = static_cast<cudaStream_t>(stream);
		// what must start at zero is cleared by the call (a memset node, not an allocation); the key slots need no
		// clearing: the scan rewrites every slot the select reads
		LB_CUDA(cudaMemsetAsync(nd.done, 0, sizeof(unsigned), s));
		const long long need = (rows + 32ll * NOALLOC_WARPS - 1) / (32ll * NOALLOC_WARPS);
		const int grid = (int)std::max(1ll, std::min<long long>(NOALLOC_BLOCKS, need));
		noalloc_scan_kernel<<<grid, NOALLOC_WARPS * 32, 0, s>>>(d_vectors, rows, dim, metric, d_query, k, d_allow,
		                                                         d_allow ? bitmap_bits : 0, nd.keys, nd.done);
		LB_CUDA(cudaGetLastError());
		noalloc_select_kernel<<<1, NOALLOC_WARPS * 32, 0, s>>>(grid * NOALLOC_WARPS, k, label_base, d_out_labels,
		                                                        d_out_distances, d_out_count, nd.keys, nd.done);
		LB_CUDA(cudaGetLastError());
		return 0;
	} catch (const std::exception &e) {
		return fail(std::string("flat_search_noalloc failed: ") + e.what());
	} catch (...) {
		return fail("flat_search_noalloc failed: unknown error");
	}
}
