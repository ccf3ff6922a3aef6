// scan1_bench.cu — development harness of the single-query scan (flat_scan1.cuh): correctness against the library's
// multi-query scan (C ABI) and CUDA-event timings with the L2 flushed between launches.  Not part of the product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr \
//        -I duckdb-lancedb_b200/csrc scripts/dev/scan1_bench.cu -o gpurun_out/scan1_bench -L duckdb-lancedb_b200 -llance_b200
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define SCAN1_TRACE 1
#include "flat_scan1.cuh"
#include "../../include/lance_b200.h"

using namespace lb;

__global__ void fill_kernel(float *x, size_t n, uint32_t seed) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	uint32_t h = (uint32_t)i * 2654435761u + seed;
	h ^= h >> 16, h *= 0x85ebca6bu, h ^= h >> 13, h *= 0xc2b2ae35u, h ^= h >> 16;
	uint32_t g = h * 1664525u + 1013904223u;
	// sum of two uniforms, centred: good enough for timing and for distinct distances
	x[i] = ((h >> 8) * (1.0f / 16777216.0f) + (g >> 8) * (1.0f / 16777216.0f)) - 1.0f;
}
__global__ void empty_kernel(int *p) {
	if (p && threadIdx.x == 9999) *p = 1;
}

template <int METRIC>
static float time_scan1(Scan1Params &p, int grid, size_t smem, void *flush, int iters, bool do_flush, cudaStream_t s) {
	auto kern = flat_scan1_kernel<METRIC>;
	cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	std::vector<float> ts;
	cudaEvent_t a, b;
	cudaEventCreate(&a), cudaEventCreate(&b);
	for (int it = 0; it < iters + 3; it++) {
		if (do_flush) cudaMemsetAsync(flush, it, 256 << 20, s);
		cudaStreamSynchronize(s);
		cudaEventRecord(a, s);
		kern<<<grid, SCAN1_WARPS * 32, smem, s>>>(p);
		cudaEventRecord(b, s);
		cudaEventSynchronize(b);
		float ms;
		cudaEventElapsedTime(&ms, a, b);
		if (it >= 3) ts.push_back(ms * 1e3f);
	}
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) printf("CUDA error %s\n", cudaGetErrorString(e));
	std::sort(ts.begin(), ts.end());
	return ts[ts.size() / 2];
}

int main(int argc, char **argv) {
	long long n = argc > 1 ? atoll(argv[1]) : 100000;
	int d = argc > 2 ? atoi(argv[2]) : 128;
	int k = argc > 3 ? atoi(argv[3]) : 10;
	double sel = argc > 4 ? atof(argv[4]) : 1.0;
	const char *metric = argc > 5 ? argv[5] : "l2";
	const int met = metric[0] == 'c' ? METRIC_COSINE : (metric[0] == 'd' ? METRIC_DOT : METRIC_L2);
	cudaStream_t s;
	cudaStreamCreate(&s);
	cudaDeviceProp prop;
	cudaGetDeviceProperties(&prop, 0);
	const int sms = prop.multiProcessorCount;

	char err[2048];
	void *h = lance_create_detached("", d, metric, "t", err, 2048);
	if (!h) return printf("create: %s\n", err), 1;
	lance_b200_reserve(h, n, err, 2048);
	float *X;
	cudaMalloc(&X, (size_t)n * d * 4);
	fill_kernel<<<(unsigned)(((size_t)n * d + 255) / 256), 256, 0, s>>>(X, (size_t)n * d, 12345u);
	cudaStreamSynchronize(s);
	// the library's store gets a copy (old path); the new kernel reads X directly
	for (long long r = 0; r < n; r += 1 << 22) {
		long long c = std::min<long long>(1 << 22, n - r);
		if (lance_b200_add_batch_device(h, X + (size_t)r * d, c, d, s, err, 2048) < 0) return printf("add: %s\n", err), 1;
	}
	std::vector<float> q(d);
	for (int i = 0; i < d; i++) q[i] = sinf(0.37f * i) * 0.8f;
	float *dq;
	cudaMalloc(&dq, d * 4);
	cudaMemcpy(dq, q.data(), d * 4, cudaMemcpyHostToDevice);
	uint32_t *allow = nullptr;
	std::vector<uint32_t> hallow;
	if (sel < 1.0) {
		hallow.assign((size_t)((n + 31) / 32), 0u);
		uint64_t st = 88172645463325252ull;
		for (long long i = 0; i < n; i++) {
			st ^= st << 13, st ^= st >> 7, st ^= st << 17;
			if ((st >> 11) * (1.0 / 9007199254740992.0) < sel) hallow[(size_t)(i >> 5)] |= 1u << (i & 31);
		}
		cudaMalloc(&allow, hallow.size() * 4);
		cudaMemcpy(allow, hallow.data(), hallow.size() * 4, cudaMemcpyHostToDevice);
	}
	// reference result through the library (device entry)
	int64_t *oL;
	float *oD;
	int32_t *oC;
	cudaMalloc(&oL, k * 8), cudaMalloc(&oD, k * 4), cudaMalloc(&oC, 4);
	void *flush;
	cudaMalloc(&flush, 256 << 20);
	std::vector<float> told;
	cudaEvent_t a, b;
	cudaEventCreate(&a), cudaEventCreate(&b);
	for (int it = 0; it < 13; it++) {
		cudaMemsetAsync(flush, it, 256 << 20, s);
		cudaStreamSynchronize(s);
		cudaEventRecord(a, s);
		if (lance_b200_search_device(h, dq, 1, k, 20, 1, allow, allow ? (int64_t)hallow.size() * 32 : 0, oL, oD, oC, s, err, 2048) < 0)
			return printf("search: %s\n", err), 1;
		cudaEventRecord(b, s);
		cudaEventSynchronize(b);
		float ms;
		cudaEventElapsedTime(&ms, a, b);
		if (it >= 3) told.push_back(ms * 1e3f);
	}
	std::sort(told.begin(), told.end());
	std::vector<int64_t> refL(k);
	std::vector<float> refD(k);
	int refC;
	cudaMemcpy(refL.data(), oL, k * 8, cudaMemcpyDeviceToHost), cudaMemcpy(refD.data(), oD, k * 4, cudaMemcpyDeviceToHost);
	cudaMemcpy(&refC, oC, 4, cudaMemcpyDeviceToHost);

	// empty launch
	std::vector<float> te;
	for (int it = 0; it < 13; it++) {
		cudaStreamSynchronize(s);
		cudaEventRecord(a, s);
		empty_kernel<<<sms * 2, 256, 0, s>>>(nullptr);
		cudaEventRecord(b, s);
		cudaEventSynchronize(b);
		float ms;
		cudaEventElapsedTime(&ms, a, b);
		if (it >= 3) te.push_back(ms * 1e3f);
	}
	std::sort(te.begin(), te.end());
	printf("n=%lld d=%d k=%d sel=%.3f %s | library scan %.2f us | empty launch %.2f us\n", n, d, k, sel, metric, told[5], te[5]);

	// new kernel
	Scan1Params p{};
	p.X = X, p.n = n, p.row_begin = 0, p.d = d, p.k = k, p.cap = warpsel_cap(k);
	p.allow = allow, p.allow_bits = allow ? (long long)hallow.size() * 32 : 0, p.deleted = nullptr, p.label_base = 0;
	u64 *hdr, *slots;
	cudaMalloc(&hdr, SCAN1_HDR_WORDS * 8);
	u64 hinit[SCAN1_HDR_WORDS] = {~0ull, 0, 0, 0};
	cudaMemcpy(hdr, hinit, sizeof(hinit), cudaMemcpyHostToDevice);
	p.hdr = hdr;
	int64_t *nL;
	float *nD;
	int32_t *nC;
	cudaMalloc(&nL, k * 8), cudaMalloc(&nD, k * 4), cudaMalloc(&nC, 4);
	p.out_labels = nL, p.out_dist = nD, p.out_counts = nC, p.out_keys = nullptr, p.done_flag = nullptr;
	const size_t smem = scan1_smem(d, k).total;
	for (int inl = 0; inl < 2; inl++) {
		if (inl && d > SCAN1_QINLINE_MAX) break;
		p.q_dev = inl ? nullptr : dq;
		if (inl) memcpy(p.qinline, q.data(), d * 4);
		for (int per_sm = 1; per_sm <= 1; per_sm++) {
			const long long units = (n + 7) / 8;
			int grid = (int)std::min<long long>((units + SCAN1_WARPS - 1) / SCAN1_WARPS, (long long)sms * per_sm);
			cudaMalloc(&slots, (size_t)grid * scan1_slot_keys(k) * 8);
			p.slots = slots;
			p.unit_shift = getenv("SCAN1_UNIT_SHIFT") ? atoi(getenv("SCAN1_UNIT_SHIFT")) : scan1_unit_shift(n, grid);
			float t8 = 0, t16 = 0;
			for (int flushed = 1; flushed >= 0; flushed--) {
				switch (met) {
				case METRIC_COSINE: t8 = time_scan1<METRIC_COSINE>(p, grid, smem, flush, 15, flushed, s); break;
				case METRIC_DOT: t8 = time_scan1<METRIC_DOT>(p, grid, smem, flush, 15, flushed, s); break;
				default: t8 = time_scan1<METRIC_L2>(p, grid, smem, flush, 15, flushed, s);
				}
				t16 = t8;
				std::vector<int64_t> L(k);
				std::vector<float> D(k);
				int C;
				cudaMemcpy(L.data(), nL, k * 8, cudaMemcpyDeviceToHost), cudaMemcpy(D.data(), nD, k * 4, cudaMemcpyDeviceToHost);
				cudaMemcpy(&C, nC, 4, cudaMemcpyDeviceToHost);
				bool ok = C == refC;
				for (int i = 0; i < refC && ok; i++) ok = L[i] == refL[i] && D[i] == refD[i];
				const double bytes = (double)n * d * 4 * sel;
				printf("  scan1 %s grid=%4d (%d/SM) %s: RB8 %.2f us (%.0f GB/s)  RB16 %.2f us (%.0f GB/s)  result %s\n",
				       inl ? "inline-q" : "device-q", grid, per_sm, flushed ? "L2 flushed" : "L2 warm   ", t8, bytes / t8 / 1e3, t16,
				       bytes / t16 / 1e3, ok ? "== library" : "MISMATCH");
				if (!ok) {
					for (int i = 0; i < k; i++) printf("    %d: %lld %g | %lld %g\n", i, (long long)L[i], D[i], (long long)refL[i], refD[i]);
				}
			}
			if (!inl) {
				unsigned long long *tr;
				cudaMalloc(&tr, (size_t)grid * 64);
				cudaMemset(tr, 0, (size_t)grid * 64);
				p.trace = tr;
				long long *tr2;
				cudaMalloc(&tr2, (size_t)grid * SCAN1_WARPS * 32);
				cudaMemset(tr2, 0, (size_t)grid * SCAN1_WARPS * 32);
				p.trace2 = tr2;
				cudaMemsetAsync(flush, 7, 256 << 20, s);
				cudaStreamSynchronize(s);
				if (met == METRIC_L2) flat_scan1_kernel<METRIC_L2><<<grid, SCAN1_WARPS * 32, smem, s>>>(p);
				else if (met == METRIC_COSINE) flat_scan1_kernel<METRIC_COSINE><<<grid, SCAN1_WARPS * 32, smem, s>>>(p);
				else flat_scan1_kernel<METRIC_DOT><<<grid, SCAN1_WARPS * 32, smem, s>>>(p);
				cudaStreamSynchronize(s);
				std::vector<unsigned long long> ht((size_t)grid * 8);
				cudaMemcpy(ht.data(), tr, ht.size() * 8, cudaMemcpyDeviceToHost);
				unsigned long long t0 = ~0ull;
				for (int b2 = 0; b2 < grid; b2++) t0 = std::min(t0, ht[(size_t)b2 * 8]);
				const char *names[7] = {"start", "query in smem", "scan done", "block sorted", "ticket taken", "last: gathered", "last: written"};
				for (int sl = 0; sl < 7; sl++) {
					double mn = 1e30, mx = 0, sum = 0;
					int cntb = 0;
					for (int b2 = 0; b2 < grid; b2++) {
						unsigned long long v = ht[(size_t)b2 * 8 + sl];
						if (!v) continue;
						double us = (double)(v - t0) * 1e-3;
						mn = std::min(mn, us), mx = std::max(mx, us), sum += us, cntb++;
					}
					if (cntb) printf("      trace %-16s blocks=%4d  min %7.2f  mean %7.2f  max %7.2f us\n", names[sl], cntb, mn, sum / cntb, mx);
				}
				{
					std::vector<long long> h2((size_t)grid * SCAN1_WARPS * 4);
					cudaMemcpy(h2.data(), tr2, h2.size() * 8, cudaMemcpyDeviceToHost);
					double sl = 0, sw = 0, sf = 0, ss = 0, mxl = 0;
					for (size_t i = 0; i < h2.size(); i += 4) sl += h2[i], sw += h2[i + 1], sf += h2[i + 2], ss += h2[i + 3], mxl = std::max(mxl, (double)h2[i]);
					const double nw = grid * (double)SCAN1_WARPS;
					printf("      per warp: loop %.0f cyc (max %.0f), of which waiting on mbarriers %.0f, first stage after %.0f, stages %.1f -> %.0f cyc/stage busy\n",
					       sl / nw, mxl, sw / nw, sf / nw, ss / nw, (sl - sw) / std::max(ss, 1.0));
				}
				p.trace = nullptr, p.trace2 = nullptr;
				cudaFree(tr), cudaFree(tr2);
			}
			cudaFree(slots);
		}
	}
	lance_free_detached(h);
	return 0;
}
