// kernels.h — host-side launch interfaces of the CUDA kernels (internal to liblance_b200).
#pragma once
#include "common.cuh"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace lb {

// Optional CUDA-event bracket around the dominant kernels (bench.py's roofline.achieved): events are
// recorded on the launching stream, read back after a synchronise.
enum KernelClass : int { KC_FLAT_SCAN = 0, KC_FLAT_GEMM = 1, KC_ADC_SCAN = 2, KC_RERANK = 3, KC_COUNT = 4 };
struct KernelTimer {
	bool on = false;
	std::mutex mu;
	struct Rec {
		cudaEvent_t a, b;
		int cls;
	};
	std::vector<Rec> recs;
	std::vector<cudaEvent_t> free_events;
	cudaEvent_t get_event() {
		if (!free_events.empty()) {
			cudaEvent_t e = free_events.back();
			free_events.pop_back();
			return e;
		}
		cudaEvent_t e;
		LB_CUDA(cudaEventCreate(&e));
		return e;
	}
	// returns an index to pass to end(), or -1 when off
	int begin(int cls, cudaStream_t s) {
		if (!on) return -1;
		std::lock_guard<std::mutex> g(mu);
		Rec r{get_event(), get_event(), cls};
		LB_CUDA(cudaEventRecord(r.a, s));
		recs.push_back(r);
		return (int)recs.size() - 1;
	}
	void end(int idx, cudaStream_t s) {
		if (idx < 0) return;
		std::lock_guard<std::mutex> g(mu);
		LB_CUDA(cudaEventRecord(recs[idx].b, s));
	}
	// sums and clears; the caller has synchronised the device
	void read(double ms[KC_COUNT], int64_t n[KC_COUNT]) {
		std::lock_guard<std::mutex> g(mu);
		const bool dump = getenv("LANCE_B200_TIMER_DUMP") != nullptr; // per-launch durations on stderr (tuning aid)
		for (int i = 0; i < KC_COUNT; i++) ms[i] = 0, n[i] = 0;
		for (auto &r : recs) {
			float t = 0.f;
			if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) {
				ms[r.cls] += t;
				n[r.cls]++;
				if (dump) fprintf(stderr, "[lance_b200 timer] class %d  %.3f us\n", r.cls, t * 1e3);
			}
			free_events.push_back(r.a);
			free_events.push_back(r.b);
		}
		recs.clear();
	}
	~KernelTimer() {
		for (auto &r : recs) {
			cudaEventDestroy(r.a);
			cudaEventDestroy(r.b);
		}
		for (auto e : free_events) cudaEventDestroy(e);
	}
};

// ------------------------------------------------------------------ flat_scan.cu
struct FlatScanArgs {
	const float *X = nullptr; // [n][d] fp32, device
	long long n = 0;
	int d = 0;
	int metric = METRIC_L2;
	const float *dQ = nullptr; // [nq][d] fp32, device
	int nq = 0;
	int k = 0;
	const uint32_t *allow = nullptr; // device bitmap or null
	long long allow_bits = 0;
	const uint32_t *deleted = nullptr; // device tombstones or null
	long long label_base = 0;
	int64_t *out_labels = nullptr; // [nq][k] device (nullable)
	float *out_dist = nullptr;     // [nq][k] device (nullable)
	int32_t *out_counts = nullptr; // [nq] device (nullable)
	u64 *out_keys = nullptr;       // [nq][k] device (nullable): raw keys, row-relative
	long long row_begin = 0;       // only rows >= row_begin are scanned (IVF_PQ's unindexed tail)
	KernelTimer *timer = nullptr;
	// Optional persistent scan header (SCAN_HDR_BYTES, device) + its host-side "is all-ones" flag: the last block of a
	// pass leaves the header as the next pass needs it, so a pass is one launch instead of memset + launch.
	u64 *hdr = nullptr;
	bool *hdr_clean = nullptr;
	// Single-query latency path (flat_scan1.cuh), used when nq == 1 and scan1_eligible():
	const float *hQ = nullptr;       // the query in HOST memory: it then rides in the kernel parameters (no H2D copy)
	u64 *scan1_hdr = nullptr;        // persistent SCAN1 header of the workspace (initialised once, handed back clean)
	volatile uint32_t *done_flag = nullptr; // completion word in mapped pinned host memory (nullable)
	uint32_t done_value = 0;
};
static constexpr size_t SCAN_HDR_BYTES = 128 * sizeof(u64);

struct FlatScanPlan {
	int QB = 1;     // queries per pass (1, 2 or 4)
	int nwarps = 8; // warps per block
	int cap = 64;   // per-warp selection buffer entries
	int grid = 1;
	size_t smem = 0;
	bool vec4 = true;
	const void *kernel = nullptr;
};

// single-query latency path: eligibility and scratch (slots of the grid's blocks), launch
bool scan1_eligible(const float *X, int d, int k, int nq);
size_t scan1_scratch_bytes(int k, int num_sms);
void init_scan1_header(u64 *hdr, cudaStream_t s);
void run_flat_scan1(const FlatScanArgs &a, void *scratch, int num_sms, cudaStream_t s, int64_t *launches);

FlatScanPlan plan_flat_scan(const float *X, int metric, int d, int k, int nq, long long n, int num_sms);
size_t flat_scan_scratch_bytes(const FlatScanPlan &p, int k);
void run_flat_scan(const FlatScanArgs &a, const FlatScanPlan &p, void *scratch, cudaStream_t s, int64_t *launches);

// keys[nq][m] -> sorted lists; any out pointer may be null
void launch_topk_merge(const u64 *keys, int nq, long long m, int k, long long label_base, int64_t *out_labels,
                       float *out_dist, int32_t *out_counts, u64 *out_keys, cudaStream_t s, int64_t *launches);
void fill_u64(u64 *p, long long n, u64 v, cudaStream_t s, int64_t *launches);

// ------------------------------------------------------------------ store_kernels.cu
void launch_gather_rows(const float *X, const int64_t *d_rows, long long m, int d, float *dst, cudaStream_t s,
                        int64_t *launches);
void launch_read_pass(const void *p, size_t bytes, unsigned *sink, cudaStream_t s);
void launch_set_bits(uint32_t *bitmap, const int64_t *d_rows, int count, long long nrows, cudaStream_t s,
                     int64_t *launches);
// Final select over G gathered per-shard lists in one launch.  Shard g's labels / dists / counts start at
// labels + g*ls, dists + g*ds, counts + g*cs (element strides): [G][nq][k] arrays (ls = ds = nq*k, cs = nq) or a
// gathered buffer of per-rank packed records.  Keys carry the index g*k+i, which orders ties like the label does
// because shard g holds a lower label range than shard g+1 and every list is already (distance, label)-sorted.
void launch_merge_lists(const int64_t *labels, const float *dists, const int32_t *counts, long long ls, long long ds,
                        long long cs, int G, int nq, int k, int64_t *out_labels, float *out_dists, int32_t *out_counts,
                        cudaStream_t s, int64_t *launches);

} // namespace lb
