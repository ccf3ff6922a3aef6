// flat_gemm.h — batched flat search on the tensor cores: bf16 tcgen05 candidate GEMM with a fused
// threshold filter, then an exact fp32 re-rank (see flat_gemm.cu for the design).
#pragma once
#include "store.h"

namespace lb {

class FlatGemm {
public:
	FlatGemm(int device, int num_sms, int dim, int metric);
	~FlatGemm();

	// is the tensor path implemented for this shape at all
	static bool supported(int dim, int metric, int k);
	// is it the faster choice (auto mode)
	static bool worthwhile(int nq, long long n, int dim, int k);

	// the bf16 shadow copy of X and the per-row norms follow the store: convert rows [done, n)
	void sync_rows(const float *X, long long n, long long cap_rows, cudaStream_t s, int64_t *launches);
	void invalidate(); // X was reallocated

	void search(const float *X, long long n, const uint32_t *deleted, const float *d_queries, int nq, int k,
	            const uint32_t *d_allow, long long allow_bits, long long label_base, int64_t *d_out_labels,
	            float *d_out_dist, int32_t *d_out_counts, Workspace &ws, cudaStream_t s, int64_t *launches,
	            int64_t *fallback_queries, KernelTimer *timer, bool defer = false);
	// defer = true: when the batch is a single pass, the candidate-list overflow check (one 4-byte read-back and a
	// stream synchronise) is left to finish(); until then the results of overflowed queries (degenerate data only)
	// are not final and the caller's buffers must stay valid.
	// One pending search per STREAM; finish settles the caller's stream only.
	int finish(Workspace &ws, cudaStream_t s, int64_t *launches, int64_t *fallback_queries);
	bool has_pending(cudaStream_t s) const;
	const uint32_t *pending_total(cudaStream_t s) const; // device word: number of overflowed queries of the pending search

	// Nearest row of `table` for every point (the k-means assignment of the IVF_PQ trainer, ivfpq_train.cuh): the
	// seeding variant of the candidate kernels over the WHOLE table — per (half) tile the best bf16 score and its row —
	// then the exact fp32 distance of each point to the rows of its four best tiles, argmin.  The metric of this object
	// is l2 or dot (cosine trains on unit rows under l2).  `table` must be what sync_rows last converted, or the object
	// invalidated.  False when the shape has no tensor-memory kernel (the caller keeps its CUDA-core path).
	bool nearest(const float *table, long long table_rows, const float *d_points, long long npts, int32_t *d_out_idx,
	             Workspace &ws, cudaStream_t s, int64_t *launches);
	static bool nearest_supported(int dim, int metric, long long table_rows);

	struct Impl; // opaque (public only so that file-local helpers can name it)

private:
	void rerun_flagged(const float *X, long long n, const uint32_t *deleted, const float *Qchunk, int nq, int k,
	                   const uint32_t *d_allow, long long allow_bits, long long label_base, int64_t *d_out_labels,
	                   float *d_out_dist, int32_t *d_out_counts, const uint32_t *flag_out, Workspace &ws, cudaStream_t s,
	                   int64_t *launches, int64_t *fallback_queries);
	Impl *impl_;
};

} // namespace lb
