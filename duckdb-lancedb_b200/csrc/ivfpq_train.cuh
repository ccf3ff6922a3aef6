// ivfpq_train.cuh — GPU trainer behind lance_detached_create_index (included at the end of ivfpq.cu).
//
// Stands in for `IvfPqIndexBuilder` (rust_lib/src/lance_manager.rs:483-515; k-means + PQ training run
// inside Lance, not in the reference tree).  Same scheme as the checker's oracle_ivfpq_train
// (oracle/oracle.c): Lloyd k-means for the P coarse centroids (on unit rows for cosine), residuals
// (l2 / cosine) or raw rows (dot), one 256-centre k-means per sub-space, encode, partition-major order
// with ascending labels inside a partition.  Differences, all outside the parity contract because the
// trained index is exported to the checker (lance_b200_get_ivfpq) rather than re-derived: training runs
// on a strided sample of at most 256*P rows, and centre sums use fp32 atomics.
#pragma once

// `tc` (nullable, whole-row k-means only): the assignment step runs on the tensor cores (FlatGemm::nearest) — the
// centres are re-converted to their bf16 shadow every iteration, the points stream through as query tiles
static void kmeans_device(const float *X, long long n, int ldx, int dd, int Z, float *C, int nc, int is_dot,
                          const float *sub_base, const int32_t *sub_idx, int iters, int32_t *assign, float *sums,
                          int *counts, cudaStream_t s, int64_t *launches, FlatGemm *tc = nullptr, Workspace *tc_ws = nullptr) {
	for (int it = 0; it < iters; it++) {
		bool assigned = false;
		if (tc && Z == 1 && ldx == dd && !sub_base) {
			tc->invalidate();
			assigned = tc->nearest(C, nc, X, n, assign, *tc_ws, s, launches);
		}
		if (!assigned)
			nearest_center_kernel<<<dim3(blocks_for(n, NC_BM), 1, Z), 256, 0, s>>>(X, n, ldx, dd, C, nc, is_dot, sub_base,
			                                                                        sub_idx, assign, n);
		LB_CUDA(cudaGetLastError());
		LB_CUDA(cudaMemsetAsync(sums, 0, (size_t)Z * nc * dd * sizeof(float), s));
		LB_CUDA(cudaMemsetAsync(counts, 0, (size_t)Z * nc * sizeof(int), s));
		accumulate_centers_kernel<<<dim3(blocks_for(n * dd, 256), 1, Z), 256, 0, s>>>(X, n, ldx, dd, sub_base, sub_idx,
		                                                                               assign, n, nc, sums, counts);
		LB_CUDA(cudaGetLastError());
		finish_centers_kernel<<<blocks_for((long long)Z * nc * dd, 256), 256, 0, s>>>(C, sums, counts,
		                                                                              (long long)Z * nc * dd, dd);
		LB_CUDA(cudaGetLastError());
		if (launches) *launches += 3;
	}
}

// initial PQ centres: sub-vectors (residual or raw) of sample rows i*ns/256
__global__ void init_codebook_kernel(const float *__restrict__ S, long long ns, int d, int M, int dsub,
                                     const float *__restrict__ sub_base, const int32_t *__restrict__ sub_idx,
                                     float *__restrict__ codebook) {
	long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (long long)M * 256 * dsub) return;
	int j = (int)(t % dsub);
	int c = (int)((t / dsub) % 256);
	int m = (int)(t / ((long long)dsub * 256));
	long long row = ns >= 256 ? (long long)c * ns / 256 : c % ns;
	float v = S[(size_t)row * d + m * dsub + j];
	if (sub_base) v -= sub_base[(size_t)sub_idx[row] * d + m * dsub + j];
	codebook[t] = v;
}

void Store::create_index(int num_partitions, int num_sub_vectors) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	const long long n = n_rows_;
	const int d = dim_;
	if (n <= 0) throw Error("cannot train an IVF_PQ index on an empty table");
	int P = num_partitions > 0 ? num_partitions : (int)std::max(1.0, std::floor(std::sqrt((double)n) + 0.5));
	if (P > n) P = (int)n;
	if (n >= (1ll << 31) - 64) throw Error("IVF_PQ training supports fewer than 2^31 rows per device store");
	int M = num_sub_vectors > 0 ? num_sub_vectors : default_sub_vectors(d);
	check_shape(d, P, M);
	const int dsub = d / M;
	const bool is_dot = metric_ == METRIC_DOT;
	const bool normalise = metric_ == METRIC_COSINE;
	int iters = 10;
	if (const char *e = getenv("LANCE_B200_KMEANS_ITERS")) iters = std::max(1, atoi(e));
	cudaStream_t s = copy_stream_;
	int64_t launches = 0;

	std::unique_ptr<IvfPqIndexFull> ix(new IvfPqIndexFull());
	ix->metric = metric_;
	ix->dim = d, ix->P = P, ix->M = M, ix->dsub = dsub, ix->n_indexed = n;
	ix->centroids.alloc((size_t)P * d * sizeof(float));
	ix->codebook.alloc((size_t)M * 256 * dsub * sizeof(float));

	// ---- training sample (strided), unit rows for cosine
	const long long ns = std::min<long long>(n, 256ll * P);
	const long long stride = n / ns;
	DevBuf S, assign, sums, counts;
	S.alloc((size_t)ns * d * sizeof(float));
	gather_rows_kernel<<<blocks_for(ns, 8), 256, 0, s>>>(X_.as<float>(), d, ns, stride, normalise ? 1 : 0, S.as<float>());
	LB_CUDA(cudaGetLastError());
	// initial centroids: sample rows i*ns/P
	gather_rows_kernel<<<blocks_for(P, 8), 256, 0, s>>>(S.as<float>(), d, P, ns / P, 0, ix->centroids.as<float>());
	LB_CUDA(cudaGetLastError());
	launches += 2;
	const long long ns_pq = std::min<long long>(ns, 65536);
	assign.alloc((size_t)std::max<long long>(ns, (long long)M * ns_pq) * sizeof(int32_t));
	sums.alloc((size_t)std::max<long long>((long long)P * d, (long long)M * 256 * dsub) * sizeof(float));
	counts.alloc((size_t)std::max<long long>(P, (long long)M * 256) * sizeof(int));
	// The partition-level assignments (k-means over the sample, then every row of the table) are 2 n P d flops — 64 + 63
	// TFLOP at C3, 7.8 s of an 8.5 s build on the CUDA cores: they run on the tensor cores when the shape has a
	// tensor-memory kernel (FlatGemm::nearest: bf16 tile maxima, then the exact fp32 distance to the rows of the four best
	// tiles).  LANCE_B200_TRAIN_TC=0 keeps the CUDA-core kernel.
	std::unique_ptr<FlatGemm> tc;
	std::unique_ptr<Workspace> tc_ws;
	{
		const char *e = getenv("LANCE_B200_TRAIN_TC");
		const int train_metric = is_dot ? METRIC_DOT : METRIC_L2;
		if (!(e && atoi(e) == 0) && d >= 64 && FlatGemm::nearest_supported(d, train_metric, P)) {
			tc.reset(new FlatGemm(device_, num_sms_, d, train_metric));
			tc_ws.reset(new Workspace(false));
		}
	}
	kmeans_device(S.as<float>(), ns, d, d, 1, ix->centroids.as<float>(), P, is_dot, nullptr, nullptr, iters,
	              assign.as<int32_t>(), sums.as<float>(), counts.as<int>(), s, &launches, tc.get(), tc_ws.get());
	if (tc) tc->invalidate(); // the last iteration moved the centres once more

	// ---- PQ codebooks on residuals (l2 / cosine) or raw rows (dot)
	DevBuf part_of_sample;
	part_of_sample.alloc((size_t)ns_pq * sizeof(int32_t));
	nearest_center_kernel<<<dim3(blocks_for(ns_pq, NC_BM), 1, 1), 256, 0, s>>>(S.as<float>(), ns_pq, d, d,
	                                                                           ix->centroids.as<float>(), P, is_dot, nullptr,
	                                                                           nullptr, part_of_sample.as<int32_t>(), ns_pq);
	LB_CUDA(cudaGetLastError());
	const float *sub_base = is_dot ? nullptr : ix->centroids.as<float>();
	const int32_t *sub_idx = is_dot ? nullptr : part_of_sample.as<int32_t>();
	init_codebook_kernel<<<blocks_for((long long)M * 256 * dsub, 256), 256, 0, s>>>(S.as<float>(), ns_pq, d, M, dsub, sub_base,
	                                                                                sub_idx, ix->codebook.as<float>());
	LB_CUDA(cudaGetLastError());
	launches += 2;
	kmeans_device(S.as<float>(), ns_pq, d, dsub, M, ix->codebook.as<float>(), 256, 0, sub_base, sub_idx, iters,
	              assign.as<int32_t>(), sums.as<float>(), counts.as<int>(), s, &launches);
	LB_CUDA(cudaStreamSynchronize(s));
	S.release();

	// ---- encode every row, in chunks
	const long long CH = std::min<long long>(n, 262144);
	DevBuf part_assign, sub_assign, codes_rm, chunk_n;
	part_assign.alloc((size_t)n * sizeof(int32_t));
	sub_assign.alloc((size_t)M * CH * sizeof(int32_t));
	codes_rm.alloc((size_t)n * M);
	if (normalise) chunk_n.alloc((size_t)CH * d * sizeof(float));
	for (long long r0 = 0; r0 < n; r0 += CH) {
		const long long c = std::min(CH, n - r0);
		const float *Xc = X_.as<float>() + (size_t)r0 * d;
		if (normalise) {
			gather_rows_kernel<<<blocks_for(c, 8), 256, 0, s>>>(Xc, d, c, 1, 1, chunk_n.as<float>());
			LB_CUDA(cudaGetLastError());
			Xc = chunk_n.as<float>();
			launches++;
		}
		int32_t *pa = part_assign.as<int32_t>() + r0;
		if (!(tc && tc->nearest(ix->centroids.as<float>(), P, Xc, c, pa, *tc_ws, s, &launches)))
			nearest_center_kernel<<<dim3(blocks_for(c, NC_BM), 1, 1), 256, 0, s>>>(Xc, c, d, d, ix->centroids.as<float>(), P,
			                                                                       is_dot, nullptr, nullptr, pa, c);
		LB_CUDA(cudaGetLastError());
		nearest_center_kernel<<<dim3(blocks_for(c, NC_BM), 1, M), 256, 0, s>>>(Xc, c, d, dsub, ix->codebook.as<float>(), 256, 0,
		                                                                       is_dot ? nullptr : ix->centroids.as<float>(),
		                                                                       is_dot ? nullptr : pa, sub_assign.as<int32_t>(), c);
		LB_CUDA(cudaGetLastError());
		pack_codes_kernel<<<blocks_for(c * M, 256), 256, 0, s>>>(sub_assign.as<int32_t>(), c, M,
		                                                         codes_rm.as<uint8_t>() + (size_t)r0 * M);
		LB_CUDA(cudaGetLastError());
		launches += 3;
	}

	// ---- partition-major order (stable sort: ascending rows inside a partition, as the checker's trainer)
	DevBuf order_in, order_out, keys_out, hist, tmp;
	order_in.alloc((size_t)n * sizeof(int32_t));
	order_out.alloc((size_t)n * sizeof(int32_t));
	keys_out.alloc((size_t)n * sizeof(int32_t));
	hist.alloc((size_t)P * sizeof(int));
	iota_kernel<<<blocks_for(n, 256), 256, 0, s>>>(order_in.as<int32_t>(), n, 0);
	LB_CUDA(cudaGetLastError());
	size_t tmp_bytes = 0;
	int end_bit = 1;
	while ((1ll << end_bit) < P) end_bit++;
	LB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, part_assign.as<int32_t>(), keys_out.as<int32_t>(),
	                                        order_in.as<int32_t>(), order_out.as<int32_t>(), (int)n, 0, end_bit, s));
	tmp.alloc(tmp_bytes + 16);
	LB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, part_assign.as<int32_t>(), keys_out.as<int32_t>(),
	                                        order_in.as<int32_t>(), order_out.as<int32_t>(), (int)n, 0, end_bit, s));
	LB_CUDA(cudaMemsetAsync(hist.p, 0, (size_t)P * sizeof(int), s));
	histogram_kernel<<<blocks_for(n, 256), 256, 0, s>>>(part_assign.as<int32_t>(), n, hist.as<int>());
	LB_CUDA(cudaGetLastError());
	std::vector<int> hh(P);
	LB_CUDA(cudaMemcpyAsync(hh.data(), hist.p, (size_t)P * sizeof(int), cudaMemcpyDeviceToHost, s));
	LB_CUDA(cudaStreamSynchronize(s));
	ix->part_offsets.assign(P + 1, 0);
	for (int p = 0; p < P; p++) ix->part_offsets[p + 1] = ix->part_offsets[p] + hh[p];
	ix->ex.part_offsets_dev.alloc((size_t)(P + 1) * sizeof(int64_t));
	LB_CUDA(cudaMemcpyAsync(ix->ex.part_offsets_dev.p, ix->part_offsets.data(), (size_t)(P + 1) * sizeof(int64_t),
	                        cudaMemcpyHostToDevice, s));
	ix->ex.row_ids.alloc((size_t)n * sizeof(int64_t));
	ix->ex.codes.alloc((size_t)n * M);
	gather_sorted_kernel<<<blocks_for(n * M, 256), 256, 0, s>>>(codes_rm.as<uint8_t>(), order_out.as<int32_t>(), n, M,
	                                                            label_base_, ix->ex.codes.as<uint8_t>(),
	                                                            ix->ex.row_ids.as<int64_t>());
	LB_CUDA(cudaGetLastError());
	launches += 4;
	finalize_index(*ix, ix->ex, label_base_, s, &launches);
	add_launches(launches);
	LB_CUDA(cudaDeviceSynchronize());
	index_.reset(ix.release()); // .replace(true): a new index supersedes the old one (lance_manager.rs:508)
}
