// ivfpq_host.cuh — host side of the IVF_PQ path (included at the end of ivfpq.cu, inside namespace lb):
// index install from the exchange layout, the search pipeline, the trainer.
#pragma once

static inline unsigned blocks_for(long long n, int per) { return (unsigned)((n + per - 1) / per); }

// ------------------------------------------------------------------ install
// Builds the scan layout from exchange-layout arrays that already sit on the device in ix.
struct IvfPqExchange {
	DevBuf row_ids; // int64 [n_indexed] labels, partition-major
	DevBuf codes;   // u8 [n_indexed][M], partition-major
	DevBuf part_offsets_dev; // int64 [P+1]
};

static void finalize_index(IvfPqIndex &ix, IvfPqExchange &ex, long long label_base, cudaStream_t s, int64_t *launches) {
	const int P = ix.P, M = ix.M;
	ix.gran = (M % 16 == 0) ? 16 : (M % 4 == 0 ? 4 : 1);
	if (M % 32 == 0) ix.gran = 32; // rotated granules + transposed table: no shared-memory bank conflicts in the scan
	if (const char *e = getenv("LANCE_B200_ADC_ROT")) // tuning hook (A/B of the two layouts)
		if (atoi(e) == 0 && ix.gran == 32) ix.gran = 16;
	std::vector<long long> pad(P + 1, 0);
	ix.max_part_rows = 0;
	for (int p = 0; p < P; p++) {
		long long sz = ix.part_offsets[p + 1] - ix.part_offsets[p];
		if (sz < 0) throw Error("part_offsets must be non-decreasing");
		ix.max_part_rows = std::max<long long>(ix.max_part_rows, sz);
		pad[p + 1] = pad[p] + ((sz + 31) & ~31ll);
	}
	ix.n_pad = pad[P];
	ix.pad_offsets.alloc((size_t)(P + 1) * sizeof(long long));
	LB_CUDA(cudaMemcpyAsync(ix.pad_offsets.p, pad.data(), (size_t)(P + 1) * sizeof(long long), cudaMemcpyHostToDevice, s));
	const long long np = std::max<long long>(ix.n_pad, 32);
	ix.slot_rows.alloc((size_t)np * sizeof(int32_t));
	ix.codes_il.alloc((size_t)np * M);
	ix.row_term.alloc((size_t)np * sizeof(float));
	DevBuf slot_src;
	slot_src.alloc((size_t)np * sizeof(long long));
	if (ix.n_pad > 0) {
		build_slots_kernel<<<blocks_for(ix.n_pad, 256), 256, 0, s>>>(
		    ex.part_offsets_dev.as<long long>(), ix.pad_offsets.as<long long>(), P, ex.row_ids.as<int64_t>(), label_base,
		    ix.n_pad, ix.slot_rows.as<int32_t>(), slot_src.as<long long>());
		LB_CUDA(cudaGetLastError());
		interleave_codes_kernel<<<blocks_for(ix.n_pad * M, 256), 256, 0, s>>>(ex.codes.as<uint8_t>(), slot_src.as<long long>(),
		                                                                    M, ix.gran, ix.n_pad, ix.codes_il.as<uint8_t>());
		LB_CUDA(cudaGetLastError());
		if (ix.metric != METRIC_DOT) {
			row_term_kernel<<<blocks_for(ix.n_pad, 8), 256, 0, s>>>(ex.codes.as<uint8_t>(), slot_src.as<long long>(),
			                                                       ix.pad_offsets.as<long long>(), P, ix.centroids.as<float>(),
			                                                       ix.codebook.as<float>(), ix.dim, M, ix.dsub, ix.n_pad,
			                                                       ix.row_term.as<float>());
			LB_CUDA(cudaGetLastError());
		} else {
			LB_CUDA(cudaMemsetAsync(ix.row_term.p, 0, (size_t)np * sizeof(float), s));
		}
		if (launches) *launches += 3;
	}
	LB_CUDA(cudaStreamSynchronize(s));
}

// exchange copies are kept so that lance_b200_get_ivfpq can hand the index back (oracle parity tests)
struct IvfPqIndexFull : IvfPqIndex {
	IvfPqExchange ex;
};

static void check_shape(int dim, int P, int M) {
	if (P <= 0 || M <= 0) throw Error("num_partitions and num_sub_vectors must be positive");
	if (dim % M != 0) throw Error("dimension " + std::to_string(dim) + " is not divisible by num_sub_vectors " + std::to_string(M));
	if ((size_t)M * 1024 + 8192 > 200 * 1024) throw Error("num_sub_vectors > 192 does not fit the shared-memory lookup table");
	// lut_kernel keeps one sub-quantizer's codebook slice [256][dsub + 1] and LUT_QPB query sub-vectors in shared memory
	const int dsub = dim / M;
	if (((size_t)256 * (dsub + 1) + (size_t)LUT_QPB * dsub) * sizeof(float) > 220 * 1024)
		throw Error("sub-vector dimension " + std::to_string(dsub) + " (dimension / num_sub_vectors) is too large: use at least " +
		            std::to_string((dim + 199) / 200) + " sub-vectors");
}

// num_sub_vectors = 0: lancedb's rule (dim/16, else dim/8, else 1: lance_manager.rs:483-515) when the device scan
// supports the result, else the largest sub-vector width >= 16 that divides the dimension and fits
static int default_sub_vectors(int d) {
	int M = d % 16 == 0 ? d / 16 : (d % 8 == 0 ? d / 8 : 1);
	auto fits = [&](int m) {
		const int dsub = d / m;
		return (size_t)m * 1024 + 8192 <= 200 * 1024 && ((size_t)256 * (dsub + 1) + (size_t)LUT_QPB * dsub) * sizeof(float) <= 220 * 1024;
	};
	if (fits(M)) return M;
	for (int dsub = 200; dsub >= 1; dsub--)
		if (d % dsub == 0 && fits(d / dsub)) return d / dsub;
	return M; // check_shape reports why
}

void Store::set_index(int P, int M, int64_t n_indexed, const float *centroids, const float *codebook,
                      const int64_t *part_offsets, const int64_t *row_ids, const uint8_t *codes) {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	install_index_locked(P, M, n_indexed, centroids, codebook, part_offsets, row_ids, codes);
}

// row_ids are labels of the implicit numbering (label_base_ + row); on a store with explicit labels (compacted /
// sharded) label_base_ is 0 and the exchange layout is in ROW numbers
void Store::install_index_locked(int P, int M, int64_t n_indexed, const float *centroids, const float *codebook,
                                 const int64_t *part_offsets, const int64_t *row_ids, const uint8_t *codes) {
	check_shape(dim_, P, M);
	if (n_indexed < 0 || n_indexed > n_rows_) throw Error("index covers more rows than the table holds");
	if (part_offsets[0] != 0 || part_offsets[P] != n_indexed) throw Error("part_offsets must span [0, n_indexed]");
	{
		// the tail scan assumes the indexed rows are exactly [0, n_indexed): row_ids must be a permutation of
		// [label_base, label_base + n_indexed)
		std::vector<bool> seen((size_t)n_indexed, false);
		for (int64_t i = 0; i < n_indexed; i++) {
			const int64_t r = row_ids[i] - label_base_;
			if (r < 0 || r >= n_rows_) throw Error("row id outside the table");
			if (r >= n_indexed || seen[(size_t)r])
				throw Error("row_ids must be a permutation of the first n_indexed labels of the table");
			seen[(size_t)r] = true;
		}
	}
	std::unique_ptr<IvfPqIndexFull> ix(new IvfPqIndexFull());
	ix->metric = metric_;
	ix->dim = dim_, ix->P = P, ix->M = M, ix->dsub = dim_ / M, ix->n_indexed = n_indexed;
	ix->part_offsets.assign(part_offsets, part_offsets + P + 1);
	cudaStream_t s = copy_stream_;
	ix->centroids.alloc((size_t)P * dim_ * sizeof(float));
	ix->codebook.alloc((size_t)M * 256 * ix->dsub * sizeof(float));
	ix->ex.row_ids.alloc(std::max<size_t>((size_t)n_indexed, 1) * sizeof(int64_t));
	ix->ex.codes.alloc(std::max<size_t>((size_t)n_indexed * M, 1));
	ix->ex.part_offsets_dev.alloc((size_t)(P + 1) * sizeof(int64_t));
	LB_CUDA(cudaMemcpyAsync(ix->centroids.p, centroids, (size_t)P * dim_ * sizeof(float), cudaMemcpyHostToDevice, s));
	LB_CUDA(cudaMemcpyAsync(ix->codebook.p, codebook, (size_t)M * 256 * ix->dsub * sizeof(float), cudaMemcpyHostToDevice, s));
	LB_CUDA(cudaMemcpyAsync(ix->ex.row_ids.p, row_ids, (size_t)n_indexed * sizeof(int64_t), cudaMemcpyHostToDevice, s));
	LB_CUDA(cudaMemcpyAsync(ix->ex.codes.p, codes, (size_t)n_indexed * M, cudaMemcpyHostToDevice, s));
	LB_CUDA(cudaMemcpyAsync(ix->ex.part_offsets_dev.p, part_offsets, (size_t)(P + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
	int64_t l = 0;
	finalize_index(*ix, ix->ex, label_base_, s, &l);
	add_launches(l);
	index_.reset(ix.release());
}

// The index restricted to the rows that survive a compaction, in the exchange layout with ROW numbers of the compacted
// table (order inside a partition is kept).  Exclusive lock held; nothing is modified.
void Store::extract_index_for_compaction(const std::vector<int64_t> &new_row_of, CompactedIndex &out) {
	IvfPqIndexFull &ix = static_cast<IvfPqIndexFull &>(*index_);
	out.present = true;
	out.P = ix.P, out.M = ix.M;
	out.centroids.resize((size_t)ix.P * ix.dim);
	out.codebook.resize((size_t)ix.M * 256 * ix.dsub);
	std::vector<int64_t> ids((size_t)ix.n_indexed);
	std::vector<uint8_t> codes((size_t)ix.n_indexed * ix.M);
	LB_CUDA(cudaMemcpy(out.centroids.data(), ix.centroids.p, out.centroids.size() * sizeof(float), cudaMemcpyDeviceToHost));
	LB_CUDA(cudaMemcpy(out.codebook.data(), ix.codebook.p, out.codebook.size() * sizeof(float), cudaMemcpyDeviceToHost));
	if (ix.n_indexed > 0) {
		LB_CUDA(cudaMemcpy(ids.data(), ix.ex.row_ids.p, ids.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
		LB_CUDA(cudaMemcpy(codes.data(), ix.ex.codes.p, codes.size(), cudaMemcpyDeviceToHost));
	}
	out.part_offsets.assign((size_t)ix.P + 1, 0);
	out.row_ids.reserve(ids.size());
	out.codes.reserve(codes.size());
	for (int p = 0; p < ix.P; p++) {
		for (int64_t i = ix.part_offsets[(size_t)p]; i < ix.part_offsets[(size_t)p + 1]; i++) {
			const int64_t nr = new_row_of[(size_t)(ids[(size_t)i] - label_base_)];
			if (nr < 0) continue;
			out.row_ids.push_back(nr);
			out.codes.insert(out.codes.end(), codes.begin() + (size_t)i * ix.M, codes.begin() + (size_t)(i + 1) * ix.M);
		}
		out.part_offsets[(size_t)p + 1] = (int64_t)out.row_ids.size();
	}
	out.n_indexed = (int64_t)out.row_ids.size();
}

void Store::get_index(lance_b200_ivfpq_info *info, float *centroids, float *codebook, int64_t *part_offsets,
                      int64_t *row_ids, uint8_t *codes) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (!index_) throw Error("no IVF_PQ index on this handle");
	IvfPqIndexFull &ix = static_cast<IvfPqIndexFull &>(*index_);
	if (info) {
		info->metric = ix.metric, info->dim = ix.dim, info->P = ix.P, info->M = ix.M, info->n_indexed = ix.n_indexed;
	}
	if (centroids) LB_CUDA(cudaMemcpy(centroids, ix.centroids.p, (size_t)ix.P * ix.dim * sizeof(float), cudaMemcpyDeviceToHost));
	if (codebook) LB_CUDA(cudaMemcpy(codebook, ix.codebook.p, (size_t)ix.M * 256 * ix.dsub * sizeof(float), cudaMemcpyDeviceToHost));
	if (part_offsets) memcpy(part_offsets, ix.part_offsets.data(), (size_t)(ix.P + 1) * sizeof(int64_t));
	if (row_ids) LB_CUDA(cudaMemcpy(row_ids, ix.ex.row_ids.p, (size_t)ix.n_indexed * sizeof(int64_t), cudaMemcpyDeviceToHost));
	if (codes) LB_CUDA(cudaMemcpy(codes, ix.ex.codes.p, (size_t)ix.n_indexed * ix.M, cudaMemcpyDeviceToHost));
}

// ------------------------------------------------------------------ persistence of the index (store.cu: save / open)
// int64 ihdr[4] = {P, M, n_indexed, metric} | centroids fp32 [P][dim] | codebook fp32 [M][256][dim/M]
// | part_offsets int64 [P+1] | row_ids int64 [n_indexed] | codes u8 [n_indexed][M]   (the exchange layout)
static void write_device_array(FILE *f, const void *d_ptr, size_t bytes) {
	const size_t chunk = 64u << 20;
	std::vector<unsigned char> buf(std::min(bytes, chunk));
	for (size_t off = 0; off < bytes; off += chunk) {
		const size_t c = std::min(chunk, bytes - off);
		LB_CUDA(cudaMemcpy(buf.data(), static_cast<const unsigned char *>(d_ptr) + off, c, cudaMemcpyDeviceToHost));
		if (fwrite(buf.data(), 1, c, f) != c) throw Error("short write of the IVF_PQ index");
	}
}

void Store::save_index(FILE *f) {
	IvfPqIndexFull &ix = static_cast<IvfPqIndexFull &>(*index_);
	int64_t ihdr[4] = {ix.P, ix.M, ix.n_indexed, ix.metric};
	if (fwrite(ihdr, sizeof(int64_t), 4, f) != 4) throw Error("short write of the IVF_PQ index");
	write_device_array(f, ix.centroids.p, (size_t)ix.P * ix.dim * sizeof(float));
	write_device_array(f, ix.codebook.p, (size_t)ix.M * 256 * ix.dsub * sizeof(float));
	if (fwrite(ix.part_offsets.data(), sizeof(int64_t), (size_t)ix.P + 1, f) != (size_t)ix.P + 1)
		throw Error("short write of the IVF_PQ index");
	write_device_array(f, ix.ex.row_ids.p, (size_t)ix.n_indexed * sizeof(int64_t));
	write_device_array(f, ix.ex.codes.p, (size_t)ix.n_indexed * ix.M);
}

void Store::load_index(FILE *f, const std::string &path) {
	int64_t ihdr[4];
	if (fread(ihdr, sizeof(int64_t), 4, f) != 4) throw Error("truncated IVF_PQ index: " + path);
	const int64_t P = ihdr[0], M = ihdr[1], ni = ihdr[2];
	if (P <= 0 || P > (1 << 24) || M <= 0 || M > 4096 || ni < 0 || ni > n_rows_ || dim_ % M != 0)
		throw Error("corrupt IVF_PQ index header: " + path);
	const int dsub = dim_ / (int)M;
	std::vector<float> cen((size_t)P * dim_), cb((size_t)M * 256 * dsub);
	std::vector<int64_t> off((size_t)P + 1), ids((size_t)ni);
	std::vector<uint8_t> codes((size_t)ni * M);
	bool ok = fread(cen.data(), sizeof(float), cen.size(), f) == cen.size() && fread(cb.data(), sizeof(float), cb.size(), f) == cb.size() &&
	          fread(off.data(), sizeof(int64_t), off.size(), f) == off.size() && fread(ids.data(), sizeof(int64_t), ids.size(), f) == ids.size() &&
	          fread(codes.data(), 1, codes.size(), f) == codes.size();
	if (!ok) throw Error("truncated IVF_PQ index: " + path);
	// the same entry the checker-trained indexes come through: validates offsets and row ids, builds the scan layout
	set_index((int)P, (int)M, ni, cen.data(), cb.data(), off.data(), ids.data(), codes.data());
}

void Store::drop_index() {
	std::unique_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	LB_CUDA(cudaDeviceSynchronize());
	index_.reset();
}

// ------------------------------------------------------------------ search
template <int GRAN, int NB = 0>
static void launch_adc(const AdcParams &ap, int nwarps, int S, int nq, size_t smem, cudaStream_t s) {
	auto kern = adc_scan_kernel<GRAN, NB>;
	if (smem > 48 * 1024) LB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	kern<<<dim3(S, nq), nwarps * 32, smem, s>>>(ap);
	LB_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------ one query, one cooperative launch (ivfpq1.cuh)
template <int NB>
static void launch_ivfpq1(const Ivf1Params &p, int grid, size_t smem, cudaStream_t s) {
	static std::mutex mu;
	static size_t granted_by_device[64] = {0}; // the attribute is per device (in-process sharded handles)
	{
		std::lock_guard<std::mutex> g(mu);
		int dev = 0;
		LB_CUDA(cudaGetDevice(&dev));
		size_t &granted = granted_by_device[dev & 63];
		if (granted == 0) granted = 48 * 1024;
		if (smem > granted) {
			LB_CUDA(cudaFuncSetAttribute(ivfpq1_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
			granted = smem;
		}
	}
	void *args[] = {const_cast<Ivf1Params *>(&p)};
	LB_CUDA(cudaLaunchCooperativeKernel((const void *)ivfpq1_kernel<NB>, dim3(grid), dim3(IVF1_NT), args, smem, s));
}

static bool ivfpq1_eligible(const IvfPqIndex &ix, int nq, int np, long long kc, int d, bool has_tail) {
	static const bool off = [] { // A/B hook
		const char *e = getenv("LANCE_B200_IVF1");
		return e && atoi(e) == 0;
	}();
	if (off || nq != 1 || has_tail) return false;
	if (ix.gran != 32 || ix.M / 32 > 4) return false;       // rotated layout, table <= 128 KB
	if (ix.P > IVF1_MAXKEYS * IVF1_NT || np > 32 || kc > 32) return false;
	return ivf1_smem(d, ix.M).total <= 220 * 1024;
}

bool Store::ivfpq_takes_inline_query(int k, int nprobes, int refine) {
	const IvfPqIndex &ix = *index_;
	const int np = std::min(std::max(nprobes, 1), ix.P);
	const long long kc = (long long)k * std::max(refine, 1);
	return ivfpq1_eligible(ix, 1, np, kc, dim_, ix.n_indexed < n_rows_);
}

void Store::ivfpq_search_locked(const float *d_queries, int nq, int k, int nprobes, int refine, const uint32_t *d_allow,
                                int64_t allow_bits, int64_t *d_out_labels, float *d_out_dist, int32_t *d_out_counts,
                                Workspace &ws, cudaStream_t s, int32_t *d_dbg_parts, u64 *d_dbg_cand, int32_t *d_dbg_ncand,
                                HostHints *hints) {
	IvfPqIndex &ix = *index_;
	int64_t launches = 0;
	const int d = dim_, M = ix.M;
	int np = nprobes < 1 ? 1 : nprobes; // oracle_ivfpq_search clamps the same way
	if (np > ix.P) np = ix.P;
	const int rf = refine < 1 ? 1 : refine;
	const long long kc_ll = (long long)k * rf;
	if (kc_ll > 8192) throw Error("k * refine_factor > 8192 is not supported by the device ADC scan");
	const int kc = (int)kc_ll;
	const uint32_t *deleted = n_deleted_ > 0 ? deleted_.as<uint32_t>() : nullptr;

	if (ivfpq1_eligible(ix, nq, np, kc_ll, d, ix.n_indexed < n_rows_)) {
		// the single query of the SQL surface: probe, table, scan, refine and select in ONE cooperative launch
		Ivf1Params p;
		const bool inline_q = hints && !hints->q_uploaded && hints->h_queries && d <= SCAN1_QINLINE_MAX;
		p.q_dev = inline_q ? nullptr : d_queries;
		if (inline_q) memcpy(p.qinline, hints->h_queries, (size_t)d * sizeof(float));
		p.d = d, p.metric = ix.metric;
		p.centroids = ix.centroids.as<float>(), p.P = ix.P, p.M = M, p.dsub = ix.dsub, p.codebook = ix.codebook.as<float>();
		p.pad_offsets = ix.pad_offsets.as<long long>(), p.slot_rows = ix.slot_rows.as<int32_t>();
		p.codes_il = ix.codes_il.as<uint8_t>(), p.row_term = ix.row_term.as<float>();
		p.allow = d_allow, p.allow_bits = allow_bits, p.deleted = deleted;
		p.nprobes = np, p.k = k, p.kc = kc, p.refine = refine >= 1 ? 1 : 0;
		const int grid = num_sms_;
		ws.work2.ensure(64 * sizeof(u64));                      // the sorted probe keys (read by the debug entry)
		ws.work3.ensure((size_t)256 * M * sizeof(float));
		ws.work4.ensure(((size_t)ix.P + (size_t)grid * 32 + 64) * sizeof(u64));
		if (!ws.ivf1_hdr.p) {
			ws.ivf1_hdr.alloc(64);
			LB_CUDA(cudaMemsetAsync(ws.ivf1_hdr.p, 0, 64, s));
		}
		p.probe_sorted = ws.work2.as<u64>();
		p.lut_g = ws.work3.as<float>();
		p.probe_keys = ws.work4.as<u64>();
		p.slots = p.probe_keys + ix.P;
		p.cand = p.slots + (size_t)grid * 32;
		p.hdr = ws.ivf1_hdr.as<u64>();
		p.X = X_.as<float>(), p.label_base = label_base_;
		p.out_labels = d_out_labels, p.out_dist = d_out_dist, p.out_counts = d_out_counts;
		p.done_flag = hints ? hints->done_flag : nullptr;
		p.done_value = hints ? hints->done_value : 0;
		if (hints) hints->armed = hints->done_flag != nullptr;
		const size_t smem = ivf1_smem(d, M).total;
		int tix = timer_.begin(KC_ADC_SCAN, s);
		switch (M / 32) {
		case 1: launch_ivfpq1<1>(p, grid, smem, s); break;
		case 2: launch_ivfpq1<2>(p, grid, smem, s); break;
		case 3: launch_ivfpq1<3>(p, grid, smem, s); break;
		default: launch_ivfpq1<4>(p, grid, smem, s);
		}
		timer_.end(tix, s);
		launches++;
		if (d_dbg_cand) LB_CUDA(cudaMemcpyAsync(d_dbg_cand, p.cand, (size_t)kc * sizeof(u64), cudaMemcpyDeviceToDevice, s));
		std::lock_guard<std::mutex> g(stat_mu_);
		ivfpq_batches_++;
		launches_ += launches;
		return;
	}

	// 1. q' and |q'|^2
	ws.q.ensure((size_t)nq * d * sizeof(float) + (size_t)nq * sizeof(float) + 256);
	float *Qp = ws.q.as<float>();
	float *qn2 = Qp + (size_t)nq * d;
	prep_queries_kernel<<<blocks_for(nq, 8), 256, 0, s>>>(d_queries, nq, d, metric_ == METRIC_COSINE ? 1 : 0, Qp, qn2);
	LB_CUDA(cudaGetLastError());
	launches++;

	// 2. probe: the nprobes nearest centroids = a flat scan over the centroid table (a12)
	ws.work2.ensure((size_t)nq * np * sizeof(u64));
	u64 *probe_keys = ws.work2.as<u64>();
	if (nq >= 8) {
		// batches: tiled fp32 contraction against the (L2-resident) centroid table, then one select per query
		const int chunk = (int)std::max<long long>(64, std::min<long long>(nq, (256ll << 20) / ((long long)ix.P * 8)));
		ws.probe_mat.ensure((size_t)std::min(nq, chunk) * ix.P * sizeof(u64));
		for (int q0 = 0; q0 < nq; q0 += chunk) {
			const int nc = std::min(chunk, nq - q0);
			dim3 grid((ix.P + 63) / 64, (nc + 63) / 64);
			if (metric_ == METRIC_DOT)
				probe_tile_kernel<1><<<grid, 256, 0, s>>>(Qp + (size_t)q0 * d, nc, ix.centroids.as<float>(), ix.P, d, ws.probe_mat.as<u64>());
			else
				probe_tile_kernel<0><<<grid, 256, 0, s>>>(Qp + (size_t)q0 * d, nc, ix.centroids.as<float>(), ix.P, d, ws.probe_mat.as<u64>());
			LB_CUDA(cudaGetLastError());
			launches++;
			launch_topk_merge(ws.probe_mat.as<u64>(), nc, ix.P, np, 0, nullptr, nullptr, nullptr, probe_keys + (size_t)q0 * np, s, &launches);
		}
	} else {
		// a few queries: one warp per centroid scores all of them, then one select per query
		ws.probe_mat.ensure((size_t)nq * ix.P * sizeof(u64));
		const size_t qsm = (size_t)nq * d * sizeof(float);
		if (qsm > 200 * 1024) throw Error("dimension too large for the IVF probe");
		const int blocks = (ix.P + 7) / 8;
		if (metric_ == METRIC_DOT) {
			if (qsm > 48 * 1024) LB_CUDA(cudaFuncSetAttribute(probe_rows_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsm));
			probe_rows_kernel<1><<<blocks, 256, qsm, s>>>(Qp, nq, ix.centroids.as<float>(), ix.P, d, ws.probe_mat.as<u64>());
		} else {
			if (qsm > 48 * 1024) LB_CUDA(cudaFuncSetAttribute(probe_rows_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)qsm));
			probe_rows_kernel<0><<<blocks, 256, qsm, s>>>(Qp, nq, ix.centroids.as<float>(), ix.P, d, ws.probe_mat.as<u64>());
		}
		LB_CUDA(cudaGetLastError());
		launches++;
		launch_topk_merge(ws.probe_mat.as<u64>(), nq, ix.P, np, 0, nullptr, nullptr, nullptr, probe_keys, s, &launches);
	}

	// 3. per-query lookup table T1
	ws.work3.ensure((size_t)nq * M * 256 * sizeof(float));
	float *lut = ws.work3.as<float>();
	{
		size_t smem = ((size_t)256 * (ix.dsub + 1) + (size_t)LUT_QPB * ix.dsub) * sizeof(float);
		if (smem > 48 * 1024) LB_CUDA(cudaFuncSetAttribute(lut_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		lut_kernel<<<dim3(M, blocks_for(nq, LUT_QPB)), 256, smem, s>>>(Qp, nq, d, M, ix.dsub, ix.codebook.as<float>(),
		                                                               metric_ == METRIC_DOT ? 1 : 0, ix.gran == 32 ? 1 : 0, lut);
		LB_CUDA(cudaGetLastError());
		launches++;
	}

	// 4. ADC scan (a13) -> per (query, slice) candidate lists
	const int cap = warpsel_cap(kc);
	int nwarps = nq >= 64 ? 16 : 8; // batches: more warps per table copy to hide the code-load latency
	if (const char *e = getenv("LANCE_B200_ADC_WARPS")) nwarps = std::max(1, std::min(16, atoi(e))); // tuning hook
	while (nwarps > 1 && (size_t)M * 1024 + 64 + (size_t)nwarps * cap * 8 > 220 * 1024) nwarps >>= 1;
	const size_t smem = (size_t)M * 1024 + 64 + (size_t)nwarps * cap * 8;
	if (smem > 226 * 1024) throw Error("k * refine_factor too large for the ADC scan's shared memory");
	int per_sm = (int)((227 * 1024) / (smem + 1024));
	if (per_sm < 1) per_sm = 1;
	// one resident wave for a lone query (every CTA pays a 96 KB table load), two for batches
	long long want = (nq == 1 ? 1ll : 2ll) * num_sms_ * per_sm;
	if (const char *e = getenv("LANCE_B200_ADC_CTAS")) want = std::max(1, atoi(e)); // tuning hook
	const long long want_q = std::max<long long>(1, (want + nq - 1) / nq); // CTAs wanted per query
	const int PS = (int)std::min<long long>(np, want_q);                   // probe slots
	const int J = (int)std::max<long long>(1, std::min<long long>(16, want_q / PS)); // sub-slices of every list
	const int S = PS * J;
	const int tail_k = (ix.n_indexed < n_rows_) ? k : 0;
	const int fin_stride = kc + tail_k;
	ws.work4.ensure(((size_t)nq * S * kc + (size_t)nq * kc + (size_t)nq * fin_stride + (size_t)nq * std::max(tail_k, 1)) * sizeof(u64) +
	                (size_t)nq * sizeof(unsigned) + 64);
	u64 *partials = ws.work4.as<u64>();
	u64 *cand = partials + (size_t)nq * S * kc;
	u64 *fin = cand + (size_t)nq * kc;
	u64 *tailk = fin + (size_t)nq * fin_stride;
	// arrival tickets: persistent and zero between launches (the last CTA of a query hands its ticket back), so the
	// scan is one launch; zeroed here only when the buffer is new
	if (ws.adc_tickets.bytes < (size_t)nq * sizeof(unsigned)) {
		ws.adc_tickets.ensure(std::max<size_t>((size_t)nq * sizeof(unsigned), 4096));
		LB_CUDA(cudaMemsetAsync(ws.adc_tickets.p, 0, ws.adc_tickets.bytes, s));
	}
	unsigned *tickets = ws.adc_tickets.as<unsigned>();
	{
		// 4-6 in one launch: the last CTA of a query merges the slices' lists (5), refines (6, a14) and, when there is no
		// unindexed tail, writes the sorted top-k (8)
		AdcParams ap;
		ap.lut_g = lut, ap.M = M, ap.probe_keys = probe_keys, ap.nprobes_eff = np, ap.probe_stride = np, ap.qn2 = qn2;
		ap.metric = ix.metric, ap.pad_offsets = ix.pad_offsets.as<long long>(), ap.slot_rows = ix.slot_rows.as<int32_t>();
		ap.codes_il = ix.codes_il.as<uint8_t>(), ap.row_term = ix.row_term.as<float>();
		ap.allow = d_allow, ap.allow_bits = allow_bits, ap.deleted = deleted, ap.kc = kc, ap.cap = cap, ap.PS = PS;
		ap.partials = partials, ap.tickets = tickets, ap.cand = cand;
		ap.X = X_.as<float>(), ap.d = d, ap.Q = d_queries, ap.refine = refine >= 1 ? 1 : 0, ap.fin = fin, ap.fin_stride = fin_stride;
		ap.fuse_final = tail_k == 0 ? 1 : 0, ap.k = k, ap.label_base = label_base_;
		ap.out_labels = d_out_labels, ap.out_dist = d_out_dist, ap.out_counts = d_out_counts;
		int tix = timer_.begin(KC_ADC_SCAN, s);
		switch (ix.gran) {
		case 32: // the table must fit in shared memory, so M <= 192
			switch (M / 32) {
			case 1: launch_adc<32, 1>(ap, nwarps, S, nq, smem, s); break;
			case 2: launch_adc<32, 2>(ap, nwarps, S, nq, smem, s); break;
			case 3: launch_adc<32, 3>(ap, nwarps, S, nq, smem, s); break;
			case 4: launch_adc<32, 4>(ap, nwarps, S, nq, smem, s); break;
			case 5: launch_adc<32, 5>(ap, nwarps, S, nq, smem, s); break;
			case 6: launch_adc<32, 6>(ap, nwarps, S, nq, smem, s); break;
			default: throw Error("IVF_PQ: sub-vector count too large for the ADC scan's shared memory");
			}
			break;
		case 16: launch_adc<16>(ap, nwarps, S, nq, smem, s); break;
		case 4: launch_adc<4>(ap, nwarps, S, nq, smem, s); break;
		default: launch_adc<1>(ap, nwarps, S, nq, smem, s);
		}
		timer_.end(tix, s);
		launches++;
	}
	if (d_dbg_cand) {
		LB_CUDA(cudaMemcpyAsync(d_dbg_cand, cand, (size_t)nq * kc * sizeof(u64), cudaMemcpyDeviceToDevice, s));
	}
	(void)d_dbg_parts;
	(void)d_dbg_ncand;

	// 7. rows appended after the index was built: exact scan of the tail, merged (Lance's unindexed fragments)
	if (tail_k > 0) {
		FlatScanArgs a;
		a.X = X_.as<float>();
		a.n = n_rows_;
		a.row_begin = ix.n_indexed;
		a.d = d;
		a.metric = metric_;
		a.dQ = d_queries;
		a.nq = nq;
		a.k = k;
		a.allow = d_allow;
		a.allow_bits = allow_bits;
		a.deleted = deleted;
		a.out_keys = tailk;
		FlatScanPlan p = plan_flat_scan(a.X, a.metric, d, k, nq, n_rows_ - ix.n_indexed, num_sms_);
		ws.work.ensure(flat_scan_scratch_bytes(p, k));
		run_flat_scan(a, p, ws.work.p, s, &launches);
		copy_keys_strided_kernel<<<blocks_for((long long)nq * k, 256), 256, 0, s>>>(tailk, k, k, nq, fin, fin_stride, kc);
		LB_CUDA(cudaGetLastError());
		launches++;
	}

	// 8. final select (already done by the scan kernel when there is no tail)
	if (tail_k > 0)
		launch_topk_merge(fin, nq, fin_stride, k, label_base_, d_out_labels, d_out_dist, d_out_counts, nullptr, s, &launches);
	std::lock_guard<std::mutex> g(stat_mu_);
	ivfpq_batches_++;
	launches_ += launches;
}

void Store::ivfpq_search_debug(const float *query, int k, int nprobes, int refine, const uint32_t *allow,
                               int64_t allow_bits, int32_t *out_parts, int64_t *out_cand_labels, float *out_cand_adc,
                               int32_t *out_ncand, int64_t *out_labels, float *out_dist, int32_t *out_count) {
	std::shared_lock<std::shared_mutex> lk(mu_);
	LB_CUDA(cudaSetDevice(device_));
	if (!index_) throw Error("no IVF_PQ index on this handle");
	if (k <= 0) {
		*out_count = 0;
		if (out_ncand) *out_ncand = 0;
		return;
	}
	auto io = acquire_io();
	try {
		cudaStream_t s = io->stream;
		const int rf = refine < 1 ? 1 : refine;
		const int kc = k * rf;
		int np = std::min(std::max(nprobes, 1), index_->P);
		io->q.ensure((size_t)dim_ * sizeof(float));
		io->out_labels.ensure((size_t)(k + kc) * sizeof(int64_t));
		io->out_dist.ensure((size_t)(k + kc) * sizeof(float));
		io->out_counts.ensure(4 * sizeof(int32_t));
		io->work.ensure((size_t)kc * sizeof(u64));
		LB_CUDA(cudaMemcpyAsync(io->q.p, query, (size_t)dim_ * sizeof(float), cudaMemcpyHostToDevice, s));
		const uint32_t *d_allow = nullptr;
		if (allow) {
			size_t words = (size_t)((allow_bits + 31) / 32);
			io->allow.ensure(std::max<size_t>(words, 1) * sizeof(uint32_t));
			LB_CUDA(cudaMemcpyAsync(io->allow.p, allow, words * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
			d_allow = io->allow.as<uint32_t>();
		}
		Workspace *ws = scratch_for(s);
		std::lock_guard<std::mutex> wg(ws->mu);
		ivfpq_search_locked(io->q.as<float>(), 1, k, nprobes, refine, d_allow, allow_bits, io->out_labels.as<int64_t>(),
		                    io->out_dist.as<float>(), io->out_counts.as<int32_t>(), *ws, s, nullptr, io->work.as<u64>(), nullptr);
		// decode the debug lists on the device
		keys_to_debug_kernel<<<blocks_for(kc, 256), 256, 0, s>>>(io->work.as<u64>(), kc, label_base_,
		                                                         io->out_labels.as<int64_t>() + k, io->out_dist.as<float>() + k,
		                                                         io->out_counts.as<int32_t>() + 1);
		LB_CUDA(cudaGetLastError());
		std::vector<u64> pk(np);
		LB_CUDA(cudaMemcpyAsync(pk.data(), ws->work2.p, (size_t)np * sizeof(u64), cudaMemcpyDeviceToHost, s));
		std::vector<int64_t> hl(k + kc);
		std::vector<float> hd(k + kc);
		int32_t hc[2];
		LB_CUDA(cudaMemcpyAsync(hl.data(), io->out_labels.p, (size_t)(k + kc) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
		LB_CUDA(cudaMemcpyAsync(hd.data(), io->out_dist.p, (size_t)(k + kc) * sizeof(float), cudaMemcpyDeviceToHost, s));
		LB_CUDA(cudaMemcpyAsync(hc, io->out_counts.p, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
		LB_CUDA(cudaStreamSynchronize(s));
		for (int i = 0; i < np && out_parts; i++) out_parts[i] = pk[i] == KEY_INVALID ? -1 : (int32_t)key_row(pk[i]);
		*out_count = hc[0];
		if (out_ncand) *out_ncand = hc[1];
		for (int i = 0; i < hc[0]; i++) out_labels[i] = hl[i], out_dist[i] = hd[i];
		for (int i = 0; i < hc[1]; i++) {
			if (out_cand_labels) out_cand_labels[i] = hl[k + i];
			if (out_cand_adc) out_cand_adc[i] = hd[k + i];
		}
	} catch (...) {
		release_io(std::move(io));
		throw;
	}
	release_io(std::move(io));
}
