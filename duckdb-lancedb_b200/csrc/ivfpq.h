// ivfpq.h — device-resident IVF_PQ index (8-bit codes, 256 centroids per sub-quantizer).
// Search-side restatement of what the reference delegates to lancedb's IvfPq index
// (rust_lib/src/lance_manager.rs:483-515 build, :411-419 query with nprobes / refine_factor).
#pragma once
#include <vector>

#include "store.h"

namespace lb {

struct IvfPqIndex {
	virtual ~IvfPqIndex() {}
	int metric = METRIC_L2;
	int dim = 0, P = 0, M = 0, dsub = 0;
	int64_t n_indexed = 0; // rows [0, n_indexed) are covered; later rows are scanned exactly
	int gran = 16;         // interleave granule in bytes (16, 4 or 1); 32 = 16-byte granules, lane-rotated (M % 32 == 0)
	int64_t n_pad = 0;     // padded slots (every partition rounded up to 32 rows)
	int64_t max_part_rows = 0;

	DevBuf centroids;   // fp32 [P][dim]
	DevBuf codebook;    // fp32 [M][256][dsub]
	DevBuf pad_offsets; // int64 [P+1] into the padded slot space (multiples of 32)
	DevBuf slot_rows;   // int32 [n_pad] row of each slot, -1 = padding
	DevBuf codes_il;    // u8, interleaved: [n_pad/32][M/gran][32][gran]
	DevBuf row_term;    // fp32 [n_pad]: 2 * sum_m <centroid_p[m], codebook[m][code_m]> (l2/cosine), 0 for dot

	// host mirrors in the exchange layout (oracle/oracle.h oracle_ivfpq_t)
	std::vector<int64_t> part_offsets; // [P+1]
};

} // namespace lb
