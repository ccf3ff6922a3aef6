// ref_shim.cpp — compiles the REFERENCE's own distance kernels (DuckDB core_functions, the path
// that actually executes `ORDER BY array_distance(...) LIMIT k` today) where they lie
// under /root/reference, into oracle/_ref/libref_array_kernels.so.  No reference source is copied:
// the header is #included from the read-only tree at build time (see oracle/Makefile).
// TEST INFRASTRUCTURE ONLY — used to validate oracle.c's restatement.
#include "core_functions/array_kernels.hpp" // /root/reference/duckdb/extension/core_functions/include
#include <cstdint>

extern "C" {
// array_distance(l, r)            array_kernels.hpp:97-105 (sqrt of the squared fold :76-95)
float ref_array_distance(const float *l, const float *r, int64_t n) {
	return duckdb::DistanceOp::Operation<float>(l, r, static_cast<duckdb::idx_t>(n));
}
float ref_array_distance_squared(const float *l, const float *r, int64_t n) {
	return duckdb::DistanceSquaredOp::Operation<float>(l, r, static_cast<duckdb::idx_t>(n));
}
// array_cosine_distance(l, r)     array_kernels.hpp:66-74
float ref_array_cosine_distance(const float *l, const float *r, int64_t n) {
	return duckdb::CosineDistanceOp::Operation<float>(l, r, static_cast<duckdb::idx_t>(n));
}
// array_inner_product(l, r)       array_kernels.hpp:11-30
float ref_array_inner_product(const float *l, const float *r, int64_t n) {
	return duckdb::InnerProductOp::Operation<float>(l, r, static_cast<duckdb::idx_t>(n));
}
}
