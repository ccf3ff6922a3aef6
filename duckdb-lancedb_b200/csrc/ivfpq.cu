// placeholder until the IVF_PQ kernels land
#include "ivfpq.h"
namespace lb {
void Store::create_index(int, int) { throw Error("IVF_PQ not built"); }
void Store::set_index(int, int, int64_t, const float *, const float *, const int64_t *, const int64_t *, const uint8_t *) { throw Error("IVF_PQ not built"); }
void Store::get_index(lance_b200_ivfpq_info *, float *, float *, int64_t *, int64_t *, uint8_t *) { throw Error("IVF_PQ not built"); }
void Store::drop_index() { std::unique_lock<std::shared_mutex> lk(mu_); index_.reset(); }
void Store::ivfpq_search_debug(const float *, int, int, int, const uint32_t *, int64_t, int32_t *, int64_t *, float *, int32_t *, int64_t *, float *, int32_t *) { throw Error("IVF_PQ not built"); }
void Store::ivfpq_search_locked(const float *, int, int, int, int, const uint32_t *, int64_t, int64_t *, float *, int32_t *, Workspace &, cudaStream_t, int32_t *, u64 *, int32_t *) { throw Error("IVF_PQ not built"); }
} // namespace lb
