// Stage B6 / C / D entry points (being filled in; see DESIGN.md for status).
#include "../../include/sphb.h"
#include "common.cuh"

using namespace sphb;

extern "C" {

SPHB_API int sphb_symmetrize(int64_t, const int64_t*, const uint32_t*, const float*, int, int64_t*, uint32_t*, float*, int64_t, int64_t*) {
    return fail(SPHB_ERR_UNSUPPORTED, "sphb_symmetrize: not implemented yet");
}
SPHB_API int sphb_random_walks(int64_t, const int64_t*, const uint32_t*, const float*, uint64_t, uint64_t, int, float, uint64_t, uint64_t, int, int,
                               int64_t*, uint32_t*, float*, int64_t, int64_t*) {
    return fail(SPHB_ERR_UNSUPPORTED, "sphb_random_walks: not implemented yet");
}
SPHB_API int sphb_rw_merge(int64_t, const int64_t*, const uint32_t*, const float*, const uint64_t*, int64_t, int, int, int64_t*, uint32_t*, float*,
                           int64_t, int64_t*) {
    return fail(SPHB_ERR_UNSUPPORTED, "sphb_rw_merge: not implemented yet");
}
SPHB_API int sphb_create_similarities(int64_t, const int64_t*, const uint32_t*, const float*, int64_t, float, int, const int64_t*, uint32_t*, float*,
                                      int32_t*) {
    return fail(SPHB_ERR_UNSUPPORTED, "sphb_create_similarities: not implemented yet");
}
SPHB_API int sphb_pair_sims(int64_t, const int64_t*, const uint32_t*, const float*, const int64_t*, const int64_t*, int64_t, int, float*) {
    return fail(SPHB_ERR_UNSUPPORTED, "sphb_pair_sims: not implemented yet");
}
}
