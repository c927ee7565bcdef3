// TEST INFRASTRUCTURE ONLY — stands in for csrc/api.cu in the CPU emulation build (tests/emu/README.md): context,
// workspace and error plumbing on the host, the Stage-B host entry points, and "unsupported" stubs for the tensor-core
// kNN entry points (tcgen05 / TMA cannot be emulated; they are verified on hardware).
#include "../../include/sphb.h"
#include "common.cuh"

#include <string>

namespace sphb {
static Context g_ctx;
static std::string g_err;
static int64_t g_launches = 0;
Context& ctx() { return g_ctx; }
void set_error(const std::string& msg) { g_err = msg; }
int fail(int code, const std::string& msg) { g_err = msg; return code; }
void count_launch(int64_t n) { g_launches += n; }
int ensure_init() { return sphb_init(-1); }
cudaError_t ws_alloc(void** p, size_t bytes, cudaStream_t) {
    *p = bytes ? std::malloc(bytes) : nullptr;
    return (bytes && !*p) ? cudaErrorUnknown : cudaSuccess;
}
void ws_free(void* p, cudaStream_t) { std::free(p); }
void ws_trim() {}
void sims_force_slab(int on);
void walks_force_scan(int on);
int prow_graph_launch(const int64_t*, const float*, int64_t, int, float, int, uint32_t*, float*, double*, cudaStream_t);
int prow_inplace_launch(int64_t, const int64_t*, float*, int, float, int, cudaStream_t);
int prow_flat_launch(const float*, int64_t, int, double, float*, cudaStream_t);
}  // namespace sphb
using namespace sphb;

static int unsupported() { return fail(SPHB_ERR_UNSUPPORTED, "emulation build: the tensor-core kNN path runs on hardware only"); }

extern "C" {
SPHB_API int sphb_init(int) {
    g_ctx.device = 0;
    g_ctx.num_sms = 2;
    g_ctx.ready = true;
    return SPHB_OK;
}
SPHB_API void sphb_shutdown(void) {}
SPHB_API const char* sphb_last_error(void) { return g_err.c_str(); }
SPHB_API int sphb_version(void) { return -1; }  // marks the emulation build
SPHB_API int64_t sphb_launch_count(void) { return g_launches; }
SPHB_API void sphb_reset_launch_count(void) { g_launches = 0; }
SPHB_API void* sphb_stream(void) { return nullptr; }
SPHB_API void sphb_debug_force_knn_path(int) {}
SPHB_API void sphb_debug_force_sims_slab(int on) { sims_force_slab(on); }
SPHB_API void sphb_debug_force_walk_scan(int on) { walks_force_scan(on); }
SPHB_API int sphb_knn_exact(const float*, int64_t, int, int, int, int, int64_t*, float*) { return unsupported(); }
SPHB_API int sphb_knn_exact_slab(const float*, int64_t, int, int, int, int, int64_t, int64_t, int64_t*, float*) { return unsupported(); }
SPHB_API int sphb_knn_exact_dev(const float*, int64_t, int, int, int, int, int64_t, int64_t, int64_t*, float*, void*) { return unsupported(); }
SPHB_API int sphb_knn_last_stats(int64_t*) { return unsupported(); }
SPHB_API int sphb_knn_last_timing(float*) { return unsupported(); }
SPHB_API int sphb_knn_graph_prow(const float*, int64_t, int, int, int, int, float, int64_t, int64_t, int64_t*, float*, uint32_t*, float*) { return unsupported(); }
SPHB_API int sphb_prow_gaussian_graph_dev(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity, uint32_t* cols, float* vals,
                                          double* sigma, void*) {
    return prow_graph_launch(ids, dist, n, k, perplexity, 0, cols, vals, sigma, nullptr);
}
SPHB_API int sphb_prow_gaussian_graph(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity, uint32_t* cols, float* vals, double* sigma) {
    if (!ids || !dist || n < 0 || k < 2) return fail(SPHB_ERR_INVALID, "prow_graph: bad arguments");
    return n ? prow_graph_launch(ids, dist, n, k, perplexity, 0, cols, vals, sigma, nullptr) : SPHB_OK;
}
SPHB_API int sphb_prow_linear_graph(const int64_t* ids, const float* dist, int64_t n, int k, uint32_t* cols, float* vals) {
    if (!ids || !dist || n < 0 || k < 2) return fail(SPHB_ERR_INVALID, "prow_graph: bad arguments");
    return n ? prow_graph_launch(ids, dist, n, k, 0.f, 1, cols, vals, nullptr, nullptr) : SPHB_OK;
}
SPHB_API int sphb_prow_gaussian_inplace(int64_t n, const int64_t* rowptr, float* val, float perplexity, int ignore) {
    if (n < 0 || !rowptr) return fail(SPHB_ERR_INVALID, "prow_inplace: bad arguments");
    int64_t max_nnz = 0;
    for (int64_t i = 0; i < n; ++i) max_nnz = std::max(max_nnz, rowptr[i + 1] - rowptr[i]);
    if (n == 0 || rowptr[n] == 0) return SPHB_OK;
    return prow_inplace_launch(n, rowptr, val, static_cast<int>(max_nnz), perplexity, ignore, nullptr);
}
SPHB_API int sphb_prow_gaussian_flat(const float* dist, int64_t n, int k, double perplexity, float* temp) {
    if (!dist || !temp || n < 0 || k < 1) return fail(SPHB_ERR_INVALID, "prow_flat: bad arguments");
    return n ? prow_flat_launch(dist, n, k, perplexity, temp, nullptr) : SPHB_OK;
}
}
