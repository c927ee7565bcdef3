// C-ABI entry points (include/sphb.h): context, Stage A and Stage B wrappers.
#include "../../include/sphb.h"
#include "common.cuh"

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>
#include <algorithm>

namespace sphb {

// One context per HOST THREAD: a thread binds a device with sphb_init (or lazily to the current device) and owns a stream; the
// multi-device entry point below drives every GPU from its own worker thread, each with its own context.
static thread_local Context g_ctx;
static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

Context& ctx() { return g_ctx; }
void set_error(const std::string& msg) { g_err = msg; }
int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
void count_launch(int64_t n) { g_launches += n; }

int ensure_init() {
    if (g_ctx.ready) return 0;
    return sphb_init(-1);
}

// ---- workspace cache (see common.cuh) -----------------------------------------------------------
namespace {
struct WsBlock {
    void* p;
    size_t bytes;
    cudaStream_t stream;
};
std::mutex g_ws_mu;
std::vector<WsBlock> g_ws_free;                     // idle blocks
std::vector<WsBlock> g_ws_live;                     // blocks handed out (to find their size at release)
size_t g_ws_idle_bytes = 0;
constexpr size_t WS_IDLE_CAP = 64ull << 30;
}  // namespace

cudaError_t ws_alloc(void** p, size_t bytes, cudaStream_t stream) {
    if (bytes == 0) { *p = nullptr; return cudaSuccess; }
    {
        std::lock_guard<std::mutex> lk(g_ws_mu);
        int best = -1;
        for (int i = 0; i < static_cast<int>(g_ws_free.size()); ++i) {
            const WsBlock& b = g_ws_free[i];
            if (b.stream != stream || b.bytes < bytes || b.bytes > bytes + bytes / 4 + (1u << 20)) continue;
            if (best < 0 || b.bytes < g_ws_free[best].bytes) best = i;
        }
        if (best >= 0) {
            WsBlock b = g_ws_free[best];
            g_ws_free.erase(g_ws_free.begin() + best);
            g_ws_idle_bytes -= b.bytes;
            g_ws_live.push_back(b);
            *p = b.p;
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMallocAsync(p, bytes, stream);
    if (e == cudaErrorMemoryAllocation) {
        // the idle blocks of the cache (any stream, any size class) may be what is missing: give them back and try once more
        cudaGetLastError();
        ws_trim();
        cudaStreamSynchronize(stream);
        int dev = 0;
        cudaMemPool_t pool;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
        e = cudaMallocAsync(p, bytes, stream);
    }
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(g_ws_mu);
    g_ws_live.push_back({*p, bytes, stream});
    return cudaSuccess;
}

void ws_free(void* p, cudaStream_t stream) {
    if (!p) return;
    std::lock_guard<std::mutex> lk(g_ws_mu);
    for (size_t i = 0; i < g_ws_live.size(); ++i) {
        if (g_ws_live[i].p != p) continue;
        WsBlock b = g_ws_live[i];
        g_ws_live.erase(g_ws_live.begin() + static_cast<long>(i));
        if (b.stream == stream && g_ws_idle_bytes + b.bytes <= WS_IDLE_CAP) {
            g_ws_free.push_back(b);
            g_ws_idle_bytes += b.bytes;
        } else {
            cudaFreeAsync(p, stream);
        }
        return;
    }
    cudaFreeAsync(p, stream);  // not ours (should not happen)
}

void ws_trim() {
    std::lock_guard<std::mutex> lk(g_ws_mu);
    for (const WsBlock& b : g_ws_free) cudaFreeAsync(b.p, b.stream);
    g_ws_free.clear();
    g_ws_idle_bytes = 0;
}
// only the idle blocks of one stream (a thread shutting down its own context leaves the other threads' caches alone)
static void ws_trim_stream(cudaStream_t stream) {
    std::lock_guard<std::mutex> lk(g_ws_mu);
    size_t w = 0;
    for (size_t i = 0; i < g_ws_free.size(); ++i) {
        if (g_ws_free[i].stream == stream) {
            cudaFreeAsync(g_ws_free[i].p, stream);
            g_ws_idle_bytes -= g_ws_free[i].bytes;
        } else {
            g_ws_free[w++] = g_ws_free[i];
        }
    }
    g_ws_free.resize(w);
}

// kernels / launchers implemented in the other translation units
int knn_simt_launch(const float*, int64_t, int, int, int, int, int64_t, int64_t, const int64_t*, const int*, int64_t, int64_t*, float*, cudaStream_t);
int knn_tc_launch(const float*, int64_t, int, int, int, int, int64_t, int64_t, int64_t*, float*, cudaStream_t, int);
bool knn_tc_supported(int64_t n, int d, int k);
void knn_set_stats(const int64_t* s);
void knn_get_stats(int64_t* s);
void knn_get_timing(float* ms4);
void knn_clear_overflow_rows();
const std::vector<int64_t>& knn_overflow_rows(int which);
int prow_graph_launch(const int64_t*, const float*, int64_t, int, float, int, uint32_t*, float*, double*, cudaStream_t);
int prow_inplace_launch(int64_t, const int64_t*, float*, int, float, int, cudaStream_t);
int prow_flat_launch(const float*, int64_t, int, double, float*, cudaStream_t);
int component_centers_dev(const float*, int64_t, int, const int64_t*, int64_t, float*, cudaStream_t);
int bridge_components_dev(const float*, int64_t, int, const int64_t*, int64_t, const int64_t*, int64_t, int64_t*, int64_t*, float*, cudaStream_t);
int prow_exponential_graph_launch(const int64_t*, const float*, int64_t, int, int, uint32_t*, float*, cudaStream_t);
int prow_exponential_inplace_launch(int64_t, const int64_t*, const uint32_t*, float*, int, int, cudaStream_t);
void sims_force_slab(int on);
void walks_force_scan(int on);

// testing hook: 0 auto, 1 SIMT, 2 tensor core (auto), 3 tc streaming 3-pass, 4 tc resident MH=1 3-pass, 5 tc 3-pass (auto flavour),
// 6 tc 1-pass (resident when d allows), 7 tc 1-pass streaming
static int g_force_path = 0;

// COSINE preparation (NearestNeighbors::initFaissData, sph/NearestNeighbors.cpp:208-222: copy + faiss::fvec_renorm_L2): one warp per
// row; the squared norm is the fp32 sum of x[c] * x[c] taken in column order (the published faiss scalar loop) — the lanes hold
// the row, the running sum is formed in order through shuffles.
__global__ void renorm_l2_kernel(const float* __restrict__ x, int64_t n, int d, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const float* xr = x + r * d;
        float nr = 0.f;
        for (int c0 = 0; c0 < d; c0 += 32) {
            const float v = c0 + lane < d ? xr[c0 + lane] : 0.f;
            const float sq = __fmul_rn(v, v);
            const int m = min(32, d - c0);
            for (int t = 0; t < m; ++t) nr = __fadd_rn(nr, __shfl_sync(0xffffffffu, sq, t));
        }
        const float inv = nr > 0.f ? 1.0f / sqrtf(nr) : 1.f;
        for (int c = lane; c < d; c += 32) out[r * d + c] = nr > 0.f ? __fmul_rn(xr[c], inv) : xr[c];
    }
}

int renorm_l2_dev(const float* d_x, int64_t n, int d, float* d_out, cudaStream_t s) {
    const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((n * 32 + 255) / 256, static_cast<int64_t>(ctx().num_sms) * 16));
    renorm_l2_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(d_x, n, d, d_out);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace sphb

using namespace sphb;

namespace {
// Persistent worker threads of the multi-device entry point: worker g keeps its context (device, stream, cached workspace
// blocks, enabled peer access) between calls, so a steady-state call makes no cudaMallocAsync call and creates no stream.
struct MultiWorker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::function<void()> job;
    bool has_job = false, done = true;
    MultiWorker() {
        th = std::thread([this] {
            for (;;) {
                std::function<void()> j;
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return has_job; });
                    j = std::move(job);
                }
                j();
                {
                    std::lock_guard<std::mutex> lk(mu);
                    has_job = false;
                    done = true;
                }
                cv.notify_all();
            }
        });
        th.detach();  // lives until the process exits
    }
    void post(std::function<void()> j) {
        {
            std::lock_guard<std::mutex> lk(mu);
            job = std::move(j);
            has_job = true;
            done = false;
        }
        cv.notify_all();
    }
    void wait() {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return done; });
    }
};
std::mutex g_multi_call_mu;             // one multi-device call at a time
std::vector<MultiWorker*> g_multi_pool;  // never destroyed: the threads are detached

}  // namespace

extern "C" {

SPHB_API int sphb_init(int device) {
    if (g_ctx.ready && (device < 0 || device == g_ctx.device)) return SPHB_OK;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(SPHB_ERR_NO_DEVICE, "no CUDA device available (this back-end has no CPU fallback)");
    }
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    }
    if (device >= count) return fail(SPHB_ERR_INVALID, "device index out of range");
    SPHB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    SPHB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(SPHB_ERR_UNSUPPORTED, std::string("sm_100a build, found ") + prop.name);
    if (g_ctx.stream) {
        sphb::ws_trim_stream(g_ctx.stream);  // cached blocks are tied to the stream they were released on
        cudaStreamSynchronize(g_ctx.stream);
        cudaStreamDestroy(g_ctx.stream);
    }
    SPHB_CUDA(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
    // keep freed stream-ordered allocations cached in the pool between calls
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        uint64_t thresh = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh);
    }
    g_ctx.device = device;
    g_ctx.num_sms = prop.multiProcessorCount;
    g_ctx.ready = true;
    return SPHB_OK;
}

static void shutdown_context(bool trim_pool) {
    if (!g_ctx.ready) return;
    cudaStreamSynchronize(g_ctx.stream);
    sphb::ws_trim_stream(g_ctx.stream);
    cudaStreamSynchronize(g_ctx.stream);
    cudaStreamDestroy(g_ctx.stream);
    g_ctx.stream = nullptr;
    cudaMemPool_t pool;
    if (trim_pool && cudaDeviceGetDefaultMemPool(&pool, g_ctx.device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
    g_ctx.ready = false;
}

SPHB_API void sphb_shutdown(void) {
    shutdown_context(true);
    // the worker threads of sphb_knn_graph_prow_multi hold contexts of their own
    std::vector<MultiWorker*> pool;
    {
        std::lock_guard<std::mutex> call_lock(g_multi_call_mu);
        pool = g_multi_pool;
        for (MultiWorker* w : pool) w->post([] { shutdown_context(true); });
        for (MultiWorker* w : pool) w->wait();
    }
}

SPHB_API const char* sphb_last_error(void) { return g_err.c_str(); }
SPHB_API int sphb_version(void) { return 100; }
SPHB_API int64_t sphb_launch_count(void) { return g_launches.load(); }
SPHB_API void sphb_reset_launch_count(void) { g_launches = 0; }
SPHB_API void* sphb_stream(void) { return g_ctx.ready ? static_cast<void*>(g_ctx.stream) : nullptr; }

// testing hooks (not in the public header): 0 auto, 1 force SIMT, 2 force tensor-core path; D1 accumulator home
SPHB_API void sphb_debug_force_knn_path(int path) { g_force_path = path; }
SPHB_API void sphb_debug_force_sims_slab(int on) { sphb::sims_force_slab(on); }
SPHB_API void sphb_debug_force_walk_scan(int on) { sphb::walks_force_scan(on); }

// ---------------------------------------------------------------- Stage A
SPHB_API int sphb_knn_exact_dev(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1,
                                int64_t* d_ids, float* d_dist, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n <= 0 || d <= 0 || k <= 0 || q0 < 0 || q1 > n || q0 > q1) return fail(SPHB_ERR_INVALID, "knn: bad sizes");
    if (metric != SPHB_METRIC_L2 && metric != SPHB_METRIC_INNER_PRODUCT && metric != SPHB_METRIC_COSINE) return fail(SPHB_ERR_INVALID, "knn: unknown metric");
    if (n >= (1ll << 32) - 1) return fail(SPHB_ERR_UNSUPPORTED, "knn: n must fit in uint32");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (metric == SPHB_METRIC_COSINE) {
        // KnnMetric::COSINE (sph/NearestNeighbors.cpp:81-96, 208-222): rows L2-normalised into a device copy, then inner product
        DevBuf<float> xn;
        SPHB_CUDA(xn.alloc(static_cast<size_t>(n) * d, s));
        if (int rc = renorm_l2_dev(d_x, n, d, xn.p, s)) return rc;
        const int rc = sphb_knn_exact_dev(xn.p, n, d, k, SPHB_METRIC_INNER_PRODUCT, 0, q0, q1, d_ids, d_dist, stream);
        SPHB_CUDA(cudaStreamSynchronize(s));  // the copy is released on return
        return rc;
    }
    bool use_tc = knn_tc_supported(n, d, k);
    if (g_force_path == 1) use_tc = false;
    if (g_force_path >= 2) {
        if (!use_tc) return fail(SPHB_ERR_UNSUPPORTED, "knn: tensor-core path unsupported for this shape");
    }
    if (use_tc) return knn_tc_launch(d_x, n, d, k, metric, l2_squared, q0, q1, d_ids, d_dist, s, g_force_path >= 3 ? g_force_path - 2 : 0);
    const int64_t st[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    knn_set_stats(st);
    knn_clear_overflow_rows();
    return knn_simt_launch(d_x, n, d, k, metric, l2_squared, q0, q1, nullptr, nullptr, 0, d_ids, d_dist, s);
}

SPHB_API int sphb_knn_exact_slab(const float* x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1,
                                 int64_t* ids, float* dist) {
    if (int rc = ensure_init()) return rc;
    if (!x || !ids || !dist) return fail(SPHB_ERR_INVALID, "knn: null pointer");
    if (n <= 0 || d <= 0 || k <= 0 || q0 < 0 || q1 > n || q0 > q1) return fail(SPHB_ERR_INVALID, "knn: bad sizes");
    cudaStream_t s = g_ctx.stream;
    const int64_t nq = q1 - q0;
    DevBuf<float> dx, ddist;
    DevBuf<int64_t> dids;
    SPHB_CUDA(dx.alloc(static_cast<size_t>(n) * d, s));
    SPHB_CUDA(dids.alloc(static_cast<size_t>(nq) * k, s));
    SPHB_CUDA(ddist.alloc(static_cast<size_t>(nq) * k, s));
    SPHB_CUDA(cudaMemcpyAsync(dx.p, x, static_cast<size_t>(n) * d * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = sphb_knn_exact_dev(dx.p, n, d, k, metric, l2_squared, q0, q1, dids.p, ddist.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(ids, dids.p, static_cast<size_t>(nq) * k * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(dist, ddist.p, static_cast<size_t>(nq) * k * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_knn_exact(const float* x, int64_t n, int d, int k, int metric, int l2_squared, int64_t* ids, float* dist) {
    return sphb_knn_exact_slab(x, n, d, k, metric, l2_squared, 0, n, ids, dist);
}

SPHB_API int sphb_renorm_l2(const float* x, int64_t n, int d, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!x || !out || n < 0 || d <= 0) return fail(SPHB_ERR_INVALID, "renorm_l2: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    DevBuf<float> dx, dout;
    const size_t cnt = static_cast<size_t>(n) * d;
    SPHB_CUDA(dx.alloc(cnt, s));
    SPHB_CUDA(dout.alloc(cnt, s));
    SPHB_CUDA(cudaMemcpyAsync(dx.p, x, cnt * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = renorm_l2_dev(dx.p, n, d, dout.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out, dout.p, cnt * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_knn_last_stats(int64_t* out8) {
    if (!out8) return fail(SPHB_ERR_INVALID, "null pointer");
    knn_get_stats(out8);
    return SPHB_OK;
}

SPHB_API int sphb_knn_last_overflow_rows(int which, int64_t* out, int64_t cap, int64_t* count) {
    if (!count || (cap > 0 && !out) || which < 0 || which > 1) return fail(SPHB_ERR_INVALID, "knn_last_overflow_rows: bad arguments");
    const std::vector<int64_t>& v = knn_overflow_rows(which);
    *count = static_cast<int64_t>(v.size());
    for (int64_t i = 0; i < cap && i < *count; ++i) out[i] = v[static_cast<size_t>(i)];
    return SPHB_OK;
}

SPHB_API int sphb_knn_last_timing(float* ms4) {
    if (!ms4) return fail(SPHB_ERR_INVALID, "null pointer");
    knn_get_timing(ms4);
    return SPHB_OK;
}

// Stage A + level-0 P rows in one call: x goes up once, the graph stays in HBM between the two
// stages, (ids, dist, cols, vals) come back once.
SPHB_API int sphb_knn_graph_prow(const float* x, int64_t n, int d, int k, int metric, int l2_squared, float perplexity, int64_t q0, int64_t q1,
                                 int64_t* ids, float* dist, uint32_t* cols, float* vals) {
    if (int rc = ensure_init()) return rc;
    if (!x || !ids || !dist || !cols || !vals) return fail(SPHB_ERR_INVALID, "knn_graph_prow: null pointer");
    if (n <= 0 || d <= 0 || k < 2 || q0 < 0 || q1 > n || q0 > q1) return fail(SPHB_ERR_INVALID, "knn_graph_prow: bad sizes");
    cudaStream_t s = g_ctx.stream;
    const size_t nq = static_cast<size_t>(q1 - q0);
    DevBuf<float> dx, ddist, dvals;
    DevBuf<int64_t> dids;
    DevBuf<uint32_t> dcols;
    SPHB_CUDA(dx.alloc(static_cast<size_t>(n) * d, s));
    SPHB_CUDA(dids.alloc(nq * k, s));
    SPHB_CUDA(ddist.alloc(nq * k, s));
    SPHB_CUDA(dcols.alloc(nq * (k - 1), s));
    SPHB_CUDA(dvals.alloc(nq * (k - 1), s));
    SPHB_CUDA(cudaMemcpyAsync(dx.p, x, static_cast<size_t>(n) * d * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = sphb_knn_exact_dev(dx.p, n, d, k, metric, l2_squared, q0, q1, dids.p, ddist.p, s)) return rc;
    // The graph goes back to the host on a second stream WHILE the P-row kernel runs (it only reads ids / dist): the kernel's
    // 36 ms at C4 disappear behind the 4.6 GB copy.
    cudaStream_t cs = nullptr;
    cudaEvent_t graph_done = nullptr, copies_done = nullptr;
    const bool overlap = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) == cudaSuccess &&
                         cudaEventCreateWithFlags(&graph_done, cudaEventDisableTiming) == cudaSuccess &&
                         cudaEventCreateWithFlags(&copies_done, cudaEventDisableTiming) == cudaSuccess;
    auto cleanup = [&]() {
        if (graph_done) cudaEventDestroy(graph_done);
        if (copies_done) cudaEventDestroy(copies_done);
        if (cs) cudaStreamDestroy(cs);
    };
    cudaStream_t gs = overlap ? cs : s;  // stream of the graph copies
    cudaError_t e = cudaSuccess;
    if (overlap) {
        e = cudaEventRecord(graph_done, s);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(cs, graph_done, 0);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(ids, dids.p, nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, gs);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dist, ddist.p, nq * k * sizeof(float), cudaMemcpyDeviceToHost, gs);
    if (e != cudaSuccess) {
        cudaStreamSynchronize(gs);
        cleanup();
        return fail(SPHB_ERR_CUDA, std::string("knn_graph_prow: ") + cudaGetErrorString(e));
    }
    // the rows of a slab are points q0.., but prow only uses ids for the column labels, so the slab graph can be fed as is
    int rc = prow_graph_launch(dids.p, ddist.p, static_cast<int64_t>(nq), k, perplexity, 0, dcols.p, dvals.p, nullptr, s);
    if (rc == SPHB_OK) {
        e = cudaMemcpyAsync(cols, dcols.p, nq * (k - 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(vals, dvals.p, nq * (k - 1) * sizeof(float), cudaMemcpyDeviceToHost, s);
    }
    if (overlap) {  // the graph buffers go back to the workspace cache of s only after the other stream is done with them
        cudaEventRecord(copies_done, cs);
        cudaStreamWaitEvent(s, copies_done, 0);
    }
    const cudaError_t e2 = cudaStreamSynchronize(s);
    cleanup();
    if (rc != SPHB_OK) return rc;
    if (e != cudaSuccess || e2 != cudaSuccess) return fail(SPHB_ERR_CUDA, std::string("knn_graph_prow: ") + cudaGetErrorString(e != cudaSuccess ? e : e2));
    return SPHB_OK;
}

// ---------------------------------------------------------------- Stage B
SPHB_API int sphb_prow_gaussian_graph_dev(const int64_t* d_ids, const float* d_dist, int64_t n, int k, float perplexity, uint32_t* d_cols,
                                          float* d_vals, double* d_sigma, void* stream) {
    if (int rc = ensure_init()) return rc;
    return prow_graph_launch(d_ids, d_dist, n, k, perplexity, 0, d_cols, d_vals, d_sigma, static_cast<cudaStream_t>(stream));
}

static int prow_graph_host(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity, int mode, uint32_t* cols, float* vals,
                           double* sigma) {
    if (int rc = ensure_init()) return rc;
    if (!ids || !dist || !cols || !vals) return fail(SPHB_ERR_INVALID, "prow: null pointer");
    if (n < 0 || k < 2) return fail(SPHB_ERR_INVALID, "prow: bad sizes");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    const size_t nk = static_cast<size_t>(n) * k, ne = static_cast<size_t>(n) * (k - 1);
    DevBuf<int64_t> dids;
    DevBuf<float> ddist, dvals;
    DevBuf<uint32_t> dcols;
    DevBuf<double> dsig;
    SPHB_CUDA(dids.alloc(nk, s));
    SPHB_CUDA(ddist.alloc(nk, s));
    SPHB_CUDA(dcols.alloc(ne, s));
    SPHB_CUDA(dvals.alloc(ne, s));
    if (sigma) SPHB_CUDA(dsig.alloc(n, s));
    SPHB_CUDA(cudaMemcpyAsync(dids.p, ids, nk * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(ddist.p, dist, nk * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = prow_graph_launch(dids.p, ddist.p, n, k, perplexity, mode, dcols.p, dvals.p, sigma ? dsig.p : nullptr, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(cols, dcols.p, ne * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(vals, dvals.p, ne * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (sigma) SPHB_CUDA(cudaMemcpyAsync(sigma, dsig.p, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_prow_gaussian_graph(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity, uint32_t* cols, float* vals,
                                      double* sigma) {
    return prow_graph_host(ids, dist, n, k, perplexity, 0, cols, vals, sigma);
}

SPHB_API int sphb_prow_linear_graph(const int64_t* ids, const float* dist, int64_t n, int k, uint32_t* cols, float* vals) {
    return prow_graph_host(ids, dist, n, k, 0.f, 1, cols, vals, nullptr);
}

SPHB_API int sphb_prow_gaussian_inplace(int64_t n, const int64_t* rowptr, float* val, float perplexity, int ignore) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || !rowptr || (!val && n > 0 && rowptr[n] > 0)) return fail(SPHB_ERR_INVALID, "prow_inplace: bad arguments");
    if (n == 0 || rowptr[n] == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    const int64_t nnz = rowptr[n];
    int64_t max_nnz = 0;
    for (int64_t i = 0; i < n; ++i) max_nnz = std::max(max_nnz, rowptr[i + 1] - rowptr[i]);
    if (max_nnz > 1024) return fail(SPHB_ERR_UNSUPPORTED, "prow_inplace: rows with more than 1024 entries unsupported");
    DevBuf<int64_t> drp;
    DevBuf<float> dval;
    SPHB_CUDA(drp.alloc(n + 1, s));
    SPHB_CUDA(dval.alloc(nnz, s));
    SPHB_CUDA(cudaMemcpyAsync(drp.p, rowptr, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dval.p, val, nnz * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = prow_inplace_launch(n, drp.p, dval.p, static_cast<int>(max_nnz), perplexity, ignore, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(val, dval.p, nnz * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

// §8(f)-2: NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861), device parts
SPHB_API int sphb_component_centers_dev(const float* d_x, int64_t n, int d, const int64_t* d_labels, int64_t num_components, float* d_centers,
                                        void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || d <= 0 || num_components < 0 || (n > 0 && (!d_x || !d_labels)) || (num_components > 0 && !d_centers))
        return fail(SPHB_ERR_INVALID, "component_centers_dev: bad arguments");
    if (num_components > 0x7fffffff) return fail(SPHB_ERR_UNSUPPORTED, "component_centers: more than 2^31 - 1 components");
    return component_centers_dev(d_x, n, d, d_labels, num_components, d_centers, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_bridge_components_dev(const float* d_x, int64_t n, int d, int metric, const int64_t* d_labels, int64_t num_components,
                                        const int64_t* edges, int64_t num_edges, int64_t* d_ids_a, int64_t* d_ids_b, float* d_dist, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || d <= 0 || num_components < 0 || num_edges < 0 || (n > 0 && (!d_x || !d_labels)) ||
        (num_edges > 0 && (!edges || !d_ids_a || !d_ids_b || !d_dist)))
        return fail(SPHB_ERR_INVALID, "bridge_components_dev: bad arguments");
    if (metric != SPHB_METRIC_L2) return fail(SPHB_ERR_UNSUPPORTED, "bridge_components: only SPHB_METRIC_L2 (squared distances, as faiss returns them)");
    return bridge_components_dev(d_x, n, d, d_labels, num_components, edges, num_edges, d_ids_a, d_ids_b, d_dist, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_component_centers(const float* x, int64_t n, int d, const int64_t* labels, int64_t num_components, float* centers) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || d <= 0 || num_components < 0 || (n > 0 && (!x || !labels)) || (num_components > 0 && !centers))
        return fail(SPHB_ERR_INVALID, "component_centers: bad arguments");
    if (num_components == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    DevBuf<float> dx, dc;
    DevBuf<int64_t> dl;
    SPHB_CUDA(dx.alloc(static_cast<size_t>(n) * d, s));
    SPHB_CUDA(dl.alloc(n, s));
    SPHB_CUDA(dc.alloc(static_cast<size_t>(num_components) * d, s));
    SPHB_CUDA(cudaMemcpyAsync(dx.p, x, static_cast<size_t>(n) * d * sizeof(float), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dl.p, labels, n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (int rc = sphb_component_centers_dev(dx.p, n, d, dl.p, num_components, dc.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(centers, dc.p, static_cast<size_t>(num_components) * d * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_bridge_components(const float* x, int64_t n, int d, int metric, const int64_t* labels, int64_t num_components, const int64_t* edges,
                                    int64_t num_edges, int64_t* ids_a, int64_t* ids_b, float* dist) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || d <= 0 || num_edges < 0 || (n > 0 && (!x || !labels)) || (num_edges > 0 && (!edges || !ids_a || !ids_b || !dist)))
        return fail(SPHB_ERR_INVALID, "bridge_components: bad arguments");
    if (num_edges == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    DevBuf<float> dx, dd;
    DevBuf<int64_t> dl, da, db;
    SPHB_CUDA(dx.alloc(static_cast<size_t>(n) * d, s));
    SPHB_CUDA(dl.alloc(n, s));
    SPHB_CUDA(da.alloc(num_edges, s));
    SPHB_CUDA(db.alloc(num_edges, s));
    SPHB_CUDA(dd.alloc(num_edges, s));
    SPHB_CUDA(cudaMemcpyAsync(dx.p, x, static_cast<size_t>(n) * d * sizeof(float), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dl.p, labels, n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (int rc = sphb_bridge_components_dev(dx.p, n, d, metric, dl.p, num_components, edges, num_edges, da.p, db.p, dd.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(ids_a, da.p, num_edges * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(ids_b, db.p, num_edges * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(dist, dd.p, num_edges * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

// §8(f)-3: utils::computeExponentialDistributions(graph, rows) (sph/utils/GraphNormalization.cpp:508-593)
SPHB_API int sphb_prow_exponential_graph_dev(const int64_t* d_ids, const float* d_dist, int64_t n, int k, int sort_cols, uint32_t* d_cols,
                                             float* d_vals, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || k < 2 || (n > 0 && (!d_ids || !d_dist || !d_cols || !d_vals))) return fail(SPHB_ERR_INVALID, "prow_exponential_graph_dev: bad arguments");
    return prow_exponential_graph_launch(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_prow_exponential_graph(const int64_t* ids, const float* dist, int64_t n, int k, int sort_cols, uint32_t* cols, float* vals) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || k < 2 || (n > 0 && (!ids || !dist || !cols || !vals))) return fail(SPHB_ERR_INVALID, "prow_exponential_graph: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    const size_t nk = static_cast<size_t>(n) * k, ne = static_cast<size_t>(n) * (k - 1);
    DevBuf<int64_t> dids;
    DevBuf<float> ddist, dvals;
    DevBuf<uint32_t> dcols;
    SPHB_CUDA(dids.alloc(nk, s));
    SPHB_CUDA(ddist.alloc(nk, s));
    SPHB_CUDA(dcols.alloc(ne, s));
    SPHB_CUDA(dvals.alloc(ne, s));
    SPHB_CUDA(cudaMemcpyAsync(dids.p, ids, nk * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(ddist.p, dist, nk * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = prow_exponential_graph_launch(dids.p, ddist.p, n, k, sort_cols, dcols.p, dvals.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(cols, dcols.p, ne * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(vals, dvals.p, ne * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

// §8(f)-3: utils::computeExponentialDistributions(rows, ignore) in place (sph/utils/GraphNormalization.cpp:413-506)
SPHB_API int sphb_prow_exponential_inplace_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, float* d_val, int max_row_nnz, int ignore,
                                               void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || (n > 0 && (!d_rowptr || !d_val)) || (ignore >= 0 && !d_col)) return fail(SPHB_ERR_INVALID, "prow_exponential_inplace_dev: bad arguments");
    if (max_row_nnz <= 0 || max_row_nnz > 1024) return fail(SPHB_ERR_UNSUPPORTED, "prow_exponential_inplace_dev: max_row_nnz must be in 1..1024");
    return prow_exponential_inplace_launch(n, d_rowptr, d_col, d_val, max_row_nnz, ignore, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_prow_exponential_inplace(int64_t n, const int64_t* rowptr, const uint32_t* col, float* val, int ignore) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || !rowptr || (!val && n > 0 && rowptr[n] > 0) || (ignore >= 0 && !col)) return fail(SPHB_ERR_INVALID, "prow_exponential_inplace: bad arguments");
    if (n == 0 || rowptr[n] == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    const int64_t nnz = rowptr[n];
    int64_t max_nnz = 0;
    for (int64_t i = 0; i < n; ++i) max_nnz = std::max(max_nnz, rowptr[i + 1] - rowptr[i]);
    if (max_nnz > 1024) return fail(SPHB_ERR_UNSUPPORTED, "prow_exponential_inplace: rows with more than 1024 entries unsupported");
    DevBuf<int64_t> drp;
    DevBuf<uint32_t> dcol;
    DevBuf<float> dval;
    SPHB_CUDA(drp.alloc(n + 1, s));
    SPHB_CUDA(dval.alloc(nnz, s));
    SPHB_CUDA(cudaMemcpyAsync(drp.p, rowptr, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dval.p, val, nnz * sizeof(float), cudaMemcpyHostToDevice, s));
    if (ignore >= 0) {
        SPHB_CUDA(dcol.alloc(nnz, s));
        SPHB_CUDA(cudaMemcpyAsync(dcol.p, col, nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    }
    if (int rc = prow_exponential_inplace_launch(n, drp.p, ignore >= 0 ? dcol.p : nullptr, dval.p, static_cast<int>(max_nnz), ignore, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(val, dval.p, nnz * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_prow_gaussian_flat(const float* dist, int64_t n, int k, double perplexity, float* temp) {
    if (int rc = ensure_init()) return rc;
    if (!dist || !temp || n < 0 || k < 1) return fail(SPHB_ERR_INVALID, "prow_flat: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = g_ctx.stream;
    const size_t nk = static_cast<size_t>(n) * k;
    DevBuf<float> dd, dt;
    SPHB_CUDA(dd.alloc(nk, s));
    SPHB_CUDA(dt.alloc(nk, s));
    SPHB_CUDA(cudaMemcpyAsync(dd.p, dist, nk * sizeof(float), cudaMemcpyHostToDevice, s));
    if (int rc = prow_flat_launch(dd.p, n, k, perplexity, dt.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(temp, dt.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}


// ---------------------------------------------------------------- single process, several GPUs
// The reference calls ComputeHierarchy::compute() from ONE host thread (sph/ComputeHierarchy.cpp:262-276); a drop-in for that
// thread cannot ask it to become one process per GPU.  sphb_knn_graph_prow_multi takes the same host buffers as
// sphb_knn_graph_prow plus a device list and runs one worker thread per device: worker g uploads ONLY its N/G rows of the stack
// from host memory, the workers exchange their slabs device to device (cudaMemcpyPeerAsync: NVLink where peer access exists) so
// that every GPU holds the whole database, then each searches its slab of query rows, builds its P rows and copies its four
// result slabs into its range of the caller's buffers.  Host-to-device traffic is N*d*4 bytes in total, as in the
// one-process-per-GPU path (sph_b200/multigpu.py), and there is no data-path collective on the result side.
namespace {
struct MultiBarrier {
    std::mutex mu;
    std::condition_variable cv;
    int count = 0, generation = 0, parties;
    explicit MultiBarrier(int p) : parties(p) {}
    void wait() {
        std::unique_lock<std::mutex> lk(mu);
        const int gen = generation;
        if (++count == parties) {
            count = 0;
            ++generation;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return gen != generation; });
        }
    }
};
}  // namespace

SPHB_API int sphb_knn_graph_prow_multi(const float* x, int64_t n, int d, int k, int metric, int l2_squared, float perplexity, const int* devices,
                                       int num_devices, int64_t* ids, float* dist, uint32_t* cols, float* vals) {
    if (!x || !ids || !dist || !cols || !vals || !devices) return fail(SPHB_ERR_INVALID, "knn_graph_prow_multi: null pointer");
    if (n <= 0 || d <= 0 || k < 2 || num_devices < 1 || num_devices > 64) return fail(SPHB_ERR_INVALID, "knn_graph_prow_multi: bad sizes");
    const int G = num_devices;
    const int64_t chunk = (n + G - 1) / G;
    std::vector<float*> dx(G, nullptr);
    std::vector<cudaStream_t> streams(G, nullptr);
    std::vector<int> rcs(G, SPHB_OK);
    std::vector<std::string> errs(G);
    MultiBarrier bar(G);
    auto worker = [&](int g) {
        int rc = sphb_init(devices[g]);
        const int64_t q0 = std::min<int64_t>(n, g * chunk), q1 = std::min<int64_t>(n, q0 + chunk), nq = q1 - q0;
        DevBuf<float> xall, dd, dv;
        DevBuf<int64_t> di;
        DevBuf<uint32_t> dc;
        cudaStream_t s = nullptr;
        auto check = [&](cudaError_t e, const char* what) {
            if (rc == SPHB_OK && e != cudaSuccess) rc = fail(SPHB_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
        };
        if (rc == SPHB_OK) {
            s = g_ctx.stream;
            check(xall.alloc(static_cast<size_t>(n) * d, s), "alloc stack");
            if (rc == SPHB_OK && nq > 0)
                check(cudaMemcpyAsync(xall.p + q0 * d, x + q0 * d, static_cast<size_t>(nq) * d * sizeof(float), cudaMemcpyHostToDevice, s), "H2D slab");
            check(cudaStreamSynchronize(s), "H2D slab");
        }
        dx[g] = xall.p;
        streams[g] = s;
        bar.wait();  // every worker's slab is in place (or it failed: rcs is checked after the exchange barrier)
        rcs[g] = rc;
        bar.wait();
        bool all_ok = true;
        for (int p = 0; p < G; ++p) all_ok &= rcs[p] == SPHB_OK;
        if (all_ok) {
            // all-gather by peer copies: pull every other worker's slab
            for (int o = 1; o < G && rc == SPHB_OK; ++o) {
                const int p = (g + o) % G;
                const int64_t p0 = std::min<int64_t>(n, p * chunk), p1 = std::min<int64_t>(n, p0 + chunk);
                if (p1 <= p0) continue;
                if (devices[p] != devices[g]) {
                    int can = 0;
                    cudaDeviceCanAccessPeer(&can, devices[g], devices[p]);
                    if (can) {
                        const cudaError_t e = cudaDeviceEnablePeerAccess(devices[p], 0);
                        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) check(e, "enable peer access");
                        cudaGetLastError();
                    }
                }
                check(cudaMemcpyPeerAsync(xall.p + p0 * d, devices[g], dx[p] + p0 * d, devices[p], static_cast<size_t>(p1 - p0) * d * sizeof(float), s),
                      "peer copy");
            }
            check(cudaStreamSynchronize(s), "peer copy");
        }
        bar.wait();  // nobody releases its copy of the stack while a peer still reads it
        if (all_ok && rc == SPHB_OK && nq > 0) {
            const size_t nk = static_cast<size_t>(nq) * k, ne = static_cast<size_t>(nq) * (k - 1);
            check(di.alloc(nk, s), "alloc");
            check(dd.alloc(nk, s), "alloc");
            check(dc.alloc(ne, s), "alloc");
            check(dv.alloc(ne, s), "alloc");
            if (rc == SPHB_OK) rc = sphb_knn_exact_dev(xall.p, n, d, k, metric, l2_squared, q0, q1, di.p, dd.p, s);
            if (rc == SPHB_OK) rc = prow_graph_launch(di.p, dd.p, nq, k, perplexity, 0, dc.p, dv.p, nullptr, s);
            if (rc == SPHB_OK) {
                check(cudaMemcpyAsync(ids + q0 * k, di.p, nk * sizeof(int64_t), cudaMemcpyDeviceToHost, s), "D2H");
                check(cudaMemcpyAsync(dist + q0 * k, dd.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s), "D2H");
                check(cudaMemcpyAsync(cols + q0 * (k - 1), dc.p, ne * sizeof(uint32_t), cudaMemcpyDeviceToHost, s), "D2H");
                check(cudaMemcpyAsync(vals + q0 * (k - 1), dv.p, ne * sizeof(float), cudaMemcpyDeviceToHost, s), "D2H");
                check(cudaStreamSynchronize(s), "D2H");
            }
        }
        if (rc != SPHB_OK) errs[g] = g_err;
        rcs[g] = rc;
        // buffers go back to the workspace cache of the worker's stream
        xall.release(); di.release(); dd.release(); dc.release(); dv.release();
        if (s) cudaStreamSynchronize(s);  // the context stays with the worker thread: its workspace cache serves the next call
    };
    {
        std::lock_guard<std::mutex> call_lock(g_multi_call_mu);
        while (static_cast<int>(g_multi_pool.size()) < G) g_multi_pool.push_back(new MultiWorker());
        for (int g = 0; g < G; ++g) g_multi_pool[g]->post([&worker, g] { worker(g); });
        for (int g = 0; g < G; ++g) g_multi_pool[g]->wait();
    }
    for (int g = 0; g < G; ++g)
        if (rcs[g] != SPHB_OK) return fail(rcs[g], "knn_graph_prow_multi: device " + std::to_string(devices[g]) + ": " + errs[g]);
    return SPHB_OK;
}

}  // extern "C"
