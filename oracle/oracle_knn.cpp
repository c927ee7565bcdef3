// TEST INFRASTRUCTURE ONLY — see oracle_common.hpp.
// Stage A oracle: exact kNN graph + the reference's post-passes.
//
// Follows the *call sites* of the reference (faiss itself is an external,
// un-vendored dependency — faiss 1.13.2, vcpkg-overlays/faiss/vcpkg.json:3):
//   sph/utils/Knn.cpp:56-108   computeButeForce -> faiss::knn_L2sqr / knn_inner_product
//   sph/utils/Knn.cpp:110-136  computeIndexFlat -> IndexFlat::search
//   sph/NearestNeighbors.cpp:208-222 initFaissData (COSINE renorm)
//   sph/NearestNeighbors.cpp:224-269 sqrt / eps clamp / monotone re-sort
//   sph/utils/GraphUtils.cpp:23-95   ensureClosestPointIsSelf
// Contract: the distance of a pair is the fp64 sequential
// FMA sum  acc = fma(xi[c]-xj[c], xi[c]-xj[c], acc), c = 0..d-1, rounded once
// to fp32; neighbours are ordered by (fp32 distance, index).  For the inner
// product metric the key is (-dot) with the same fp64 FMA chain.
#include "oracle_common.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <dlfcn.h>
#include <numeric>
#include <omp.h>

namespace {

enum Metric { METRIC_IP = 0, METRIC_L2 = 1 };  // faiss::MetricType values, as at the reference call sites

inline float pair_key(const float* xi, const float* xj, int d, int metric) {
    double acc = 0.0;
    if (metric == METRIC_L2) {
        for (int c = 0; c < d; ++c) {
            const double diff = static_cast<double>(xi[c]) - static_cast<double>(xj[c]);
            acc = std::fma(diff, diff, acc);
        }
        return static_cast<float>(acc);
    }
    for (int c = 0; c < d; ++c) acc = std::fma(static_cast<double>(xi[c]), static_cast<double>(xj[c]), acc);
    return static_cast<float>(acc);
}

struct Cand {
    float key;  // L2: squared distance; IP: NEGATED dot (so ascending order is right for both)
    int64_t idx;
};
inline bool cand_less(const Cand& a, const Cand& b) { return a.key < b.key || (a.key == b.key && a.idx < b.idx); }

void knn_rows(const float* x, int64_t n, int d, int k, int metric, const int64_t* rows, int64_t nrows, int64_t* ids, float* dist) {
#pragma omp parallel
    {
        std::vector<Cand> all(static_cast<size_t>(n));
#pragma omp for schedule(dynamic, 4)
        for (int64_t r = 0; r < nrows; ++r) {
            const int64_t i = rows ? rows[r] : r;
            const float* xi = x + i * d;
            for (int64_t j = 0; j < n; ++j) {
                const float v = pair_key(xi, x + j * d, d, metric);
                all[j] = Cand{metric == METRIC_L2 ? v : -v, j};
            }
            const int64_t kk = std::min<int64_t>(k, n);
            std::partial_sort(all.begin(), all.begin() + kk, all.end(), cand_less);
            for (int64_t t = 0; t < kk; ++t) {
                ids[r * k + t] = all[t].idx;
                dist[r * k + t] = metric == METRIC_L2 ? all[t].key : -all[t].key;
            }
            for (int64_t t = kk; t < k; ++t) {  // unfilled slots keep the caller's -1 / -1.0f convention
                ids[r * k + t] = -1;
                dist[r * k + t] = -1.0f;
            }
        }
    }
}

}  // namespace

// Exact kNN for all rows (queries == database, self included).
ORC_API void orc_knn_exact(const float* x, int64_t n, int d, int k, int metric, int64_t* ids, float* dist) {
    knn_rows(x, n, d, k, metric, nullptr, n, ids, dist);
}

// Exact kNN for a subset of query rows against the full database (used to
// check large configurations on a sample of rows).
ORC_API void orc_knn_exact_rows(const float* x, int64_t n, int d, int k, int metric, const int64_t* rows, int64_t nrows, int64_t* ids, float* dist) {
    knn_rows(x, n, d, k, metric, rows, nrows, ids, dist);
}

// COSINE preparation: faiss::fvec_renorm_L2 on a copy (NearestNeighbors.cpp:208-222).
// Restated: per row, nr = fp32 sum of squares (sequential), if nr > 0 scale by 1/sqrtf(nr).
ORC_API void orc_renorm_l2(const float* x, int64_t n, int d, float* out) {
    for (int64_t i = 0; i < n; ++i) {
        float nr = 0.f;
        for (int c = 0; c < d; ++c) nr += x[i * d + c] * x[i * d + c];
        if (nr > 0) {
            const float inv = 1.0f / sqrtf(nr);
            for (int c = 0; c < d; ++c) out[i * d + c] = x[i * d + c] * inv;
        } else {
            for (int c = 0; c < d; ++c) out[i * d + c] = x[i * d + c];
        }
    }
}

// Post-passes in the order NearestNeighbors::compute applies them
// (sph/NearestNeighbors.cpp:152-167): A6 eps clamp, A7 sqrt, A8 monotone
// re-sort (stable by distance), A9 self first.  Returns the number of rows A9 adjusted.
// row_ids (optional): the point id of each row (a slab of query rows); default row r <-> point r.
ORC_API int64_t orc_knn_postprocess_rows(int64_t* ids, float* dist, int64_t n, int k, int l2squared, const int64_t* row_ids);
ORC_API int64_t orc_knn_postprocess(int64_t* ids, float* dist, int64_t n, int k, int l2squared) {
    return orc_knn_postprocess_rows(ids, dist, n, k, l2squared, nullptr);
}
ORC_API int64_t orc_knn_postprocess_rows(int64_t* ids, float* dist, int64_t n, int k, int l2squared, const int64_t* row_ids) {
    const int64_t total = n * static_cast<int64_t>(k);
    // A6: ensureFloatingPointIntegrity (NearestNeighbors.cpp:233-242)
    for (int64_t e = 0; e < total; ++e)
        if (dist[e] <= FLT_EPSILON) dist[e] = 0.f;
    // A7: squareRootOfAllDistances (:224-231)
    if (!l2squared)
        for (int64_t e = 0; e < total; ++e) dist[e] = std::sqrt(dist[e]);
    // A8: checkDistancesAreNonDecreasing (:244-269) + synchronizedSort (Algorithms.hpp:57-93)
    for (int64_t p = 0; p < n; ++p) {
        float* dp = dist + p * k;
        int64_t* ip = ids + p * k;
        bool mono = true;
        for (int t = 1; t < k; ++t)
            if (dp[t] < dp[t - 1]) { mono = false; break; }
        if (!mono) {
            std::vector<int> perm(k);
            std::iota(perm.begin(), perm.end(), 0);
            std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return dp[a] < dp[b]; });
            std::vector<float> dn(k);
            std::vector<int64_t> in(k);
            for (int t = 0; t < k; ++t) { dn[t] = dp[perm[t]]; in[t] = ip[perm[t]]; }
            std::copy(dn.begin(), dn.end(), dp);
            std::copy(in.begin(), in.end(), ip);
        }
    }
    // A9: ensureClosestPointIsSelf (GraphUtils.cpp:23-95)
    int64_t adjusted = 0;
    for (int64_t row = 0; row < n; ++row) {
        const int64_t p = row_ids ? row_ids[row] : row;
        float* dp = dist + row * k;
        int64_t* ip = ids + row * k;
        if (ip[0] == p) continue;
        int swap_pos = k - 1;
        int found = -1;
        for (int t = 0; t < k; ++t)
            if (ip[t] == p) { found = t; break; }
        if (found >= 0) {
            swap_pos = found;
        } else {
            bool all_zero = true;
            for (int t = 0; t < k; ++t)
                if (dp[t] != 0.f) { all_zero = false; break; }
            if (all_zero) {
                ip[swap_pos] = p;
            } else {
                for (int t = k - 1; t > 0; --t) { ip[t] = ip[t - 1]; dp[t] = dp[t - 1]; }
                swap_pos = 0;
                ip[0] = p;
                dp[0] = 0.f;
            }
        }
        std::swap(ip[0], ip[swap_pos]);
        std::swap(dp[0], dp[swap_pos]);
        ++adjusted;
    }
    return adjusted;
}

// checkAllNeighborsExist (sph/utils/Knn.cpp:370-398): 1 if no id == -1.
ORC_API int orc_check_all_neighbors_exist(const int64_t* ids, int64_t total) {
    for (int64_t e = 0; e < total; ++e)
        if (ids[e] == -1) return 0;
    return 1;
}

// ---------------------------------------------------------------------------
// CPU *baseline* flavour of the same search: the algorithm faiss' knn_L2sqr
// uses for >= 20 queries (published faiss >= 1.7.3 behaviour, not present in
// /root/reference): row norms, blocked sgemm of X·Yᵀ (4096 queries x 1024
// database rows), d = |x|²+|y|²-2xy clamped at 0, per-row max-heap of size k,
// final ascending reorder.  Results are fp32-rounding dependent, so this
// variant is only *timed* (bench.py cpu_baseline / --impl reference); parity
// uses orc_knn_exact.  sgemm comes from the OpenBLAS bundled with numpy
// (dlopen'ed; path passed in by the caller).
// ---------------------------------------------------------------------------
namespace {
using sgemm_fn = void (*)(int order, int transA, int transB, int64_t M, int64_t N, int64_t K, float alpha, const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C, int64_t ldc);
sgemm_fn g_sgemm = nullptr;

inline void heap_replace_top(int k, float* hv, int64_t* hi, float v, int64_t id) {
    // max-heap keyed on value, 0-based sift-down
    int i = 0;
    for (;;) {
        int l = 2 * i + 1, r = l + 1, m = i;
        float mv = v;
        if (l < k && hv[l] > mv) { m = l; mv = hv[l]; }
        if (r < k && hv[r] > mv) { m = r; mv = hv[r]; }
        if (m == i) break;
        hv[i] = hv[m]; hi[i] = hi[m];
        i = m;
    }
    hv[i] = v; hi[i] = id;
}
using setnt_fn = void (*)(int);
using getnt_fn = int (*)();
setnt_fn g_set_threads = nullptr;
getnt_fn g_get_threads = nullptr;
double g_blas_stats[4] = {0, 0, 0, 0};  // sgemm seconds (mean per thread), heap seconds (mean per thread), sgemm flop, threads
}  // namespace

ORC_API int orc_blas_init(const char* openblas_path) {
    if (g_sgemm) return 0;
    void* h = dlopen(openblas_path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return -1;
    g_sgemm = reinterpret_cast<sgemm_fn>(dlsym(h, "scipy_cblas_sgemm64_"));
    if (!g_sgemm) g_sgemm = reinterpret_cast<sgemm_fn>(dlsym(h, "cblas_sgemm64_"));
    g_set_threads = reinterpret_cast<setnt_fn>(dlsym(h, "scipy_openblas_set_num_threads64_"));
    if (!g_set_threads) g_set_threads = reinterpret_cast<setnt_fn>(dlsym(h, "openblas_set_num_threads64_"));
    g_get_threads = reinterpret_cast<getnt_fn>(dlsym(h, "scipy_openblas_get_num_threads64_"));
    if (!g_get_threads) g_get_threads = reinterpret_cast<getnt_fn>(dlsym(h, "openblas_get_num_threads64_"));
    return g_sgemm ? 0 : -2;
}

// [0] sgemm seconds and [1] heap-update seconds of the last orc_knn_blas_l2 call (mean over its threads),
// [2] sgemm flop of the call, [3] threads used.
ORC_API void orc_knn_blas_last_stats(double* out4) {
    for (int i = 0; i < 4; ++i) out4[i] = g_blas_stats[i];
}

// A standalone sgemm of the arm's block shape (4096 x 1024 x d) with the library's own threading, repeated for
// about `seconds`: the host's sgemm rate the arm's achieved rate is compared with.  Returns GFLOP/s.
ORC_API double orc_sgemm_standalone_gflops(int d, double seconds) {
    if (!g_sgemm) return -1.0;
    const int64_t M = 4096, N = 1024;
    std::vector<float> A(static_cast<size_t>(M) * d, 0.5f), B(static_cast<size_t>(N) * d, 0.25f), Cm(static_cast<size_t>(M) * N);
    g_sgemm(101, 111, 112, M, N, d, 1.0f, A.data(), d, B.data(), d, 0.0f, Cm.data(), N);
    const double t0 = omp_get_wtime();
    int reps = 0;
    while (omp_get_wtime() - t0 < seconds) {
        g_sgemm(101, 111, 112, M, N, d, 1.0f, A.data(), d, B.data(), d, 0.0f, Cm.data(), N);
        ++reps;
    }
    return 2.0 * M * N * d * reps / (omp_get_wtime() - t0) / 1e9;
}

// queries = rows [q0, q1) of x; database = all n rows.  ids/dist sized (q1-q0)*k.
// faiss' blocking (4096 queries x 1024 database rows per sgemm, then per-row max-heap updates).  Threading: faiss runs
// a multi-threaded BLAS sgemm and then an OpenMP loop over the block's rows; with a pthread OpenBLAS next to libgomp
// the two pools spin against each other between the phases (round 1: 46 GFLOP/s on 16 cores).  Here the OpenMP team
// owns the whole block: each thread takes a strip of the 4096 queries, calls a SINGLE-threaded sgemm for its strip
// against the 1024-row database block and updates its own rows' heaps — same arithmetic per (query, database) pair,
// same heap admission order (database index ascending), no cross-pool hand-off.
ORC_API int orc_knn_blas_l2(const float* x, int64_t n, int d, int k, int64_t q0, int64_t q1, int64_t* ids, float* dist) {
    if (!g_sgemm) return -1;
    const int64_t nq = q1 - q0;
    std::vector<float> norms(static_cast<size_t>(n));
#pragma omp parallel for
    for (int64_t i = 0; i < n; ++i) {
        float s = 0.f;
        for (int c = 0; c < d; ++c) s += x[i * d + c] * x[i * d + c];
        norms[i] = s;
    }
    const int64_t bs_x = 4096, bs_y = 1024;
    // strip of queries per single-threaded sgemm: every thread of the team gets one (32..128 rows)
    const int64_t strip = std::max<int64_t>(32, std::min<int64_t>(128, ((std::min(nq, bs_x) + omp_get_max_threads() - 1) / omp_get_max_threads() + 15) / 16 * 16));
    std::vector<float> hv(static_cast<size_t>(nq) * k, FLT_MAX);
    std::vector<int64_t> hi(static_cast<size_t>(nq) * k, -1);
    const int prev_threads = g_get_threads ? g_get_threads() : 0;
    if (g_set_threads) g_set_threads(1);
    double t_gemm = 0.0, t_heap = 0.0;
    int threads = 1;
    for (int64_t i0 = 0; i0 < nq; i0 += bs_x) {
        const int64_t i1 = std::min(nq, i0 + bs_x);
        const int64_t nstrips = (i1 - i0 + strip - 1) / strip;
#pragma omp parallel reduction(+ : t_gemm, t_heap)
        {
            std::vector<float> ip(static_cast<size_t>(strip * bs_y));
#pragma omp single
            threads = omp_get_num_threads();
            for (int64_t j0 = 0; j0 < n; j0 += bs_y) {
                const int64_t j1 = std::min(n, j0 + bs_y);
#pragma omp for schedule(static) nowait
                for (int64_t sp = 0; sp < nstrips; ++sp) {
                    const int64_t a = i0 + sp * strip, b = std::min(i1, a + strip);
                    double t0 = omp_get_wtime();
                    // C[i][j] = <x_i, y_j>, row-major (CblasRowMajor=101, NoTrans=111, Trans=112)
                    g_sgemm(101, 111, 112, b - a, j1 - j0, d, 1.0f, x + (q0 + a) * d, d, x + j0 * d, d, 0.0f, ip.data(), j1 - j0);
                    double t1 = omp_get_wtime();
                    t_gemm += t1 - t0;
                    for (int64_t i = a; i < b; ++i) {
                        float* v = hv.data() + i * k;
                        int64_t* id = hi.data() + i * k;
                        const float* row = ip.data() + (i - a) * (j1 - j0);
                        const float ni = norms[q0 + i];
                        for (int64_t j = j0; j < j1; ++j) {
                            float dis = ni + norms[j] - 2 * row[j - j0];
                            if (dis < 0) dis = 0;
                            if (dis < v[0]) heap_replace_top(k, v, id, dis, j);
                        }
                    }
                    t_heap += omp_get_wtime() - t1;
                }
            }
        }
    }
    if (g_set_threads && prev_threads > 0) g_set_threads(prev_threads);
    g_blas_stats[0] = t_gemm / threads;
    g_blas_stats[1] = t_heap / threads;
    g_blas_stats[2] = 2.0 * static_cast<double>(nq) * static_cast<double>(n) * d;
    g_blas_stats[3] = threads;
#pragma omp parallel for
    for (int64_t i = 0; i < nq; ++i) {
        std::vector<std::pair<float, int64_t>> tmp(k);
        for (int t = 0; t < k; ++t) tmp[t] = {hv[i * k + t], hi[i * k + t]};
        std::sort(tmp.begin(), tmp.end());
        for (int t = 0; t < k; ++t) { dist[i * k + t] = tmp[t].first; ids[i * k + t] = tmp[t].second; }
    }
    return 0;
}
