// Device-pointer (`_dev`) entry points of the CSR graph passes and the fused level pipeline
// (include/sphb.h, "resident level pipeline").  The host-pointer entry points of graph_ops.cu /
// graph_merge.cu / level_ops.cu upload, call the same internal *_dev functions and download; a caller
// that keeps the matrices in HBM between the stages of LevelSimilarities::compute
// (sph/LevelSimilarities.cpp:118-182) binds these instead and pays no PCIe traffic between kernels.
//   sphb_level_prob_dist[_dev] = useRandomWalks (sph/LevelSimilarities.cpp:460-508): createSimilarities or
//   top-k conversion -> in-place Gaussian rows -> zero removal (:566-581) -> symmetrizeOutput (:589-623).
#include "../../include/sphb.h"
#include "graph_common.cuh"

#include <vector>

namespace sphb {

int symmetrize_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int umap, int64_t* d_out_rowptr,
                   uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s);
int rw_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, const uint64_t* d_parents,
                 int64_t n_merged, int norm, int weight_by_size, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                 int64_t* out_nnz, cudaStream_t s);
int random_walks_slab_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks, uint64_t walk_len,
                          int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize, int64_t q0,
                          int64_t q1, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s);
int create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                            int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s);
int create_similarities_rows_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                                 int variant, const int64_t* d_comp_sizes, int64_t r0, int64_t r1, uint32_t* d_out_cols, float* d_out_vals,
                                 int32_t* d_row_nnz, cudaStream_t s);
int find_k_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int max_nnz, int k, int top, uint32_t* d_out_cols,
               float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s);
int prow_inplace_launch(int64_t n, const int64_t* d_rowptr, float* d_val, int max_nnz, float perplexity, int ignore, cudaStream_t stream);
int remove_zeros_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t* d_out_rowptr, uint32_t* d_out_col,
                     float* d_out_val, int64_t* out_nnz, cudaStream_t s);

namespace {

inline unsigned la_grid(int64_t work, int block, int sms, int per_sm = 8) {
    const int64_t need = (work + block - 1) / block;
    return static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>(need, static_cast<int64_t>(sms) * per_sm)));
}

__global__ void la_max_row_len(int64_t nrows, const int64_t* __restrict__ rowptr, int* __restrict__ max_len) {
    int m = 0;
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r < nrows; r += static_cast<int64_t>(gridDim.x) * blockDim.x)
        m = max(m, static_cast<int>(min(rowptr[r + 1] - rowptr[r], static_cast<int64_t>(0x7fffffff))));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(max_len, m);
}
__global__ void la_widen(const int32_t* __restrict__ in, int64_t num, int64_t* __restrict__ out) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t <= num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        out[t] = t < num ? in[t] : 0;
}
// fixed-stride rows (stride k, row_nnz valid entries each) -> CSR at the scanned offsets, one warp per row
__global__ void la_fixed_to_csr(int64_t n, int k, const uint32_t* __restrict__ cols, const float* __restrict__ vals, const int64_t* __restrict__ rowptr,
                                uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t b = rowptr[r];
        const int m = static_cast<int>(rowptr[r + 1] - b);
        for (int t = lane; t < m; t += 32) {
            out_col[b + t] = cols[r * k + t];
            out_val[b + t] = vals[r * k + t];
        }
    }
}

int max_row_len(int64_t n, const int64_t* d_rowptr, int* out, cudaStream_t s) {
    DevBuf<int> m;
    SPHB_CUDA(m.alloc(1, s));
    SPHB_CUDA(cudaMemsetAsync(m.p, 0, sizeof(int), s));
    la_max_row_len<<<la_grid(n, 256, ctx().num_sms), 256, 0, s>>>(n, d_rowptr, m.p);
    count_launch();
    SPHB_CUDA(cudaMemcpyAsync(out, m.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// per-stage device times of the last sphb_level_prob_dist[_dev] call (CUDA events on its stream)
thread_local float g_level_ms[5] = {0, 0, 0, 0, 0};

struct StageTimer {
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t s;
    explicit StageTimer(cudaStream_t stream) : s(stream) { for (auto& e : ev) cudaEventCreate(&e); }
    ~StageTimer() { for (auto& e : ev) if (e) cudaEventDestroy(e); }
    void mark(int i) { cudaEventRecord(ev[i], s); }
    void collect(int stages) {
        cudaEventSynchronize(ev[stages]);
        for (int i = 0; i < 5; ++i) g_level_ms[i] = 0.f;
        for (int i = 0; i < stages; ++i) cudaEventElapsedTime(&g_level_ms[i], ev[i], ev[i + 1]);
    }
};

}  // namespace

// The pipeline behind D1 / D3: fixed-stride rows (cols padded, row_nnz) -> CSR -> B3 -> B7 -> B6.  Split off so that the rows can
// come from anywhere — one call on this device, or slabs of output rows computed on several devices and gathered.
static int level_from_rows(int64_t n, int64_t k, const uint32_t* d_fcols, const float* d_fvals, const int32_t* d_fnz, float perplexity, int scheme,
                           int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz,
                           StageTimer& tm, cudaStream_t s) {
    const int sms = ctx().num_sms;
    tm.mark(1);
    // ---- to CSR
    DevBuf<int64_t> cnt64, rp;
    SPHB_CUDA(cnt64.alloc(n + 1, s));
    SPHB_CUDA(rp.alloc(n + 1, s));
    la_widen<<<la_grid(n + 1, 256, sms), 256, 0, s>>>(d_fnz, n, cnt64.p);
    count_launch();
    if (int rc = exclusive_sum<int64_t>(cnt64.p, rp.p, n + 1, s)) return rc;
    int64_t m1 = 0;
    SPHB_CUDA(cudaMemcpyAsync(&m1, rp.p + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    DevBuf<uint32_t> c1;
    DevBuf<float> v1;
    SPHB_CUDA(c1.alloc(std::max<int64_t>(m1, 1), s));
    SPHB_CUDA(v1.alloc(std::max<int64_t>(m1, 1), s));
    la_fixed_to_csr<<<la_grid(n * 32, 256, sms), 256, 0, s>>>(n, static_cast<int>(k), d_fcols, d_fvals, rp.p, c1.p, v1.p);
    count_launch();
    tm.mark(2);
    // ---- B3: in-place Gaussian rows (rows hold at most k <= 1024 entries)
    if (scheme == 1 && m1 > 0) {
        if (int rc = prow_inplace_launch(n, rp.p, v1.p, static_cast<int>(std::min<int64_t>(k, 1024)), perplexity, -1, s)) return rc;
    }
    tm.mark(3);
    // ---- B7: zero removal
    const bool sym = symmetrize != 0 && scheme == 1;
    DevBuf<int64_t> rp2;
    DevBuf<uint32_t> c2;
    DevBuf<float> v2;
    int64_t m2 = 0;
    if (sym) {
        SPHB_CUDA(rp2.alloc(n + 1, s));
        SPHB_CUDA(c2.alloc(std::max<int64_t>(m1, 1), s));
        SPHB_CUDA(v2.alloc(std::max<int64_t>(m1, 1), s));
        if (int rc = remove_zeros_dev(n, rp.p, c1.p, v1.p, rp2.p, c2.p, v2.p, &m2, s)) return rc;
    } else {
        if (m1 > cap) return fail(SPHB_ERR_CAPACITY, "level_prob_dist: output capacity too small (n * k always suffices)");
        if (int rc = remove_zeros_dev(n, rp.p, c1.p, v1.p, d_out_rowptr, d_out_col, d_out_val, &m2, s)) return rc;
    }
    tm.mark(4);
    // ---- B6: symmetrizeTSNE
    if (sym) {
        if (int rc = symmetrize_dev(n, rp2.p, c2.p, v2.p, m2, 0, d_out_rowptr, d_out_col, d_out_val, cap, &m2, s)) return rc;
    }
    tm.mark(5);
    tm.collect(5);
    *out_nnz = m2;
    return 0;
}

// The fused level pipeline on a resident random-walk matrix.  pair_sims = 1: createSimilarities(walks, k, prune_val)
// (randomWalkPairSims, the default), 0: per-row top-k conversion (convertEigenSparseVecToHDILibSparseVec).
// scheme: 1 t-SNE (in-place Gaussian rows with `perplexity`, ignore = -1; symmetrizeTSNE), 0 none.
int level_prob_dist_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                        int variant, const int64_t* d_comp_sizes, int pair_sims, float perplexity, int scheme, int symmetrize,
                        int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    *out_nnz = 0;
    if (n <= 0) return 0;
    if (k <= 0) return fail(SPHB_ERR_INVALID, "level_prob_dist: k must be positive");
    StageTimer tm(s);
    tm.mark(0);
    // ---- D1 / D3: fixed-stride rows of at most k entries
    DevBuf<uint32_t> fcols;
    DevBuf<float> fvals;
    DevBuf<int32_t> fnz;
    SPHB_CUDA(fcols.alloc(static_cast<size_t>(n) * k, s));
    SPHB_CUDA(fvals.alloc(static_cast<size_t>(n) * k, s));
    SPHB_CUDA(fnz.alloc(n, s));
    if (pair_sims) {
        if (int rc = create_similarities_dev(n, d_rowptr, d_col, d_val, nnz, k, prune_val, variant, d_comp_sizes, fcols.p, fvals.p, fnz.p, s)) return rc;
    } else {
        int mx = 0;
        if (int rc = max_row_len(n, d_rowptr, &mx, s)) return rc;
        if (int rc = find_k_dev(n, d_rowptr, d_col, d_val, mx, static_cast<int>(k), 1, fcols.p, fvals.p, fnz.p, s)) return rc;
    }
    return level_from_rows(n, k, fcols.p, fvals.p, fnz.p, perplexity, scheme, symmetrize, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, tm, s);
}

int level_from_sims_dev(int64_t n, int64_t k, const uint32_t* d_cols, const float* d_vals, const int32_t* d_row_nnz, float perplexity, int scheme,
                        int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    *out_nnz = 0;
    if (n <= 0) return 0;
    if (k <= 0) return fail(SPHB_ERR_INVALID, "level_from_sims: k must be positive");
    StageTimer tm(s);
    tm.mark(0);
    return level_from_rows(n, k, d_cols, d_vals, d_row_nnz, perplexity, scheme, symmetrize, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, tm, s);
}

}  // namespace sphb

using namespace sphb;

extern "C" {

SPHB_API int sphb_random_walks_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks,
                                   uint64_t walk_len, int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal,
                                   int normalize, int64_t q0, int64_t q1, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                                   int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_rowptr || !out_nnz) return fail(SPHB_ERR_INVALID, "random_walks_dev: null pointer");
    return random_walks_slab_dev(n, d_rowptr, d_col, d_val, num_walks, walk_len, weighting, prune_value, prune_steps, seed, remove_diagonal, normalize,
                                 q0, q1, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_rw_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, const uint64_t* d_parents,
                               int64_t n_merged, int norm, int weight_by_size, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val,
                               int64_t cap, int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_parents || !d_out_rowptr || !out_nnz || n < 0 || n_merged <= 0) return fail(SPHB_ERR_INVALID, "rw_merge_dev: bad arguments");
    return rw_merge_dev(n, d_rowptr, d_col, d_val, nnz, d_parents, n_merged, norm, weight_by_size, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz,
                        static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                          float prune_val, int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals,
                                          int32_t* d_row_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_cols || !d_out_vals || !d_row_nnz || n < 0 || k <= 0) return fail(SPHB_ERR_INVALID, "create_similarities_dev: bad arguments");
    if (n == 0) return SPHB_OK;
    return create_similarities_dev(n, d_rowptr, d_col, d_val, nnz, k, prune_val, variant, d_comp_sizes, d_out_cols, d_out_vals, d_row_nnz,
                                   static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_create_similarities_slab_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                               float prune_val, int variant, const int64_t* d_comp_sizes, int64_t r0, int64_t r1, uint32_t* d_out_cols,
                                               float* d_out_vals, int32_t* d_row_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_cols || !d_out_vals || !d_row_nnz || n < 0 || k <= 0) return fail(SPHB_ERR_INVALID, "create_similarities_slab_dev: bad arguments");
    return create_similarities_rows_dev(n, d_rowptr, d_col, d_val, nnz, k, prune_val, variant, d_comp_sizes, r0, r1, d_out_cols, d_out_vals, d_row_nnz,
                                        static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_level_from_sims_dev(int64_t n, int64_t k, const uint32_t* d_cols, const float* d_vals, const int32_t* d_row_nnz, float perplexity,
                                      int scheme, int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                                      int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || !d_cols || !d_vals || !d_row_nnz || !d_out_rowptr || !d_out_col || !d_out_val || !out_nnz)
        return fail(SPHB_ERR_INVALID, "level_from_sims_dev: bad arguments");
    return level_from_sims_dev(n, k, d_cols, d_vals, d_row_nnz, perplexity, scheme, symmetrize, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz,
                               static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_prow_gaussian_inplace_dev(int64_t n, const int64_t* d_rowptr, float* d_val, int max_row_nnz, float perplexity, int ignore,
                                            void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || n < 0) return fail(SPHB_ERR_INVALID, "prow_inplace_dev: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int mx = max_row_nnz;
    if (mx <= 0) {
        if (int rc = max_row_len(n, d_rowptr, &mx, s)) return rc;
        if (mx == 0) return SPHB_OK;
    }
    return prow_inplace_launch(n, d_rowptr, d_val, mx, perplexity, ignore, s);
}

SPHB_API int sphb_symmetrize_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int umap,
                                 int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_rowptr || !out_nnz || n < 0) return fail(SPHB_ERR_INVALID, "symmetrize_dev: bad arguments");
    return symmetrize_dev(n, d_rowptr, d_col, d_val, nnz, umap, d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_find_k_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int max_row_nnz, int64_t k, int top,
                             uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_cols || !d_out_vals || !d_row_nnz || n < 0 || k <= 0) return fail(SPHB_ERR_INVALID, "find_k_dev: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int mx = max_row_nnz;
    if (mx <= 0) {
        if (int rc = max_row_len(n, d_rowptr, &mx, s)) return rc;
    }
    return find_k_dev(n, d_rowptr, d_col, d_val, mx, static_cast<int>(k), top, d_out_cols, d_out_vals, d_row_nnz, s);
}

SPHB_API int sphb_level_prob_dist_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k,
                                      float prune_val, int variant, const int64_t* d_comp_sizes, int pair_sims, float perplexity, int scheme,
                                      int symmetrize, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz,
                                      void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_rowptr || !out_nnz || n < 0) return fail(SPHB_ERR_INVALID, "level_prob_dist_dev: bad arguments");
    return level_prob_dist_dev(n, d_rowptr, d_col, d_val, nnz, k, prune_val, variant, d_comp_sizes, pair_sims, perplexity, scheme, symmetrize,
                               d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_level_prob_dist(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, float prune_val, int variant,
                                  const int64_t* comp_sizes, int pair_sims, float perplexity, int scheme, int symmetrize, int64_t* out_rowptr,
                                  uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (!rowptr || !out_rowptr || !out_nnz || n < 0) return fail(SPHB_ERR_INVALID, "level_prob_dist: bad arguments");
    *out_nnz = 0;
    if (n == 0) return SPHB_OK;
    cudaStream_t s = ctx().stream;
    const int64_t nnz = rowptr[n];
    DevBuf<int64_t> drp, dcs, orp;
    DevBuf<uint32_t> dcol, oc;
    DevBuf<float> dval, ov;
    SPHB_CUDA(drp.alloc(n + 1, s));
    SPHB_CUDA(dcol.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(dval.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(cudaMemcpyAsync(drp.p, rowptr, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (nnz > 0) {
        SPHB_CUDA(cudaMemcpyAsync(dcol.p, col, nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        SPHB_CUDA(cudaMemcpyAsync(dval.p, val, nnz * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    if (comp_sizes) {
        SPHB_CUDA(dcs.alloc(n, s));
        SPHB_CUDA(cudaMemcpyAsync(dcs.p, comp_sizes, n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    }
    int64_t m = 0;
    if (int rc = level_prob_dist_dev(n, drp.p, dcol.p, dval.p, nnz, k, prune_val, variant, comp_sizes ? dcs.p : nullptr, pair_sims, perplexity, scheme,
                                     symmetrize, orp.p, oc.p, ov.p, cap, &m, s))
        return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m > 0) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_level_last_timing(float* ms5) {
    if (!ms5) return fail(SPHB_ERR_INVALID, "null pointer");
    for (int i = 0; i < 5; ++i) ms5[i] = g_level_ms[i];
    return SPHB_OK;
}

}  // extern "C"
