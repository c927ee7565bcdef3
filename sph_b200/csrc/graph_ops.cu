// CSR graph passes, part 1: symmetrisation of the P matrix (B6), random-walk merge (RW2),
// batched pair similarities (D4).  All are HBM-bound gather / sort passes over the kNN-sized
// sparse matrices; the sort itself is CUB's LSD radix sort (library plumbing), the expansion,
// combination and row-pointer kernels are ours.
//   B6  sph/utils/HDILibHelper.hpp:260-302           symmetrizeTSNE / symmetrizeUMAP
//   RW2 sph/utils/SparseMatrixAlgorithms.cpp:292-352 mergeNodesRandomWalks (+ :630-640 normalise)
//   D4  sph/utils/Similarities.cpp:374-396           randomWalksBhattacharyya / SingleOverlap
#include "../../include/sphb.h"
#include "graph_common.cuh"

namespace sphb {

// ---------------------------------------------------------------------------------------------
// shared: row pointers of a sorted key array whose row id is (key >> shift)
// ---------------------------------------------------------------------------------------------
__global__ void rowptr_from_keys(const uint64_t* __restrict__ keys, int64_t num, int shift, int64_t nrows, int64_t* __restrict__ rowptr) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r <= nrows; r += static_cast<int64_t>(gridDim.x) * blockDim.x)
        rowptr[r] = r == nrows ? num : lower_bound_u64(keys, num, static_cast<uint64_t>(r) << shift);
}

static inline unsigned grid_for(int64_t work, int block, int sms, int per_sm = 8) {
    const int64_t need = (work + block - 1) / block;
    return static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>(need, static_cast<int64_t>(sms) * per_sm)));
}

// ---------------------------------------------------------------------------------------------
// B6: P <- f(P, Pᵀ) on the union pattern.
// Every entry (r, c, v) emits the record ((r, c), tag 0, v) and ((c, r), tag 1, v); after a sort by
// (row, col, tag) every position of the union pattern is a run of one or two records:
// direct value (tag 0, 0 if missing) and transposed value (tag 1, 0 if missing).
//   t-SNE: (a + b) / 2      UMAP: (a + b) - a b     (sequential reference loop reduces to this)
// ---------------------------------------------------------------------------------------------
__global__ void sym_expand(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                           uint64_t* __restrict__ keys, float* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) {
            const uint64_t c = col[e];
            const float v = val[e];
            keys[2 * e] = ((static_cast<uint64_t>(r) << 31 | c) << 1);
            vals[2 * e] = v;
            keys[2 * e + 1] = ((c << 31 | static_cast<uint64_t>(r)) << 1) | 1ull;
            vals[2 * e + 1] = v;
        }
    }
}
__global__ void sym_flag_heads(const uint64_t* __restrict__ keys, int64_t num, int* __restrict__ head) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        head[t] = (t == 0 || (keys[t] >> 1) != (keys[t - 1] >> 1)) ? 1 : 0;
}
__global__ void sym_combine(const uint64_t* __restrict__ keys, const float* __restrict__ vals, const int* __restrict__ head,
                            const int* __restrict__ pos, int64_t num, int umap, uint64_t* __restrict__ ukeys, uint32_t* __restrict__ out_col,
                            float* __restrict__ out_val) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        if (!head[t]) continue;
        float a = 0.f, b = 0.f;  // a = P[r][c], b = P[c][r]
        const uint64_t key = keys[t];
        if (key & 1ull) b = vals[t];
        else a = vals[t];
        if (t + 1 < num && (keys[t + 1] >> 1) == (key >> 1)) b = vals[t + 1];
        const float v = umap ? __fsub_rn(__fadd_rn(a, b), __fmul_rn(a, b)) : (a + b) * 0.5f;  // no FMA contraction: the reference rounds the product
        const int p = pos[t];
        ukeys[p] = key >> 1;  // (row << 31) | col
        out_col[p] = static_cast<uint32_t>((key >> 1) & 0x7fffffffull);
        out_val[p] = v;
    }
}

int symmetrize_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int umap, int64_t* d_out_rowptr,
                   uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    if (n >= (1ll << 31)) return fail(SPHB_ERR_UNSUPPORTED, "symmetrize: n must be < 2^31");
    const int sms = ctx().num_sms;
    const int64_t num = 2 * nnz;
    if (nnz == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (n + 1) * sizeof(int64_t), s));
        *out_nnz = 0;
        return 0;
    }
    DevBuf<uint64_t> k0, k1, uk;
    DevBuf<float> v0, v1;
    DevBuf<int> head, pos;
    SPHB_CUDA(k0.alloc(num, s));
    SPHB_CUDA(k1.alloc(num, s));
    SPHB_CUDA(v0.alloc(num, s));
    SPHB_CUDA(v1.alloc(num, s));
    SPHB_CUDA(head.alloc(num + 1, s));
    SPHB_CUDA(pos.alloc(num + 1, s));
    sym_expand<<<grid_for(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, k0.p, v0.p);
    count_launch();
    if (int rc = sort_pairs<float>(k0.p, k1.p, v0.p, v1.p, num, 32 + bits_for(static_cast<uint64_t>(n)), s)) return rc;  // tag 1 + col 31 + row bits
    sym_flag_heads<<<grid_for(num, 256, sms), 256, 0, s>>>(k1.p, num, head.p);
    SPHB_CUDA(cudaMemsetAsync(head.p + num, 0, sizeof(int), s));
    count_launch();
    if (int rc = exclusive_sum<int>(head.p, pos.p, num + 1, s)) return rc;
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, pos.p + num, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (total > cap) return fail(SPHB_ERR_CAPACITY, "symmetrize: output capacity too small (2*nnz always suffices)");
    SPHB_CUDA(uk.alloc(total, s));
    sym_combine<<<grid_for(num, 256, sms), 256, 0, s>>>(k1.p, v1.p, head.p, pos.p, num, umap, uk.p, d_out_col, d_out_val);
    rowptr_from_keys<<<grid_for(n + 1, 256, sms), 256, 0, s>>>(uk.p, total, 31, n, d_out_rowptr);
    count_launch(2);
    SPHB_CUDA(cudaGetLastError());
    *out_nnz = total;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// RW2: M[parent(r)][parent(c)] += A[r][c] * w_r ; rows / sum(w) ; optional L1 normalisation.
// ---------------------------------------------------------------------------------------------
__global__ void merge_expand(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                             const uint64_t* __restrict__ parents, int weight_by_size, uint64_t* __restrict__ keys, float* __restrict__ vals,
                             unsigned long long* __restrict__ row_weights) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t b = rowptr[r], e1 = rowptr[r + 1];
        const uint64_t into = parents[r];
        const long long w = weight_by_size ? static_cast<long long>(e1 - b) : 1ll;
        if (lane == 0) atomicAdd(row_weights + into, static_cast<unsigned long long>(w));
        for (int64_t e = b + lane; e < e1; e += 32) {
            keys[e] = (into << 32) | (parents[col[e]] & 0xffffffffull);
            vals[e] = val[e] * static_cast<float>(w);
        }
    }
}
// one thread per run of equal keys: sequential fp32 sum in input (row-major) order — the order of the
// reference's single-thread scatter-add, because the radix sort is stable
__global__ void merge_reduce(const uint64_t* __restrict__ keys, const float* __restrict__ vals, const int* __restrict__ head,
                             const int* __restrict__ pos, int64_t num, const unsigned long long* __restrict__ row_weights, int weight_by_size,
                             uint64_t* __restrict__ ukeys, uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        if (!head[t]) continue;
        const uint64_t key = keys[t];
        float acc = 0.f;
        for (int64_t u = t; u < num && keys[u] == key; ++u) acc += vals[u];
        if (weight_by_size) acc /= static_cast<float>(row_weights[key >> 32]);
        const int p = pos[t];
        ukeys[p] = key;
        out_col[p] = static_cast<uint32_t>(key);
        out_val[p] = acc;
    }
}
__global__ void flag_heads_exact(const uint64_t* __restrict__ keys, int64_t num, int* __restrict__ head) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        head[t] = (t == 0 || keys[t] != keys[t - 1]) ? 1 : 0;
}
// normalizeUnitSparseVector (sph/utils/SparseMatrixAlgorithms.cpp:617-628): v /= fp32 sum in storage order
__global__ void normalize_rows_l1(int64_t n, const int64_t* __restrict__ rowptr, float* __restrict__ val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t b = rowptr[r], e1 = rowptr[r + 1];
        float s = 0.f;
        if (lane == 0)
            for (int64_t e = b; e < e1; ++e) s += val[e];
        s = __shfl_sync(0xffffffffu, s, 0);
        for (int64_t e = b + lane; e < e1; e += 32) val[e] = val[e] / s;
    }
}

int rw_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, const uint64_t* d_parents,
                 int64_t n_merged, int norm, int weight_by_size, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap,
                 int64_t* out_nnz, cudaStream_t s) {
    if (n_merged >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "rw_merge: n_merged must be < 2^32");
    const int sms = ctx().num_sms;
    if (nnz == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (n_merged + 1) * sizeof(int64_t), s));
        *out_nnz = 0;
        return 0;
    }
    DevBuf<uint64_t> k0, k1, uk;
    DevBuf<float> v0, v1;
    DevBuf<int> head, pos;
    DevBuf<unsigned long long> rw;
    SPHB_CUDA(k0.alloc(nnz, s));
    SPHB_CUDA(k1.alloc(nnz, s));
    SPHB_CUDA(v0.alloc(nnz, s));
    SPHB_CUDA(v1.alloc(nnz, s));
    SPHB_CUDA(head.alloc(nnz + 1, s));
    SPHB_CUDA(pos.alloc(nnz + 1, s));
    SPHB_CUDA(rw.alloc(n_merged, s));
    SPHB_CUDA(cudaMemsetAsync(rw.p, 0, n_merged * sizeof(unsigned long long), s));
    merge_expand<<<grid_for(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, d_parents, weight_by_size, k0.p, v0.p, rw.p);
    count_launch();
    if (int rc = sort_pairs<float>(k0.p, k1.p, v0.p, v1.p, nnz, 32 + bits_for(static_cast<uint64_t>(n_merged)), s)) return rc;
    flag_heads_exact<<<grid_for(nnz, 256, sms), 256, 0, s>>>(k1.p, nnz, head.p);
    SPHB_CUDA(cudaMemsetAsync(head.p + nnz, 0, sizeof(int), s));
    count_launch();
    if (int rc = exclusive_sum<int>(head.p, pos.p, nnz + 1, s)) return rc;
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, pos.p + nnz, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (total > cap) return fail(SPHB_ERR_CAPACITY, "rw_merge: output capacity too small (nnz of the input always suffices)");
    SPHB_CUDA(uk.alloc(total, s));
    merge_reduce<<<grid_for(nnz, 256, sms), 256, 0, s>>>(k1.p, v1.p, head.p, pos.p, nnz, rw.p, weight_by_size, uk.p, d_out_col, d_out_val);
    rowptr_from_keys<<<grid_for(n_merged + 1, 256, sms), 256, 0, s>>>(uk.p, total, 32, n_merged, d_out_rowptr);
    count_launch(2);
    if (norm) {
        normalize_rows_l1<<<grid_for(n_merged * 32, 256, sms), 256, 0, s>>>(n_merged, d_out_rowptr, d_out_val);
        count_launch();
    }
    SPHB_CUDA(cudaGetLastError());
    *out_nnz = total;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// D4: batched pair similarities, one warp per pair.
//   mode 0  Bhattacharyya coefficient  sum_c sqrt(A[a][c] * A[b][c]) over the columns of row a with A[b][c] > 0
//   mode 1  single overlap             max(A[a][b], A[b][a])
// ---------------------------------------------------------------------------------------------
__global__ void pair_sims_kernel(const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                                 const int64_t* __restrict__ pa, const int64_t* __restrict__ pb, int64_t npairs, int mode, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t p = warp; p < npairs; p += nwarps) {
        const int64_t a = pa[p], b = pb[p];
        const int64_t ab = rowptr[a], ae = rowptr[a + 1], bb = rowptr[b], be = rowptr[b + 1];
        if (mode == 1) {
            if (lane == 0) {
                const int64_t e1 = find_col(col, ab, ae, static_cast<uint32_t>(b));
                const int64_t e2 = find_col(col, bb, be, static_cast<uint32_t>(a));
                const float v1 = e1 >= 0 ? val[e1] : 0.f, v2 = e2 >= 0 ? val[e2] : 0.f;
                out[p] = fmaxf(v1, v2);
            }
            continue;
        }
        // The reference accumulates fp32 in the column order of row a; a warp-wide sum would change the
        // rounding, so lanes gather the matching products into order-preserving chunks of 32 and lane 0
        // adds each chunk in order.
        float acc = 0.f;
        for (int64_t base = ab; base < ae; base += 32) {
            const int64_t e = base + lane;
            float term = 0.f;
            if (e < ae) {
                const int64_t f = find_col(col, bb, be, col[e]);
                if (f >= 0 && val[f] > 0.f) term = sqrtf(val[e] * val[f]);
            }
#pragma unroll 1
            for (int l = 0; l < 32; ++l) {
                const float t = __shfl_sync(0xffffffffu, term, l);
                if (t != 0.f) acc += t;
            }
        }
        if (lane == 0) out[p] = acc;
    }
}

int pair_sims_dev(const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, const int64_t* d_pa, const int64_t* d_pb, int64_t npairs,
                  int mode, float* d_out, cudaStream_t s) {
    if (npairs <= 0) return 0;
    pair_sims_kernel<<<grid_for(npairs * 32, 256, ctx().num_sms), 256, 0, s>>>(d_rowptr, d_col, d_val, d_pa, d_pb, npairs, mode, d_out);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

// implemented in walks.cu / sims.cu
int random_walks_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks, uint64_t walk_len,
                     int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize,
                     int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s);
int create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                            int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s);
int create_similarities_rows_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                                 int variant, const int64_t* d_comp_sizes, int64_t r0, int64_t r1, uint32_t* d_out_cols, float* d_out_vals,
                                 int32_t* d_row_nnz, cudaStream_t s);

}  // namespace sphb

using namespace sphb;

// ---------------------------------------------------------------------------------------------
// host-pointer entry points: copy the CSR up, run, copy the result back
// ---------------------------------------------------------------------------------------------
namespace {
struct DevCsr {
    DevBuf<int64_t> rowptr;
    DevBuf<uint32_t> col;
    DevBuf<float> val;
    int64_t nnz = 0;
};
int upload_csr(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, DevCsr& d, cudaStream_t s) {
    d.nnz = rowptr[n];
    SPHB_CUDA(d.rowptr.alloc(n + 1, s));
    SPHB_CUDA(d.col.alloc(d.nnz, s));
    SPHB_CUDA(d.val.alloc(d.nnz, s));
    SPHB_CUDA(cudaMemcpyAsync(d.rowptr.p, rowptr, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (d.nnz) {
        SPHB_CUDA(cudaMemcpyAsync(d.col.p, col, d.nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        SPHB_CUDA(cudaMemcpyAsync(d.val.p, val, d.nnz * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    return 0;
}
bool bad_csr(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val) {
    return n < 0 || !rowptr || rowptr[0] != 0 || rowptr[n] < 0 || (rowptr[n] > 0 && (!col || !val));
}
}  // namespace

extern "C" {

SPHB_API int sphb_symmetrize(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int umap, int64_t* out_rowptr,
                             uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || !out_rowptr || !out_nnz || cap < 0) return fail(SPHB_ERR_INVALID, "symmetrize: bad arguments");
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<int64_t> orp;
    DevBuf<uint32_t> oc;
    DevBuf<float> ov;
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = symmetrize_dev(n, in.rowptr.p, in.col.p, in.val.p, in.nnz, umap, orp.p, oc.p, ov.p, cap, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_rw_merge(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, const uint64_t* parents, int64_t n_merged,
                           int norm, int weight_by_size, int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || !parents || n_merged <= 0 || !out_rowptr || !out_nnz) return fail(SPHB_ERR_INVALID, "rw_merge: bad arguments");
    for (int64_t i = 0; i < n; ++i)
        if (parents[i] >= static_cast<uint64_t>(n_merged)) return fail(SPHB_ERR_INVALID, "rw_merge: parent id out of range");
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<uint64_t> dpar;
    DevBuf<int64_t> orp;
    DevBuf<uint32_t> oc;
    DevBuf<float> ov;
    SPHB_CUDA(dpar.alloc(std::max<int64_t>(n, 1), s));
    SPHB_CUDA(cudaMemcpyAsync(dpar.p, parents, n * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(orp.alloc(n_merged + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = rw_merge_dev(n, in.rowptr.p, in.col.p, in.val.p, in.nnz, dpar.p, n_merged, norm, weight_by_size, orp.p, oc.p, ov.p, cap, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n_merged + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_pair_sims(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, const int64_t* pa, const int64_t* pb,
                            int64_t npairs, int mode, float* out) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || npairs < 0 || (npairs > 0 && (!pa || !pb || !out)) || (mode != 0 && mode != 1))
        return fail(SPHB_ERR_INVALID, "pair_sims: bad arguments");
    for (int64_t p = 0; p < npairs; ++p)
        if (pa[p] < 0 || pa[p] >= n || pb[p] < 0 || pb[p] >= n) return fail(SPHB_ERR_INVALID, "pair_sims: row id out of range");
    if (npairs == 0) return SPHB_OK;
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<int64_t> da, db;
    DevBuf<float> dout;
    SPHB_CUDA(da.alloc(npairs, s));
    SPHB_CUDA(db.alloc(npairs, s));
    SPHB_CUDA(dout.alloc(npairs, s));
    SPHB_CUDA(cudaMemcpyAsync(da.p, pa, npairs * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(db.p, pb, npairs * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (int rc = pair_sims_dev(in.rowptr.p, in.col.p, in.val.p, da.p, db.p, npairs, mode, dout.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out, dout.p, npairs * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_random_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, uint64_t num_walks, uint64_t walk_len,
                               int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize,
                               int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || !out_rowptr || !out_nnz || cap < 0 || weighting < 0 || weighting > 4 || walk_len == 0)
        return fail(SPHB_ERR_INVALID, "random_walks: bad arguments");
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<int64_t> orp;
    DevBuf<uint32_t> oc;
    DevBuf<float> ov;
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = random_walks_dev(n, in.rowptr.p, in.col.p, in.val.p, num_walks, walk_len, weighting, prune_value, prune_steps, seed, remove_diagonal,
                                  normalize, orp.p, oc.p, ov.p, cap, &m, s))
        return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_create_similarities(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, float prune_val,
                                      int variant, const int64_t* comp_sizes, uint32_t* out_cols, float* out_vals, int32_t* row_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || k <= 0 || !out_cols || !out_vals || !row_nnz || variant < 0 || variant > 2)
        return fail(SPHB_ERR_INVALID, "create_similarities: bad arguments");
    if (n == 0) return SPHB_OK;
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<int64_t> dcs;
    DevBuf<uint32_t> oc;
    DevBuf<float> ov;
    DevBuf<int32_t> on;
    if (comp_sizes) {
        SPHB_CUDA(dcs.alloc(n, s));
        SPHB_CUDA(cudaMemcpyAsync(dcs.p, comp_sizes, n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    }
    const size_t nk = static_cast<size_t>(n) * static_cast<size_t>(k);
    SPHB_CUDA(oc.alloc(nk, s));
    SPHB_CUDA(ov.alloc(nk, s));
    SPHB_CUDA(on.alloc(n, s));
    if (int rc = create_similarities_dev(n, in.rowptr.p, in.col.p, in.val.p, in.nnz, k, prune_val, variant, comp_sizes ? dcs.p : nullptr, oc.p, ov.p,
                                         on.p, s))
        return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_cols, oc.p, nk * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_vals, ov.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(row_nnz, on.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_create_similarities_slab(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, float prune_val,
                                           int variant, const int64_t* comp_sizes, int64_t r0, int64_t r1, uint32_t* out_cols, float* out_vals,
                                           int32_t* row_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr(n, rowptr, col, val) || k <= 0 || !out_cols || !out_vals || !row_nnz || variant < 0 || variant > 2)
        return fail(SPHB_ERR_INVALID, "create_similarities: bad arguments");
    if (r0 < 0 || r1 > n || r0 > r1) return fail(SPHB_ERR_INVALID, "create_similarities_slab: row range outside [0, n]");
    if (r0 == r1) return SPHB_OK;
    const int64_t nr = r1 - r0;
    cudaStream_t s = ctx().stream;
    DevCsr in;
    if (int rc = upload_csr(n, rowptr, col, val, in, s)) return rc;
    DevBuf<int64_t> dcs;
    DevBuf<uint32_t> oc;
    DevBuf<float> ov;
    DevBuf<int32_t> on;
    if (comp_sizes) {
        SPHB_CUDA(dcs.alloc(n, s));
        SPHB_CUDA(cudaMemcpyAsync(dcs.p, comp_sizes, n * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    }
    const size_t nk = static_cast<size_t>(nr) * static_cast<size_t>(k);
    SPHB_CUDA(oc.alloc(nk, s));
    SPHB_CUDA(ov.alloc(nk, s));
    SPHB_CUDA(on.alloc(nr, s));
    if (int rc = create_similarities_rows_dev(n, in.rowptr.p, in.col.p, in.val.p, in.nnz, k, prune_val, variant, comp_sizes ? dcs.p : nullptr, r0, r1,
                                              oc.p, ov.p, on.p, s))
        return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_cols, oc.p, nk * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_vals, ov.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(row_nnz, on.p, nr * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

}  // extern "C"
