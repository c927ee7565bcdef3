// Stage D, per-level helpers:
//   D3  utils::findTopK / findBottomK  (sph/utils/SparseMatrixAlgorithms.cpp:720-776) with the admission rules of
//       SortedMax/MinPairSizeDequeR (sph/utils/MaxSizeDeque.hpp:97-200), batched over the rows of a CSR matrix —
//       what convertEigenSparseVecToHDILibSparseVec does per row when randomWalkPairSims = false
//       (sph/LevelSimilarities.cpp:481-486).
//   D5  simNeighborOverlap / representedOverlap (sph/utils/Similarities.cpp:174-228): card(U_a ∩ U_b) / min(card) with
//       U_c the sorted-unique union of the data-level kNN lists of the pixels of component c.  The unions of ALL
//       components of a level are built once (expand (component, id) keys, one radix sort, unique) instead of once per
//       pair as the reference does; each pair is then a binary-search intersection by one warp.  Counts are integers,
//       the result is exact.
#include "../../include/sphb.h"
#include "graph_common.cuh"

#include <cfloat>

namespace sphb {

// ---------------------------------------------------------------------------------------------
// D3
// ---------------------------------------------------------------------------------------------
// One warp per row, rows of at most `cap` entries (cap power of two, shared memory).  The reference feeds the row, in
// column order, through a fixed-size sorted deque that (i) admits a value only if it is STRICTLY better than the
// current worst, (ii) places it in front of equal values, (iii) evicts from the back — among equal worst values the
// one that arrived first.  Closed form of that process (T = the k-th best admissible value, "G" = better than T,
// "E" = equal to T, subsequence = the G and E entries in column order):
//   * every G is kept;
//   * the first k entries of the subsequence are all admitted (the deque is not yet full of >= T values); after that
//     every E is rejected and every G evicts the oldest E still held;
//   * so the kept E's are the E's among the first k subsequence entries minus the first (lowest-column) n of them,
//     n = number of G's beyond subsequence position k.
// Values not strictly better than the deque's initial sentinel (FLT_MIN for top, FLT_MAX for bottom) are never
// admitted.  Output column-sorted, fixed stride k, padded 0xFFFFFFFF.
__global__ void __launch_bounds__(128) find_k_kernel(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col,
                                                     const float* __restrict__ val, int k, int top, int cap, uint32_t* __restrict__ out_cols,
                                                     float* __restrict__ out_vals, int32_t* __restrict__ row_nnz, int* __restrict__ err) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t r = gw; r < n; r += nw) {
        const int64_t b = rowptr[r];
        const int len = static_cast<int>(rowptr[r + 1] - b);
        if (len > cap) {
            if (lane == 0) atomicExch(err, 1);
            continue;
        }
        auto admissible = [&](float v) { return top ? (v > FLT_MIN) : (v < FLT_MAX); };
        // order value: ascending = best first
        int n_adm = 0;
        for (int t = lane; t < cap; t += 32) {
            uint64_t key = ~0ull;
            if (t < len) {
                const float v = val[b + t];
                if (admissible(v)) { key = make_key(top ? -v : v, col[b + t]); ++n_adm; }
            }
            sm[t] = key;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n_adm += __shfl_xor_sync(0xffffffffu, n_adm, o);
        bool all = n_adm <= k;
        float T = 0.f;  // order value of the k-th best
        int later_g = 0;
        if (!all) {
            __syncwarp();
            warp_bitonic_sort(sm, cap, lane);
            T = key_val(sm[k - 1]);
            // G's in total, and G's among the first k entries of the (G or E) subsequence in column order
            int g = 0;
            for (int t = lane; t < k; t += 32) g += key_val(sm[t]) < T ? 1 : 0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) g += __shfl_xor_sync(0xffffffffu, g, o);
            int seen = 0, g_first = 0;
            for (int base = 0; base < len && seen < k; base += 32) {
                const int t = base + lane;
                bool ge = false, isg = false;
                if (t < len) {
                    const float v = val[b + t];
                    if (admissible(v)) {
                        const float ov = top ? -v : v;
                        ge = ov <= T;
                        isg = ov < T;
                    }
                }
                const unsigned mge = __ballot_sync(0xffffffffu, ge);
                const int p = seen + __popc(mge & lt_mask) + 1;  // 1-based subsequence position
                g_first += __popc(__ballot_sync(0xffffffffu, isg && p <= k));
                seen += __popc(mge);
            }
            later_g = g - g_first;
        }
        __syncwarp();
        // selection pass in column order
        int seen = 0, e_seen = 0, nout = 0;
        for (int base = 0; base < len; base += 32) {
            const int t = base + lane;
            bool ge = false, isg = false;
            float v = 0.f;
            uint32_t c = 0;
            if (t < len) {
                v = val[b + t];
                c = col[b + t];
                if (admissible(v)) {
                    const float ov = top ? -v : v;
                    ge = all || ov <= T;
                    isg = all || ov < T;
                }
            }
            const unsigned mge = __ballot_sync(0xffffffffu, ge);
            const int p = seen + __popc(mge & lt_mask) + 1;
            const bool e_first = ge && !isg && p <= k;
            const unsigned me = __ballot_sync(0xffffffffu, e_first);
            const int q = e_seen + __popc(me & lt_mask) + 1;  // 1-based rank among the admitted E's
            const bool keep = isg || (e_first && q > later_g);
            const unsigned mk = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int o = nout + __popc(mk & lt_mask);
                out_cols[r * k + o] = c;
                out_vals[r * k + o] = v;
            }
            seen += __popc(mge);
            e_seen += __popc(me);
            nout += __popc(mk);
        }
        for (int t = nout + lane; t < k; t += 32) {
            out_cols[r * k + t] = 0xFFFFFFFFu;
            out_vals[r * k + t] = 0.f;
        }
        if (lane == 0) row_nnz[r] = nout;
        __syncwarp();
    }
}

int find_k_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int max_nnz, int k, int top, uint32_t* d_out_cols,
               float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s) {
    if (n <= 0) return 0;
    int cap = 32;
    while (cap < max_nnz || cap < k) cap <<= 1;
    if (cap > 4096) return fail(SPHB_ERR_UNSUPPORTED, "find_k: rows with more than 4096 entries unsupported");
    const size_t smem = 4 * static_cast<size_t>(cap) * sizeof(uint64_t);
    SPHB_CUDA(cudaFuncSetAttribute(find_k_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    DevBuf<int> err;
    SPHB_CUDA(err.alloc(1, s));
    SPHB_CUDA(cudaMemsetAsync(err.p, 0, sizeof(int), s));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 8);
    find_k_kernel<<<static_cast<unsigned>(blocks), 128, smem, s>>>(n, d_rowptr, d_col, d_val, k, top, cap, d_out_cols, d_out_vals, d_row_nnz, err.p);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    int herr = 0;
    SPHB_CUDA(cudaMemcpyAsync(&herr, err.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (herr) return fail(SPHB_ERR_UNSUPPORTED, "find_k: row longer than the dispatched tile");
    return 0;
}

// ---------------------------------------------------------------------------------------------
// D5
// ---------------------------------------------------------------------------------------------
// keys[(e - comp_ptr[0]) * k + t] = (component << 32) | knn id, one warp per component
__global__ void overlap_expand(int64_t ncomp, const int64_t* __restrict__ comp_ptr, const int64_t* __restrict__ comp_pix,
                               const int64_t* __restrict__ knn_ids, int k, uint64_t* __restrict__ keys) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t c = warp; c < ncomp; c += nwarps) {
        for (int64_t e = comp_ptr[c]; e < comp_ptr[c + 1]; ++e) {
            const int64_t* nb = knn_ids + comp_pix[e] * k;
            for (int t = lane; t < k; t += 32) keys[e * k + t] = (static_cast<uint64_t>(c) << 32) | static_cast<uint32_t>(nb[t]);
        }
    }
}
__global__ void overlap_flag(const uint64_t* __restrict__ keys, int64_t num, int* __restrict__ head) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        head[t] = (t == 0 || keys[t] != keys[t - 1]) ? 1 : 0;
}
__global__ void overlap_compact(const uint64_t* __restrict__ keys, const int* __restrict__ head, const int* __restrict__ pos, int64_t num,
                                uint64_t* __restrict__ ukeys) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        if (head[t]) ukeys[pos[t]] = keys[t];
}
__global__ void overlap_rowptr(const uint64_t* __restrict__ ukeys, int64_t num, int64_t ncomp, int64_t* __restrict__ uptr) {
    for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c <= ncomp; c += static_cast<int64_t>(gridDim.x) * blockDim.x)
        uptr[c] = c == ncomp ? num : lower_bound_u64(ukeys, num, static_cast<uint64_t>(c) << 32);
}
// one warp per pair: the shorter union is scanned, the longer one binary-searched
__global__ void overlap_pairs(const uint64_t* __restrict__ ukeys, const int64_t* __restrict__ uptr, const int64_t* __restrict__ pa,
                              const int64_t* __restrict__ pb, int64_t npairs, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t p = warp; p < npairs; p += nwarps) {
        int64_t a = pa[p], b = pb[p];
        if (uptr[a + 1] - uptr[a] > uptr[b + 1] - uptr[b]) { const int64_t t = a; a = b; b = t; }
        const int64_t ab = uptr[a], ae = uptr[a + 1], bb = uptr[b], be = uptr[b + 1];
        const uint64_t bhi = static_cast<uint64_t>(b) << 32;
        int cnt = 0;
        for (int64_t e = ab + lane; e < ae; e += 32) {
            const uint64_t want = bhi | (ukeys[e] & 0xffffffffull);
            const int64_t f = bb + lower_bound_u64(ukeys + bb, be - bb, want);
            cnt += (f < be && ukeys[f] == want) ? 1 : 0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        if (lane == 0) {
            const int64_t mn = ae - ab;  // the smaller union
            out[p] = mn > 0 ? static_cast<float>(cnt) / static_cast<float>(static_cast<uint64_t>(mn)) : 0.f;
        }
    }
}

// The sorted-unique unions U_c of ALL components: uk = sorted (component << 32 | id) keys, uptr[ncomp + 1].
static int overlap_build_unions(const int64_t* d_knn_ids, int k, int64_t ncomp, const int64_t* d_comp_ptr, const int64_t* d_comp_pix, int64_t npix,
                                DevBuf<uint64_t>& uk, DevBuf<int64_t>& uptr, int* total_out, cudaStream_t s) {
    const int sms = ctx().num_sms;
    const int64_t num = npix * k;
    DevBuf<uint64_t> k0, k1;
    DevBuf<int> head, pos;
    SPHB_CUDA(k0.alloc(num, s));
    SPHB_CUDA(k1.alloc(num, s));
    SPHB_CUDA(head.alloc(num + 1, s));
    SPHB_CUDA(pos.alloc(num + 1, s));
    SPHB_CUDA(uptr.alloc(ncomp + 1, s));
    auto grid = [&](int64_t work) { return static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((work + 255) / 256, static_cast<int64_t>(sms) * 8))); };
    overlap_expand<<<grid(ncomp * 32), 256, 0, s>>>(ncomp, d_comp_ptr, d_comp_pix, d_knn_ids, k, k0.p);
    count_launch();
    // keys only: sort with a dummy payload-free pass (SortKeys)
    {
        size_t bytes = 0;
        const int end_bit = 32 + bits_for(static_cast<uint64_t>(ncomp));
        if (num > 0x7fffffffll) return fail(SPHB_ERR_UNSUPPORTED, "neighbor_overlap: more than 2^31 - 1 (component, id) records");
        SPHB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, bytes, k0.p, k1.p, static_cast<int>(num), 0, end_bit, s));
        DevBuf<uint8_t> tmp;
        SPHB_CUDA(tmp.alloc(bytes, s));
        SPHB_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, bytes, k0.p, k1.p, static_cast<int>(num), 0, end_bit, s));
        count_launch(3);
    }
    overlap_flag<<<grid(num), 256, 0, s>>>(k1.p, num, head.p);
    SPHB_CUDA(cudaMemsetAsync(head.p + num, 0, sizeof(int), s));
    count_launch();
    if (int rc = exclusive_sum<int>(head.p, pos.p, num + 1, s)) return rc;
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, pos.p + num, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    SPHB_CUDA(uk.alloc(std::max(total, 1), s));
    overlap_compact<<<grid(num), 256, 0, s>>>(k1.p, head.p, pos.p, num, uk.p);
    overlap_rowptr<<<grid(ncomp + 1), 256, 0, s>>>(uk.p, total, ncomp, uptr.p);
    count_launch(2);
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));  // k0 / k1 / head / pos are released on return
    *total_out = total;
    return 0;
}

int neighbor_overlap_dev(const int64_t* d_knn_ids, int64_t n, int k, int64_t ncomp, const int64_t* d_comp_ptr, const int64_t* d_comp_pix,
                         int64_t npix, const int64_t* d_pa, const int64_t* d_pb, int64_t npairs, float* d_out, cudaStream_t s) {
    if (npairs <= 0) return 0;
    if (n >= (1ll << 32) || ncomp >= (1ll << 31)) return fail(SPHB_ERR_UNSUPPORTED, "neighbor_overlap: ids must fit 32 bits");
    const int sms = ctx().num_sms;
    if (npix * k == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out, 0, npairs * sizeof(float), s));
        return 0;
    }
    DevBuf<uint64_t> uk;
    DevBuf<int64_t> uptr;
    int total = 0;
    if (int rc = overlap_build_unions(d_knn_ids, k, ncomp, d_comp_ptr, d_comp_pix, npix, uk, uptr, &total, s)) return rc;
    const unsigned grid = static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((npairs * 32 + 255) / 256, static_cast<int64_t>(sms) * 8)));
    overlap_pairs<<<grid, 256, 0, s>>>(uk.p, uptr.p, d_pa, d_pb, npairs, d_out);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// implemented in graph_merge.cu
int level_knn_sets_dev(int64_t nrows, int64_t ncols, const int64_t* d_rowptr, const uint64_t* d_rowkeys, int64_t nnz, int64_t k, int64_t* d_out_ids,
                       float* d_out_dist, cudaStream_t s);

// D7 in the neighbour-overlap metric: exact level kNN over the components' kNN-id unions.
int level_knn_overlap_dev(const int64_t* d_knn_ids, int64_t n, int k_data, int64_t ncomp, const int64_t* d_comp_ptr, const int64_t* d_comp_pix,
                          int64_t npix, int64_t k, int64_t* d_out_ids, float* d_out_dist, cudaStream_t s) {
    if (n >= (1ll << 32) || ncomp >= (1ll << 31)) return fail(SPHB_ERR_UNSUPPORTED, "level_knn_overlap: ids must fit 32 bits");
    DevBuf<uint64_t> uk;
    DevBuf<int64_t> uptr;
    int total = 0;
    if (npix * k_data > 0) {
        if (int rc = overlap_build_unions(d_knn_ids, k_data, ncomp, d_comp_ptr, d_comp_pix, npix, uk, uptr, &total, s)) return rc;
    } else {
        SPHB_CUDA(uptr.alloc(ncomp + 1, s));
        SPHB_CUDA(uk.alloc(1, s));
        SPHB_CUDA(cudaMemsetAsync(uptr.p, 0, (ncomp + 1) * sizeof(int64_t), s));
    }
    return level_knn_sets_dev(ncomp, n, uptr.p, uk.p, total, k, d_out_ids, d_out_dist, s);
}

}  // namespace sphb

using namespace sphb;

extern "C" {

SPHB_API int sphb_find_k(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t k, int top, uint32_t* out_cols,
                         float* out_vals, int32_t* row_nnz) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || !rowptr || k <= 0 || k > 4096 || (n > 0 && (!out_cols || !out_vals || !row_nnz))) return fail(SPHB_ERR_INVALID, "find_k: bad arguments");
    if (n == 0) return SPHB_OK;
    const int64_t nnz = rowptr[n];
    if (nnz > 0 && (!col || !val)) return fail(SPHB_ERR_INVALID, "find_k: null matrix");
    int64_t max_nnz = 0;
    for (int64_t i = 0; i < n; ++i) max_nnz = std::max(max_nnz, rowptr[i + 1] - rowptr[i]);
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp;
    DevBuf<uint32_t> dc, oc;
    DevBuf<float> dv, ov;
    DevBuf<int32_t> on;
    SPHB_CUDA(drp.alloc(n + 1, s));
    SPHB_CUDA(dc.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(dv.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(oc.alloc(n * k, s));
    SPHB_CUDA(ov.alloc(n * k, s));
    SPHB_CUDA(on.alloc(n, s));
    SPHB_CUDA(cudaMemcpyAsync(drp.p, rowptr, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (nnz) {
        SPHB_CUDA(cudaMemcpyAsync(dc.p, col, nnz * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
        SPHB_CUDA(cudaMemcpyAsync(dv.p, val, nnz * sizeof(float), cudaMemcpyHostToDevice, s));
    }
    if (int rc = find_k_dev(n, drp.p, dc.p, dv.p, static_cast<int>(max_nnz), static_cast<int>(k), top, oc.p, ov.p, on.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_cols, oc.p, n * k * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_vals, ov.p, n * k * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(row_nnz, on.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_level_knn_overlap(const int64_t* knn_ids, int64_t n, int k_data, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix,
                                    int64_t k, int64_t* out_ids, float* out_dist) {
    if (int rc = ensure_init()) return rc;
    if (!knn_ids || n <= 0 || k_data <= 0 || ncomp <= 0 || !comp_ptr || comp_ptr[0] != 0 || k <= 0 || !out_ids || !out_dist)
        return fail(SPHB_ERR_INVALID, "level_knn_overlap: bad arguments");
    const int64_t npix = comp_ptr[ncomp];
    if (npix < 0 || (npix > 0 && !comp_pix)) return fail(SPHB_ERR_INVALID, "level_knn_overlap: bad arguments");
    for (int64_t e = 0; e < npix; ++e)
        if (comp_pix[e] < 0 || comp_pix[e] >= n) return fail(SPHB_ERR_INVALID, "level_knn_overlap: pixel id out of range");
    for (int64_t e = 0; e < n * k_data; ++e)
        if (knn_ids[e] < 0 || knn_ids[e] >= n) return fail(SPHB_ERR_INVALID, "level_knn_overlap: neighbour id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> dk, dcp, dpx, oids;
    DevBuf<float> odist;
    const size_t nk = static_cast<size_t>(ncomp) * static_cast<size_t>(k);
    SPHB_CUDA(dk.alloc(n * k_data, s));
    SPHB_CUDA(dcp.alloc(ncomp + 1, s));
    SPHB_CUDA(dpx.alloc(std::max<int64_t>(npix, 1), s));
    SPHB_CUDA(oids.alloc(nk, s));
    SPHB_CUDA(odist.alloc(nk, s));
    SPHB_CUDA(cudaMemcpyAsync(dk.p, knn_ids, n * k_data * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dcp.p, comp_ptr, (ncomp + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (npix) SPHB_CUDA(cudaMemcpyAsync(dpx.p, comp_pix, npix * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (int rc = level_knn_overlap_dev(dk.p, n, k_data, ncomp, dcp.p, dpx.p, npix, k, oids.p, odist.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_ids, oids.p, nk * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_dist, odist.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_neighbor_overlap(const int64_t* knn_ids, int64_t n, int k, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix,
                                   const int64_t* pa, const int64_t* pb, int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!knn_ids || n <= 0 || k <= 0 || ncomp <= 0 || !comp_ptr || !comp_pix || npairs < 0 || (npairs > 0 && (!pa || !pb || !out)))
        return fail(SPHB_ERR_INVALID, "neighbor_overlap: bad arguments");
    if (npairs == 0) return SPHB_OK;
    const int64_t npix = comp_ptr[ncomp];
    for (int64_t e = 0; e < npix; ++e)
        if (comp_pix[e] < 0 || comp_pix[e] >= n) return fail(SPHB_ERR_INVALID, "neighbor_overlap: pixel id out of range");
    for (int64_t p = 0; p < npairs; ++p)
        if (pa[p] < 0 || pa[p] >= ncomp || pb[p] < 0 || pb[p] >= ncomp) return fail(SPHB_ERR_INVALID, "neighbor_overlap: component id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> dk, dcp, dpx, da, db;
    DevBuf<float> dout;
    SPHB_CUDA(dk.alloc(n * k, s));
    SPHB_CUDA(dcp.alloc(ncomp + 1, s));
    SPHB_CUDA(dpx.alloc(std::max<int64_t>(npix, 1), s));
    SPHB_CUDA(da.alloc(npairs, s));
    SPHB_CUDA(db.alloc(npairs, s));
    SPHB_CUDA(dout.alloc(npairs, s));
    SPHB_CUDA(cudaMemcpyAsync(dk.p, knn_ids, n * k * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dcp.p, comp_ptr, (ncomp + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (npix) SPHB_CUDA(cudaMemcpyAsync(dpx.p, comp_pix, npix * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(da.p, pa, npairs * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(db.p, pb, npairs * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    if (int rc = neighbor_overlap_dev(dk.p, n, k, ncomp, dcp.p, dpx.p, npix, da.p, db.p, npairs, dout.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out, dout.p, npairs * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

}  // extern "C"
