// Ragged-graph passes either side of the kNN search and between hierarchy levels.  All are HBM-bound sort / gather
// passes over kNN-sized edge lists (the radix sort is CUB's, library plumbing; expansion, reduction and write-out are ours).
//   F1  sph/utils/GraphUtils.cpp:197-317             symmetrizeGraph (min-merge) + :23-95 ensureClosestPointIsSelf
//   RW3 sph/utils/SparseMatrixAlgorithms.cpp:486-561 mergeGraphNodes, :443-484 mergeNodesDataDistances
//   D9  sph/LevelSimilarities.cpp:346-389            useRandomWalksAsKnnDistances
//   B7  sph/LevelSimilarities.cpp:566-581            zero removal
//   D6  sph/utils/Similarities.cpp:414-483           euclidDistance + sph/utils/Math.cpp:167-172 symmetricHausdorffDistance
// A graph is (rowptr int64[n+1], ids int64[nnz], dist float[nnz]) — utils::Graph's offsets / knnIndices / knnDistances.
#include "../../include/sphb.h"
#include "graph_common.cuh"
#include "knn_common.cuh"

#include <vector>

namespace sphb {

namespace {

inline unsigned gm_grid(int64_t work, int block, int sms, int per_sm = 8) {
    const int64_t need = (work + block - 1) / block;
    return static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>(need, static_cast<int64_t>(sms) * per_sm)));
}

// rowptr[r] = first record whose row id (key >> shift) is >= r, r = 0..nrows.  Records with a row id >= nrows
// (sentinels of dropped entries) are not counted: rowptr[nrows] is the number of kept records.
__global__ void gm_rowptr_lower_bound(const uint64_t* __restrict__ keys, int64_t num, int shift, int64_t nrows, int64_t* __restrict__ rowptr) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r <= nrows; r += static_cast<int64_t>(gridDim.x) * blockDim.x)
        rowptr[r] = lower_bound_u64(keys, num, static_cast<uint64_t>(r) << shift);
}
__global__ void gm_fixed_rowptr(int64_t n, int k, int64_t* __restrict__ rowptr) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r <= n; r += static_cast<int64_t>(gridDim.x) * blockDim.x) rowptr[r] = r * k;
}
__global__ void gm_flag_heads(const uint64_t* __restrict__ keys, int64_t num, int* __restrict__ head) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t <= num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        head[t] = t == num ? 0 : ((t == 0 || keys[t] != keys[t - 1]) ? 1 : 0);
}

// ---------------------------------------------------------------------------------------------
// F1 / RW3 (graph form): records keyed (row, col); value = distance bits, with a tag in bit 32 for the
// transposed record of the symmetrisation.
// ---------------------------------------------------------------------------------------------
__global__ void gm_expand(int64_t n, const int64_t* __restrict__ rowptr, const int64_t* __restrict__ ids, const float* __restrict__ dist,
                          const uint64_t* __restrict__ parents, uint64_t* __restrict__ keys, uint64_t* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const uint64_t row = parents ? parents[r] : static_cast<uint64_t>(r);
        for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) {
            const uint64_t c = static_cast<uint64_t>(ids[e]);
            const uint64_t bits = __float_as_uint(dist[e]);
            if (parents) {
                keys[e] = (row << 32) | (parents[c] & 0xffffffffull);
                vals[e] = bits;
            } else {
                keys[2 * e] = (row << 32) | c;
                vals[2 * e] = bits;
                keys[2 * e + 1] = (c << 32) | row;
                vals[2 * e + 1] = (1ull << 32) | bits;
            }
        }
    }
}
// One thread per run of equal (row, col).  The radix sort is stable and records are emitted in row-major order, so a
// run lists the stored entries in the order the reference visits them.
//   symmetrise: first direct value d_ij and first transposed value d_ji (emplace never overwrites), std::min of the two
//   merge:      keep the smaller (`if (cur > dist) cur = dist`)
// Output: key = (ordered distance bits << 32) | col, value = row.
__global__ void gm_reduce(const uint64_t* __restrict__ keys, const uint64_t* __restrict__ vals, const int* __restrict__ head,
                          const int* __restrict__ pos, int64_t num, int sym, uint64_t* __restrict__ okeys, uint32_t* __restrict__ orow) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        if (!head[t]) continue;
        const uint64_t key = keys[t];
        float r;
        if (sym) {
            bool ha = false, hb = false;
            float a = 0.f, b = 0.f;
            for (int64_t u = t; u < num && keys[u] == key && !(ha && hb); ++u) {
                const uint64_t w = vals[u];
                const float v = __uint_as_float(static_cast<uint32_t>(w));
                if ((w >> 32) == 0) {
                    if (!ha) { a = v; ha = true; }
                } else if (!hb) { b = v; hb = true; }
            }
            r = (ha && hb) ? ((b < a) ? b : a) : (ha ? a : b);
        } else {
            r = __uint_as_float(static_cast<uint32_t>(vals[t]));
            for (int64_t u = t + 1; u < num && keys[u] == key; ++u) {
                const float v = __uint_as_float(static_cast<uint32_t>(vals[u]));
                if (r > v) r = v;
            }
        }
        const int p = pos[t];
        okeys[p] = (static_cast<uint64_t>(f2ord(r)) << 32) | (key & 0xffffffffull);
        orow[p] = static_cast<uint32_t>(key >> 32);
    }
}
__global__ void gm_widen(const uint32_t* __restrict__ in, int64_t num, uint64_t* __restrict__ out) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) out[t] = in[t];
}
// One warp per output row: entries sorted by (distance, index) -> ids / dist with ensureClosestPointIsSelf fused
// (swap / overwrite-last-then-swap / shift-right, sph/utils/GraphUtils.cpp:23-95).
__device__ __forceinline__ void gm_write_row(const uint64_t* ent, int64_t len, int64_t p, int64_t* __restrict__ out_ids,
                                             float* __restrict__ out_dist, int lane) {
    int64_t s = -1;
    bool allzero = true;
    for (int64_t base = 0; base < len; base += 32) {
        const int64_t t = base + lane;
        bool hit = false, nz = false;
        if (t < len) {
            const uint64_t w = ent[t];
            hit = static_cast<int64_t>(w & 0xffffffffull) == p;
            nz = ord2f(static_cast<uint32_t>(w >> 32)) != 0.f;
        }
        const unsigned mh = __ballot_sync(0xffffffffu, hit);
        if (__any_sync(0xffffffffu, nz)) allzero = false;
        if (mh && s < 0) s = base + (__ffs(mh) - 1);
    }
    const int mode = s == 0 ? 0 : (s > 0 ? 1 : (allzero ? 2 : 3));
    for (int64_t t = lane; t < len; t += 32) {
        int64_t src = t;
        if (mode == 1) src = t == 0 ? s : (t == s ? 0 : t);
        else if (mode == 2) src = t == len - 1 ? 0 : t;
        else if (mode == 3) src = t == 0 ? 0 : t - 1;
        const uint64_t w = ent[src];
        int64_t id = static_cast<int64_t>(w & 0xffffffffull);
        float d = ord2f(static_cast<uint32_t>(w >> 32));
        if (t == 0 && mode == 2) {
            id = p;
            d = ord2f(static_cast<uint32_t>(ent[len - 1] >> 32));
        } else if (t == 0 && mode == 3) {
            id = p;
            d = 0.f;
        }
        out_ids[t] = id;
        out_dist[t] = d;
    }
}
__global__ void gm_write_rows(int64_t n, const int64_t* __restrict__ rowptr, const uint64_t* __restrict__ ent, int64_t* __restrict__ out_ids,
                              float* __restrict__ out_dist) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t p = warp; p < n; p += nwarps) {
        const int64_t b = rowptr[p], len = rowptr[p + 1] - b;
        if (len <= 0) continue;
        gm_write_row(ent + b, len, p, out_ids + b, out_dist + b, lane);
    }
}
// Short rows (the common case: <= 2k - 1 entries after a symmetrisation): the unique records come out of the reduction grouped
// by row already, so each warp orders its row by (distance, index) in shared memory and writes it — no further global sort.
constexpr int GM_SORT_WARPS = 4;
__global__ void __launch_bounds__(GM_SORT_WARPS * 32) gm_sort_write_rows(int64_t n, const int64_t* __restrict__ rowptr, const uint64_t* __restrict__ ent,
                                                                         int cap, int64_t* __restrict__ out_ids, float* __restrict__ out_dist) {
    extern __shared__ uint64_t gm_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = gm_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * GM_SORT_WARPS + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * GM_SORT_WARPS;
    for (int64_t p = gw; p < n; p += nw) {
        const int64_t b = rowptr[p], len = rowptr[p + 1] - b;
        if (len <= 0) continue;  // warp-uniform
        for (int t = lane; t < cap; t += 32) sm[t] = t < len ? ent[b + t] : ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        gm_write_row(sm, len, p, out_ids + b, out_dist + b, lane);
        __syncwarp();
    }
}
// row pointers of records grouped by an ascending uint32 row id; max_len receives the longest row
__global__ void gm_rowptr_from_rows(const uint32_t* __restrict__ rows, int64_t num, int64_t nrows, int64_t* __restrict__ rowptr) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r <= nrows; r += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        int64_t lo = 0, hi = num;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (static_cast<int64_t>(rows[mid]) < r) lo = mid + 1;
            else hi = mid;
        }
        rowptr[r] = lo;
    }
}
__global__ void gm_max_row_len(int64_t nrows, const int64_t* __restrict__ rowptr, int* __restrict__ max_len) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r < nrows; r += static_cast<int64_t>(gridDim.x) * blockDim.x)
        atomicMax(max_len, static_cast<int>(min(rowptr[r + 1] - rowptr[r], static_cast<int64_t>(0x7fffffff))));
}

}  // namespace

// mode: parents == nullptr -> symmetrise (n_out = n), else merge into n_out parent nodes.
int graph_minmerge_dev(int64_t n, const int64_t* d_rowptr, const int64_t* d_ids, const float* d_dist, int64_t nnz, const uint64_t* d_parents,
                       int64_t n_out, int64_t* d_out_rowptr, int64_t* d_out_ids, float* d_out_dist, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    if (n >= (1ll << 32) || n_out >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "graph merge: more than 2^32 - 1 nodes");
    const int sms = ctx().num_sms;
    const bool sym = d_parents == nullptr;
    const int64_t num = sym ? 2 * nnz : nnz;
    if (nnz == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (n_out + 1) * sizeof(int64_t), s));
        *out_nnz = 0;
        return 0;
    }
    DevBuf<uint64_t> k0, k1, v0, v1;
    DevBuf<int> head, pos;
    SPHB_CUDA(k0.alloc(num, s));
    SPHB_CUDA(k1.alloc(num, s));
    SPHB_CUDA(v0.alloc(num, s));
    SPHB_CUDA(v1.alloc(num, s));
    SPHB_CUDA(head.alloc(num + 1, s));
    SPHB_CUDA(pos.alloc(num + 1, s));
    gm_expand<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_ids, d_dist, d_parents, k0.p, v0.p);
    count_launch();
    if (int rc = sort_pairs<uint64_t>(k0.p, k1.p, v0.p, v1.p, num, 32 + bits_for(static_cast<uint64_t>(n_out)), s)) return rc;
    gm_flag_heads<<<gm_grid(num + 1, 256, sms), 256, 0, s>>>(k1.p, num, head.p);
    count_launch();
    if (int rc = exclusive_sum<int>(head.p, pos.p, num + 1, s)) return rc;
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, pos.p + num, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (total > cap) return fail(SPHB_ERR_CAPACITY, "graph merge: output capacity too small (2*nnz for symmetrise, nnz for merge always suffices)");
    // unique (row, col, min distance) records, emitted in (row, col) order: grouped by row already
    DevBuf<uint64_t> e0;
    DevBuf<uint32_t> w0;
    DevBuf<int> maxlen;
    SPHB_CUDA(e0.alloc(total, s));
    SPHB_CUDA(w0.alloc(total, s));
    SPHB_CUDA(maxlen.alloc(1, s));
    SPHB_CUDA(cudaMemsetAsync(maxlen.p, 0, sizeof(int), s));
    gm_reduce<<<gm_grid(num, 256, sms), 256, 0, s>>>(k1.p, v1.p, head.p, pos.p, num, sym ? 1 : 0, e0.p, w0.p);
    gm_rowptr_from_rows<<<gm_grid(n_out + 1, 256, sms), 256, 0, s>>>(w0.p, total, n_out, d_out_rowptr);
    gm_max_row_len<<<gm_grid(n_out, 256, sms), 256, 0, s>>>(n_out, d_out_rowptr, maxlen.p);
    count_launch(3);
    int hmax = 0;
    SPHB_CUDA(cudaMemcpyAsync(&hmax, maxlen.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    constexpr int GM_SHORT_ROW = 2048;
    if (hmax <= GM_SHORT_ROW) {
        // per-row order (distance, index) in shared memory, one warp per row
        int cap = 32;
        while (cap < hmax) cap <<= 1;
        const size_t smem = static_cast<size_t>(GM_SORT_WARPS) * cap * sizeof(uint64_t);
        SPHB_CUDA(cudaFuncSetAttribute(gm_sort_write_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        const unsigned grid = static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((n_out + GM_SORT_WARPS - 1) / GM_SORT_WARPS, static_cast<int64_t>(sms) * 8)));
        gm_sort_write_rows<<<grid, GM_SORT_WARPS * 32, smem, s>>>(n_out, d_out_rowptr, e0.p, cap, d_out_ids, d_out_dist);
        count_launch();
    } else {
        // long rows (a merged node that absorbed many children): stable sort by (distance, col), then by row
        DevBuf<uint64_t> e1, r0, r1;
        DevBuf<uint32_t> w1;
        SPHB_CUDA(e1.alloc(total, s));
        SPHB_CUDA(r0.alloc(total, s));
        SPHB_CUDA(r1.alloc(total, s));
        SPHB_CUDA(w1.alloc(total, s));
        if (int rc = sort_pairs<uint32_t>(e0.p, e1.p, w0.p, w1.p, total, 64, s)) return rc;
        gm_widen<<<gm_grid(total, 256, sms), 256, 0, s>>>(w1.p, total, r0.p);
        count_launch();
        if (int rc = sort_pairs<uint64_t>(r0.p, r1.p, e1.p, e0.p, total, bits_for(static_cast<uint64_t>(n_out)), s)) return rc;
        gm_write_rows<<<gm_grid(n_out * 32, 256, sms), 256, 0, s>>>(n_out, d_out_rowptr, e0.p, d_out_ids, d_out_dist);
        count_launch();
        SPHB_CUDA(cudaGetLastError());
        SPHB_CUDA(cudaStreamSynchronize(s));  // e1 / r0 / r1 / w1 are released at the end of this block
    }
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));  // e0 / w0 are released on return
    *out_nnz = total;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// RW3 (sparse form): mergeNodesDataDistances.  Eigen's coeff() reads a missing entry as 0 and the reference treats a
// 0 as "empty": it is overwritten by the next value whatever that is; otherwise the smaller value stays.
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void md_expand(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                          const uint64_t* __restrict__ parents, uint64_t* __restrict__ keys, float* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const uint64_t into = parents[r];
        for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) {
            keys[e] = (into << 32) | (parents[col[e]] & 0xffffffffull);
            vals[e] = val[e];
        }
    }
}
__global__ void md_reduce(const uint64_t* __restrict__ keys, const float* __restrict__ vals, const int* __restrict__ head, const int* __restrict__ pos,
                          int64_t num, uint64_t* __restrict__ ukeys, uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        if (!head[t]) continue;
        const uint64_t key = keys[t];
        float cur = vals[t];
        for (int64_t u = t + 1; u < num && keys[u] == key; ++u) {
            const float v = vals[u];
            if (cur == 0.f) cur = v;
            else if (cur > v) cur = v;
        }
        const int p = pos[t];
        ukeys[p] = key;
        out_col[p] = static_cast<uint32_t>(key);
        out_val[p] = cur;
    }
}
}  // namespace

int merge_data_distances_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, const uint64_t* d_parents,
                             int64_t n_merged, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz,
                             cudaStream_t s) {
    if (n_merged >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "merge_data_distances: n_merged must be < 2^32");
    const int sms = ctx().num_sms;
    if (nnz == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (n_merged + 1) * sizeof(int64_t), s));
        *out_nnz = 0;
        return 0;
    }
    DevBuf<uint64_t> k0, k1, uk;
    DevBuf<float> v0, v1;
    DevBuf<int> head, pos;
    SPHB_CUDA(k0.alloc(nnz, s));
    SPHB_CUDA(k1.alloc(nnz, s));
    SPHB_CUDA(v0.alloc(nnz, s));
    SPHB_CUDA(v1.alloc(nnz, s));
    SPHB_CUDA(head.alloc(nnz + 1, s));
    SPHB_CUDA(pos.alloc(nnz + 1, s));
    md_expand<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, d_parents, k0.p, v0.p);
    count_launch();
    if (int rc = sort_pairs<float>(k0.p, k1.p, v0.p, v1.p, nnz, 32 + bits_for(static_cast<uint64_t>(n_merged)), s)) return rc;
    gm_flag_heads<<<gm_grid(nnz + 1, 256, sms), 256, 0, s>>>(k1.p, nnz, head.p);
    count_launch();
    if (int rc = exclusive_sum<int>(head.p, pos.p, nnz + 1, s)) return rc;
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, pos.p + nnz, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (total > cap) return fail(SPHB_ERR_CAPACITY, "merge_data_distances: output capacity too small (nnz of the input always suffices)");
    SPHB_CUDA(uk.alloc(total, s));
    md_reduce<<<gm_grid(nnz, 256, sms), 256, 0, s>>>(k1.p, v1.p, head.p, pos.p, nnz, uk.p, d_out_col, d_out_val);
    gm_rowptr_lower_bound<<<gm_grid(n_merged + 1, 256, sms), 256, 0, s>>>(uk.p, total, 32, n_merged, d_out_rowptr);
    count_launch(2);
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = total;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// D9: random-walk rows as a distance graph.  distance = 1 - value, exact zeros dropped, rows stable-sorted by distance
// (the LSD radix sort is stable and the records are emitted in column order, so ties keep column order).
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void rwd_expand(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                           uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) {
            const float v = val[e];
            keys[e] = v != 0.f ? ((static_cast<uint64_t>(r) << 32) | f2ord(__fsub_rn(1.f, v))) : ~0ull;
            vals[e] = col[e];
        }
    }
}
__global__ void rwd_write(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t num, int64_t* __restrict__ out_ids,
                          float* __restrict__ out_dist) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        out_ids[t] = static_cast<int64_t>(vals[t]);
        out_dist[t] = ord2f(static_cast<uint32_t>(keys[t]));
    }
}
}  // namespace

int rw_as_knn_distances_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t* d_out_rowptr,
                            int64_t* d_out_ids, float* d_out_dist, int64_t* out_nnz, cudaStream_t s) {
    if (n >= (1ll << 32) - 1) return fail(SPHB_ERR_UNSUPPORTED, "rw_as_knn_distances: n must be < 2^32 - 1");
    const int sms = ctx().num_sms;
    if (nnz == 0) {
        SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (n + 1) * sizeof(int64_t), s));
        *out_nnz = 0;
        return 0;
    }
    DevBuf<uint64_t> k0, k1;
    DevBuf<uint32_t> v0, v1;
    SPHB_CUDA(k0.alloc(nnz, s));
    SPHB_CUDA(k1.alloc(nnz, s));
    SPHB_CUDA(v0.alloc(nnz, s));
    SPHB_CUDA(v1.alloc(nnz, s));
    rwd_expand<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, k0.p, v0.p);
    count_launch();
    if (int rc = sort_pairs<uint32_t>(k0.p, k1.p, v0.p, v1.p, nnz, 64, s)) return rc;
    gm_rowptr_lower_bound<<<gm_grid(n + 1, 256, sms), 256, 0, s>>>(k1.p, nnz, 32, n, d_out_rowptr);
    count_launch();
    int64_t kept = 0;
    SPHB_CUDA(cudaMemcpyAsync(&kept, d_out_rowptr + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (kept > 0) {
        rwd_write<<<gm_grid(kept, 256, sms), 256, 0, s>>>(k1.p, v1.p, kept, d_out_ids, d_out_dist);
        count_launch();
    }
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = kept;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// B7: erase exact zeros from every CSR row (order preserved).
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void rz_count(int64_t n, const int64_t* __restrict__ rowptr, const float* __restrict__ val, int64_t* __restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r <= n; r += nwarps) {
        int c = 0;
        if (r < n)
            for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) c += val[e] != 0.f ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) cnt[r] = c;
    }
}
__global__ void rz_scatter(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                           const int64_t* __restrict__ out_rowptr, uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        int64_t out = out_rowptr[r];
        const int64_t end = rowptr[r + 1];
        for (int64_t base = rowptr[r]; base < end; base += 32) {
            const int64_t e = base + lane;
            const bool keep = e < end && val[e] != 0.f;
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int64_t o = out + __popc(mask & ((1u << lane) - 1u));
                out_col[o] = col[e];
                out_val[o] = val[e];
            }
            out += __popc(mask);
        }
    }
}
}  // namespace

int remove_zeros_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t* d_out_rowptr, uint32_t* d_out_col,
                     float* d_out_val, int64_t* out_nnz, cudaStream_t s) {
    const int sms = ctx().num_sms;
    DevBuf<int64_t> cnt;
    SPHB_CUDA(cnt.alloc(n + 1, s));
    rz_count<<<gm_grid((n + 1) * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_val, cnt.p);
    count_launch();
    if (int rc = exclusive_sum<int64_t>(cnt.p, d_out_rowptr, n + 1, s)) return rc;
    int64_t kept = 0;
    SPHB_CUDA(cudaMemcpyAsync(&kept, d_out_rowptr + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (kept > 0) {
        rz_scatter<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, d_out_rowptr, d_out_col, d_out_val);
        count_launch();
    }
    SPHB_CUDA(cudaGetLastError());
    *out_nnz = kept;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// RW4: the post-merge step of Hierarchy::updateRandomWalks (sph/utils/Hierarchy.cpp:287-307):
//   removeDiagonalElements(keepSingleEntry = true) (sph/utils/SparseMatrixAlgorithms.cpp:704-718) — a row with more than one
//   entry loses its diagonal AND every stored exact zero (removeElement: coeffRef(i) = 0, then prune(0, eps)), then
//   ONEDIM: every row / its fp32 storage-order sum (normalizeUnitSparseVector :666-673)
//   TWODIM: every value / the fp32 sum of the row sums taken in row order (normalizeSparseMatrix :617-628).
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void rd_count(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val, int remove,
                         int64_t* __restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r <= n; r += nwarps) {
        int c = 0;
        if (r < n) {
            const int64_t b = rowptr[r], e1 = rowptr[r + 1];
            const bool strip = remove && e1 - b > 1;
            for (int64_t e = b + lane; e < e1; e += 32) c += (strip && (static_cast<int64_t>(col[e]) == r || val[e] == 0.f)) ? 0 : 1;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) cnt[r] = c;
    }
}
__global__ void rd_scatter(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val, int remove,
                           const int64_t* __restrict__ out_rowptr, uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        int64_t out = out_rowptr[r];
        const int64_t b = rowptr[r], end = rowptr[r + 1];
        const bool strip = remove && end - b > 1;
        for (int64_t base = b; base < end; base += 32) {
            const int64_t e = base + lane;
            const bool keep = e < end && !(strip && (static_cast<int64_t>(col[e]) == r || val[e] == 0.f));
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int64_t o = out + __popc(mask & ((1u << lane) - 1u));
                out_col[o] = col[e];
                out_val[o] = val[e];
            }
            out += __popc(mask);
        }
    }
}
// fp32 storage-order sum of every row (Eigen SparseVector::sum()); ONEDIM divides in place
__global__ void rd_row_sums(int64_t n, const int64_t* __restrict__ rowptr, float* __restrict__ val, int divide, float* __restrict__ sums) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        const int64_t b = rowptr[r], e1 = rowptr[r + 1];
        float sacc = 0.f;
        if (lane == 0)
            for (int64_t e = b; e < e1; ++e) sacc += val[e];
        sacc = __shfl_sync(0xffffffffu, sacc, 0);
        if (sums && lane == 0) sums[r] = sacc;
        if (divide)
            for (int64_t e = b + lane; e < e1; e += 32) val[e] = val[e] / sacc;
    }
}
__global__ void rd_total(int64_t n, const float* __restrict__ sums, float* __restrict__ total) {  // one thread: the reference's row-order fp32 sum
    float t = 0.f;
    for (int64_t r = 0; r < n; ++r) t += sums[r];
    *total = t;
}
__global__ void rd_scale(int64_t nnz, float* __restrict__ val, const float* __restrict__ total) {
    const float t = *total;
    for (int64_t e = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < nnz; e += static_cast<int64_t>(gridDim.x) * blockDim.x) val[e] = val[e] / t;
}
}  // namespace

// norm: 0 none, 1 ONEDIM, 2 TWODIM
int rw_post_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int remove_diagonal, int norm,
                      int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t* out_nnz, cudaStream_t s) {
    const int sms = ctx().num_sms;
    DevBuf<int64_t> cnt;
    SPHB_CUDA(cnt.alloc(n + 1, s));
    rd_count<<<gm_grid((n + 1) * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, remove_diagonal, cnt.p);
    count_launch();
    if (int rc = exclusive_sum<int64_t>(cnt.p, d_out_rowptr, n + 1, s)) return rc;
    int64_t kept = 0;
    SPHB_CUDA(cudaMemcpyAsync(&kept, d_out_rowptr + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = kept;
    if (kept == 0) return 0;
    rd_scatter<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, remove_diagonal, d_out_rowptr, d_out_col, d_out_val);
    count_launch();
    if (norm == 1) {
        rd_row_sums<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_out_rowptr, d_out_val, 1, nullptr);
        count_launch();
    } else if (norm == 2) {
        DevBuf<float> sums, total;
        SPHB_CUDA(sums.alloc(n, s));
        SPHB_CUDA(total.alloc(1, s));
        rd_row_sums<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_out_rowptr, d_out_val, 0, sums.p);
        rd_total<<<1, 1, 0, s>>>(n, sums.p, total.p);
        rd_scale<<<gm_grid(kept, 256, sms), 256, 0, s>>>(kept, d_out_val, total.p);
        count_launch(3);
        SPHB_CUDA(cudaStreamSynchronize(s));  // sums / total are released on return
    }
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// D6: symmetric Hausdorff distance between the pixel sets of two components, one CTA per pair.
// Every pixel-pair distance is the Stage-A contract distance (fp64 FMA chain, one rounding, sqrtf), so a pair of
// single-pixel components is the plain L2 distance of the level-0 branch.  The distance matrix is never stored: the
// row minima and the column minima are two sweeps over the pairs (max_i min_j, then max_j min_i).
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int EH_WARPS = 4;
__device__ __forceinline__ float eh_directed(const float* __restrict__ x, int d, const int64_t* __restrict__ pix, int64_t a0, int64_t a1, int64_t b0,
                                             int64_t b1, int warp, int lane) {
    float best = -INFINITY;
    for (int64_t i = a0 + warp; i < a1; i += EH_WARPS) {
        const int64_t pi = pix[i];
        float m = INFINITY;
        for (int64_t j = b0 + lane; j < b1; j += 32) m = fminf(m, sqrtf(exact_key(x, pi, pix[j], d, METRIC_L2)));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
        best = fmaxf(best, m);
    }
    return best;
}
__global__ void __launch_bounds__(EH_WARPS * 32) euclid_hausdorff_kernel(const float* __restrict__ x, int d, const int64_t* __restrict__ comp_ptr,
                                                                         const int64_t* __restrict__ comp_pix, const int64_t* __restrict__ pa,
                                                                         const int64_t* __restrict__ pb, int64_t npairs, float* __restrict__ out) {
    __shared__ float s_best[EH_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t p = blockIdx.x; p < npairs; p += gridDim.x) {
        const int64_t a0 = comp_ptr[pa[p]], a1 = comp_ptr[pa[p] + 1], b0 = comp_ptr[pb[p]], b1 = comp_ptr[pb[p] + 1];
        float best = 0.f;
        if (a1 > a0 && b1 > b0) {
            best = eh_directed(x, d, comp_pix, a0, a1, b0, b1, warp, lane);
            best = fmaxf(best, eh_directed(x, d, comp_pix, b0, b1, a0, a1, warp, lane));
        }
        if (lane == 0) s_best[warp] = best;
        __syncthreads();
        if (threadIdx.x == 0) {
            float m = s_best[0];
            for (int w = 1; w < EH_WARPS; ++w) m = fmaxf(m, s_best[w]);
            out[p] = (a1 > a0 && b1 > b0) ? m : 0.f;
        }
        __syncthreads();
    }
}
}  // namespace

int euclid_hausdorff_dev(const float* d_x, int d, const int64_t* d_comp_ptr, const int64_t* d_comp_pix, const int64_t* d_pa, const int64_t* d_pb,
                         int64_t npairs, float* d_out, cudaStream_t s) {
    if (npairs <= 0) return 0;
    const unsigned grid = static_cast<unsigned>(std::min<int64_t>(npairs, static_cast<int64_t>(ctx().num_sms) * 16));
    euclid_hausdorff_kernel<<<grid, EH_WARPS * 32, 0, s>>>(d_x, d, d_comp_ptr, d_comp_pix, d_pa, d_pb, npairs, d_out);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

// D7 in the Euclid metric (computeExactKnn with EuclidDistSpace, sph/LevelSimilarities.cpp:211-252, :414;
// sph/utils/EuclidDistSpace.hpp:29-39): every component against every component through the Hausdorff kernel above, row block
// by row block, and the k smallest (distance, index) of every row picked on the device — no host round trip per block.
__global__ void eh_block_pairs(int64_t r0, int64_t rows, int64_t ncomp, int64_t* __restrict__ pa, int64_t* __restrict__ pb) {
    const int64_t total = rows * ncomp;
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        pa[t] = r0 + t / ncomp;
        pb[t] = t % ncomp;
    }
}

constexpr int BK_WARPS = 4;
__global__ void __launch_bounds__(BK_WARPS * 32) bottomk_rows_kernel(const float* __restrict__ dmat, int64_t rows, int64_t ncols, int k, int cap,
                                                                     int64_t* __restrict__ ids, float* __restrict__ dist) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * BK_WARPS + warp; r < rows; r += static_cast<int64_t>(gridDim.x) * BK_WARPS) {
        const float* row = dmat + r * ncols;
        int cnt = 0;
        uint64_t kth = ~0ull;
        for (int64_t base = 0; base < ncols; base += 32) {
            const int64_t j = base + lane;
            const uint64_t key = j < ncols ? make_key(row[j], static_cast<uint32_t>(j)) : ~0ull;
            const bool pass = j < ncols && key <= kth;
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(sm, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? sm[k - 1] : ~0ull;
                }
            }
        }
        for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        for (int t = lane; t < k; t += 32) {
            const uint64_t key = sm[t];
            ids[r * k + t] = key == ~0ull ? -1 : static_cast<int64_t>(key_idx(key));
            dist[r * k + t] = key == ~0ull ? -1.0f : key_val(key);
        }
        __syncwarp();
    }
}

int level_knn_euclid_dev(const float* d_x, int d, int64_t ncomp, const int64_t* d_comp_ptr, const int64_t* d_comp_pix, int k, int64_t* d_ids,
                         float* d_dist, cudaStream_t s) {
    if (ncomp <= 0) return 0;
    int cap = 64;
    while (cap < k + 32) cap <<= 1;
    const size_t smem = static_cast<size_t>(BK_WARPS) * cap * sizeof(uint64_t);
    if (smem > 200 * 1024) return fail(SPHB_ERR_UNSUPPORTED, "level_knn_euclid: k too large");
    SPHB_CUDA(cudaFuncSetAttribute(bottomk_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t block = std::max<int64_t>(1, std::min<int64_t>(ncomp, (int64_t{1} << 24) / ncomp));  // <= 16 M pairs per block
    DevBuf<int64_t> pa, pb;
    DevBuf<float> dm;
    SPHB_CUDA(pa.alloc(block * ncomp, s));
    SPHB_CUDA(pb.alloc(block * ncomp, s));
    SPHB_CUDA(dm.alloc(block * ncomp, s));
    const int sms = ctx().num_sms;
    for (int64_t r0 = 0; r0 < ncomp; r0 += block) {
        const int64_t rows = std::min<int64_t>(block, ncomp - r0);
        eh_block_pairs<<<gm_grid(rows * ncomp, 256, sms), 256, 0, s>>>(r0, rows, ncomp, pa.p, pb.p);
        count_launch();
        if (int rc = euclid_hausdorff_dev(d_x, d, d_comp_ptr, d_comp_pix, pa.p, pb.p, rows * ncomp, dm.p, s)) return rc;
        bottomk_rows_kernel<<<static_cast<unsigned>(std::min<int64_t>((rows + BK_WARPS - 1) / BK_WARPS, static_cast<int64_t>(sms) * 16)), BK_WARPS * 32, smem,
                              s>>>(dm.p, rows, ncomp, k, cap, d_ids + r0 * k, d_dist + r0 * k);
        count_launch();
    }
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// D7: exact level kNN in a random-walk component metric (LevelSimilarities.cpp:211-252 with
// NeighborWalksBhattacharyyaSpace / NeighborWalksSingleOverlapSpace).  The spaces' distance functions do NOT go through
// componentDistance: d(i, j) = 1 - sim(i, j) for every j, j == i included
//   mode 0  sim = sum over the columns c of row i, in column order, of sqrt(A[i][c] * A[j][c]) where A[j][c] > 0
//           (sph/utils/NeighborWalksBhattacharyyaSpace.hpp:28-41, Similarities.cpp:379-396)
//   mode 1  sim = A[i][j], the stored entry of ROW i only (NeighborWalksSingleOverlapSpace.hpp:26-36, Similarities.cpp:374-377)
// and per row the k smallest by (distance, index) (the reference's std::sort leaves ties unordered).
// The reference evaluates all n² pairs; here row i is a Gustavson expansion through the transposed matrix: one CTA per
// row walks the entries of row i IN ORDER (a barrier between columns), the threads of the CTA take the rows j of that
// column, so every acc[j] is summed in exactly the order of the per-pair loop — bit-exact, no atomics on values.
// Rows that share no column with i sit at distance exactly 1 and are never touched: they enter the result only when
// fewer than k touched rows are closer, smallest index first.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int LK_WARPS = 8;
__global__ void lk_expand(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                          uint64_t* __restrict__ tkeys, float* __restrict__ tvals) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps)
        for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) {
            tkeys[e] = (static_cast<uint64_t>(col[e]) << 32) | static_cast<uint64_t>(r);
            tvals[e] = val[e];
        }
}
__global__ void lk_swap_halves(const uint64_t* __restrict__ in, int64_t num, uint64_t* __restrict__ out) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        out[t] = (in[t] << 32) | (in[t] >> 32);
}
__device__ __forceinline__ void lk_block_sort(uint64_t* sm, int cap) {
    for (int k2 = 2; k2 <= cap; k2 <<= 1) {
        for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
            for (int t = threadIdx.x; t < (cap >> 1); t += blockDim.x) {
                const int i = ((t & ~(j2 - 1)) << 1) | (t & (j2 - 1));
                const int l = i | j2;
                const uint64_t a = sm[i], b = sm[l];
                const bool up = (i & k2) == 0;
                if ((a > b) == up) {
                    sm[i] = b;
                    sm[l] = a;
                }
            }
            __syncthreads();
        }
    }
}
struct LkParams {
    int64_t n;
    const int64_t* rowptr;
    const uint32_t* col;
    const float* val;
    const uint64_t* rowkeys; // mode 2: the rows as sorted (row << 32 | column) keys instead of (col, val)
    const int64_t* colptr;   // transpose: column c holds tkeys/tval[colptr[c] .. colptr[c+1]), row = low 32 bits of the key
    const uint64_t* tkeys;
    const float* tval;
    int mode, k, cap;        // cap: per-warp list capacity, power of two >= k + 32
    float* acc;              // [grid][n], zero on entry and on exit
    uint32_t* touched;       // [grid][n]
    int64_t* out_ids;        // [n][k]
    float* out_dist;         // [n][k]
};
__global__ void __launch_bounds__(LK_WARPS * 32) level_knn_walks_kernel(const LkParams P) {
    extern __shared__ uint64_t lk_keys[];  // LK_WARPS * cap
    __shared__ unsigned int s_touched, s_nsel, s_nlt;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* acc = P.acc + static_cast<size_t>(blockIdx.x) * P.n;
    uint32_t* touched = P.touched + static_cast<size_t>(blockIdx.x) * P.n;
    uint64_t* mylist = lk_keys + static_cast<size_t>(warp) * P.cap;
    const int k = P.k, cap = P.cap;
    const int kk = static_cast<int>(min(static_cast<int64_t>(k), P.n));
    const uint64_t key_one = make_key(1.f, 0u);  // smallest key at distance exactly 1
    for (int64_t i = blockIdx.x; i < P.n; i += gridDim.x) {
        if (threadIdx.x == 0) { s_touched = 0; s_nsel = 0; s_nlt = 0; }
        __syncthreads();
        const int64_t rb = P.rowptr[i], re = P.rowptr[i + 1];
        // ---- 1. accumulate sim(i, j) for every j that shares a column with i
        if (P.mode == 0) {
            for (int64_t e = rb; e < re; ++e) {
                const uint32_t c = P.col[e];
                const float a = P.val[e];
                const int64_t cb = P.colptr[c], ce = P.colptr[c + 1];
                for (int64_t f = cb + threadIdx.x; f < ce; f += blockDim.x) {
                    const float b = P.tval[f];
                    if (b > 0.f) {
                        const float term = sqrtf(__fmul_rn(a, b));
                        if (term != 0.f) {
                            const uint32_t j = static_cast<uint32_t>(P.tkeys[f]);
                            const float old = acc[j];
                            acc[j] = __fadd_rn(old, term);
                            if (old == 0.f) touched[atomicAdd(&s_touched, 1u)] = j;
                        }
                    }
                }
                __syncthreads();  // the next column may reach the same j from another thread
            }
        } else if (P.mode == 1) {
            for (int64_t e = rb + threadIdx.x; e < re; e += blockDim.x) {
                const float v = P.val[e];
                if (v != 0.f) {
                    const uint32_t j = P.col[e];
                    acc[j] = v;
                    touched[atomicAdd(&s_touched, 1u)] = j;
                }
            }
            __syncthreads();
        } else {
            // set intersection sizes: integer counts, the order of the additions is irrelevant (exact in fp32 below 2^24)
            for (int64_t e = rb + warp; e < re; e += LK_WARPS) {
                const uint32_t c = static_cast<uint32_t>(P.rowkeys[e]);
                const int64_t cb = P.colptr[c], ce = P.colptr[c + 1];
                for (int64_t f = cb + lane; f < ce; f += 32) {
                    const uint32_t j = static_cast<uint32_t>(P.tkeys[f]);
                    const float old = atomicAdd(acc + j, 1.f);
                    if (old == 0.f) touched[atomicAdd(&s_touched, 1u)] = j;
                }
            }
            __syncthreads();
        }
        const unsigned int T = s_touched;
        // ---- 2. per-warp k smallest of (distance, index) over the touched rows
        int cnt = 0;
        uint64_t kth = ~0ull;
        for (unsigned int base = warp * 32; base < T; base += LK_WARPS * 32) {
            const unsigned int t = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (t < T) {
                const uint32_t j = touched[t];
                float sim = acc[j];
                if (P.mode == 2) {  // card(U_i ∩ U_j) / min(card U_i, card U_j)  (Similarities.cpp:216-228)
                    const int64_t lj = P.rowptr[j + 1] - P.rowptr[j];
                    sim = __fdiv_rn(sim, static_cast<float>(static_cast<uint64_t>(min(re - rb, lj))));
                }
                const float dv = __fsub_rn(1.f, sim);
                key = make_key(dv, j);
                pass = key <= kth;
            }
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) mylist[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int u = cnt + lane; u < cap; u += 32) mylist[u] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(mylist, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? mylist[k - 1] : ~0ull;
                }
            }
        }
        for (int u = cnt + lane; u < cap; u += 32) mylist[u] = ~0ull;
        __syncthreads();
        // ---- 3. merge the warps' lists; the first k keys are the touched part of the row
        lk_block_sort(lk_keys, LK_WARPS * cap);
        {
            int m = 0, lt = 0;
            for (int t = threadIdx.x; t < k; t += blockDim.x) {
                const uint64_t key = lk_keys[t];
                m += key != ~0ull ? 1 : 0;
                lt += key < key_one ? 1 : 0;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                m += __shfl_xor_sync(0xffffffffu, m, o);
                lt += __shfl_xor_sync(0xffffffffu, lt, o);
            }
            if (lane == 0 && m) atomicAdd(&s_nsel, static_cast<unsigned int>(m));
            if (lane == 0 && lt) atomicAdd(&s_nlt, static_cast<unsigned int>(lt));
        }
        __syncthreads();
        const int nsel = static_cast<int>(s_nsel), nlt = static_cast<int>(s_nlt);
        // entries closer than 1 are final
        for (int t = threadIdx.x; t < nlt; t += blockDim.x) {
            const uint64_t key = lk_keys[t];
            P.out_ids[i * k + t] = static_cast<int64_t>(key_idx(key));
            P.out_dist[i * k + t] = key_val(key);
        }
        // the rest: merge of the touched keys at distance >= 1 with the untouched rows (distance exactly 1, ascending index)
        if (threadIdx.x == 0) {
            int o = nlt, li = nlt;
            int64_t j = 0;
            while (o < kk) {
                while (j < P.n && acc[j] != 0.f) ++j;
                const uint64_t key_u = j < P.n ? make_key(1.f, static_cast<uint32_t>(j)) : ~0ull;
                const uint64_t key_l = li < nsel ? lk_keys[li] : ~0ull;
                if (key_l == ~0ull && key_u == ~0ull) break;
                uint64_t key;
                if (key_l <= key_u) { key = key_l; ++li; }
                else { key = key_u; ++j; }
                P.out_ids[i * k + o] = static_cast<int64_t>(key_idx(key));
                P.out_dist[i * k + o] = key_val(key);
                ++o;
            }
            for (; o < k; ++o) {  // n < k: the caller's "missing neighbour" convention
                P.out_ids[i * k + o] = -1;
                P.out_dist[i * k + o] = -1.f;
            }
        }
        __syncthreads();
        // ---- 4. reset the accumulator where it was touched
        for (unsigned int t = threadIdx.x; t < T; t += blockDim.x) acc[touched[t]] = 0.f;
        __syncthreads();
    }
}
}  // namespace

int level_knn_walks_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int mode, int64_t k,
                        int64_t* d_out_ids, float* d_out_dist, cudaStream_t s) {
    if (n >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "level_knn: n must be < 2^32");
    if (k > 992) return fail(SPHB_ERR_UNSUPPORTED, "level_knn: k > 992 unsupported");
    const int sms = ctx().num_sms;
    DevBuf<int64_t> colptr;
    DevBuf<uint64_t> tk0, tk1;
    DevBuf<float> tv0, tv1;
    SPHB_CUDA(colptr.alloc(n + 1, s));
    SPHB_CUDA(tk0.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(tk1.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(tv0.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(tv1.alloc(std::max<int64_t>(nnz, 1), s));
    if (nnz > 0) {
        lk_expand<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_col, d_val, tk0.p, tv0.p);
        count_launch();
        if (int rc = sort_pairs<float>(tk0.p, tk1.p, tv0.p, tv1.p, nnz, 32 + bits_for(static_cast<uint64_t>(n)), s)) return rc;
    }
    gm_rowptr_lower_bound<<<gm_grid(n + 1, 256, sms), 256, 0, s>>>(tk1.p, nnz, 32, n, colptr.p);
    count_launch();
    int cap = 64;
    while (cap < k + 32) cap <<= 1;
    const size_t smem = static_cast<size_t>(LK_WARPS) * cap * sizeof(uint64_t);
    SPHB_CUDA(cudaFuncSetAttribute(level_knn_walks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int64_t grid = std::min<int64_t>(n, static_cast<int64_t>(sms) * 2);
    while (grid > 1 && static_cast<double>(grid) * n * 8.0 > 8.0 * (1ull << 30)) grid /= 2;  // accumulator slabs <= 8 GiB
    DevBuf<float> acc;
    DevBuf<uint32_t> touched;
    SPHB_CUDA(acc.alloc(static_cast<size_t>(grid) * n, s));
    SPHB_CUDA(touched.alloc(static_cast<size_t>(grid) * n, s));
    SPHB_CUDA(cudaMemsetAsync(acc.p, 0, static_cast<size_t>(grid) * n * sizeof(float), s));
    LkParams P{};
    P.n = n; P.rowptr = d_rowptr; P.col = d_col; P.val = d_val; P.colptr = colptr.p; P.tkeys = tk1.p; P.tval = tv1.p;
    P.mode = mode; P.k = static_cast<int>(k); P.cap = cap; P.acc = acc.p; P.touched = touched.p; P.out_ids = d_out_ids; P.out_dist = d_out_dist;
    level_knn_walks_kernel<<<static_cast<unsigned>(grid), LK_WARPS * 32, smem, s>>>(P);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// D7 in the neighbour-overlap metric (NeighborOverlapSpace, sph/utils/NeighborOverlapSpace.hpp:31-43): the rows are the
// sorted-unique unions U_c as (component << 32 | data id) keys (built by level_ops.cu for D5); distance 1 - |U_i ∩ U_j| /
// min(|U_i|, |U_j|), per row the k smallest by (distance, index).
int level_knn_sets_dev(int64_t nrows, int64_t ncols, const int64_t* d_rowptr, const uint64_t* d_rowkeys, int64_t nnz, int64_t k, int64_t* d_out_ids,
                       float* d_out_dist, cudaStream_t s) {
    if (nrows >= (1ll << 32) || ncols >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "level_knn: ids must fit 32 bits");
    if (k > 992) return fail(SPHB_ERR_UNSUPPORTED, "level_knn: k > 992 unsupported");
    if (nnz > 0x7fffffffll) return fail(SPHB_ERR_UNSUPPORTED, "level_knn: more than 2^31 - 1 set elements");
    const int sms = ctx().num_sms;
    DevBuf<int64_t> colptr;
    DevBuf<uint64_t> tk0, tk1;
    SPHB_CUDA(colptr.alloc(ncols + 1, s));
    SPHB_CUDA(tk0.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(tk1.alloc(std::max<int64_t>(nnz, 1), s));
    if (nnz > 0) {
        lk_swap_halves<<<gm_grid(nnz, 256, sms), 256, 0, s>>>(d_rowkeys, nnz, tk0.p);
        count_launch();
        size_t bytes = 0;
        const int end_bit = 32 + bits_for(static_cast<uint64_t>(ncols));
        SPHB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, bytes, tk0.p, tk1.p, static_cast<int>(nnz), 0, end_bit, s));
        DevBuf<uint8_t> tmp;
        SPHB_CUDA(tmp.alloc(bytes, s));
        SPHB_CUDA(cub::DeviceRadixSort::SortKeys(tmp.p, bytes, tk0.p, tk1.p, static_cast<int>(nnz), 0, end_bit, s));
        count_launch(3);
    }
    gm_rowptr_lower_bound<<<gm_grid(ncols + 1, 256, sms), 256, 0, s>>>(tk1.p, nnz, 32, ncols, colptr.p);
    count_launch();
    int cap = 64;
    while (cap < k + 32) cap <<= 1;
    const size_t smem = static_cast<size_t>(LK_WARPS) * cap * sizeof(uint64_t);
    SPHB_CUDA(cudaFuncSetAttribute(level_knn_walks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    int64_t grid = std::min<int64_t>(nrows, static_cast<int64_t>(sms) * 2);
    while (grid > 1 && static_cast<double>(grid) * nrows * 8.0 > 8.0 * (1ull << 30)) grid /= 2;
    DevBuf<float> acc;
    DevBuf<uint32_t> touched;
    SPHB_CUDA(acc.alloc(static_cast<size_t>(grid) * nrows, s));
    SPHB_CUDA(touched.alloc(static_cast<size_t>(grid) * nrows, s));
    SPHB_CUDA(cudaMemsetAsync(acc.p, 0, static_cast<size_t>(grid) * nrows * sizeof(float), s));
    LkParams P{};
    P.n = nrows; P.rowptr = d_rowptr; P.rowkeys = d_rowkeys; P.colptr = colptr.p; P.tkeys = tk1.p;
    P.mode = 2; P.k = static_cast<int>(k); P.cap = cap; P.acc = acc.p; P.touched = touched.p; P.out_ids = d_out_ids; P.out_dist = d_out_dist;
    level_knn_walks_kernel<<<static_cast<unsigned>(grid), LK_WARPS * 32, smem, s>>>(P);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// F2: connected components of the (symmetrised or plain) kNN graph — utils::labelGraphWeakComponents /
// labelGraphStrongComponents on a symmetric graph (sph/utils/GraphUtils.cpp:371-423), the labelling behind
// NearestNeighbors::computeConnectedComponents (sph/NearestNeighbors.cpp:318-409).  Edges are taken as undirected.
// Min-label hooking + pointer jumping: label[x] <= x always (a forest of decreasing pointers), every edge hooks the
// larger of its two labels under the smaller (atomicMin), jumping flattens the trees, repeat until no edge joins two
// labels.  The root of a component is its lowest vertex; components are numbered in ascending order of that vertex,
// which is the order boost::strong_components reaches them on a symmetric graph.
// ---------------------------------------------------------------------------------------------
namespace {
__global__ void cc_init(int64_t n, unsigned int* __restrict__ label) {
    for (int64_t v = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < n; v += static_cast<int64_t>(gridDim.x) * blockDim.x)
        label[v] = static_cast<unsigned int>(v);
}
__global__ void cc_hook(int64_t n, const int64_t* __restrict__ rowptr, const int64_t* __restrict__ ids, unsigned int* label, int* __restrict__ changed) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t v = warp; v < n; v += nwarps) {
        for (int64_t e = rowptr[v] + lane; e < rowptr[v + 1]; e += 32) {
            const unsigned int a = label[v], b = label[ids[e]];
            if (a != b) {
                atomicMin(label + max(a, b), min(a, b));
                *changed = 1;
            }
        }
    }
}
__global__ void cc_jump(int64_t n, unsigned int* label) {
    for (int64_t v = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < n; v += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        unsigned int l = label[v];
        while (true) {
            const unsigned int ll = label[l];
            if (ll == l) break;
            l = ll;
        }
        label[v] = l;
    }
}
__global__ void cc_flag_roots(int64_t n, const unsigned int* __restrict__ label, int* __restrict__ root) {
    for (int64_t v = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v <= n; v += static_cast<int64_t>(gridDim.x) * blockDim.x)
        root[v] = (v < n && label[v] == static_cast<unsigned int>(v)) ? 1 : 0;
}
__global__ void cc_number(int64_t n, const unsigned int* __restrict__ label, const int* __restrict__ rank, int64_t* __restrict__ out) {
    for (int64_t v = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < n; v += static_cast<int64_t>(gridDim.x) * blockDim.x)
        out[v] = rank[label[v]];
}
}  // namespace

int weak_components_dev(int64_t n, const int64_t* d_rowptr, const int64_t* d_ids, int64_t* d_labels, int64_t* num_components, cudaStream_t s) {
    if (n >= (1ll << 31)) return fail(SPHB_ERR_UNSUPPORTED, "weak_components: n must be < 2^31");
    const int sms = ctx().num_sms;
    DevBuf<unsigned int> label;
    DevBuf<int> changed, root, rank;
    SPHB_CUDA(label.alloc(n, s));
    SPHB_CUDA(changed.alloc(1, s));
    SPHB_CUDA(root.alloc(n + 1, s));
    SPHB_CUDA(rank.alloc(n + 1, s));
    cc_init<<<gm_grid(n, 256, sms), 256, 0, s>>>(n, label.p);
    count_launch();
    bool done = false;
    for (int iter = 0; iter < 4096 && !done; ++iter) {  // each round at least halves the number of trees per component
        SPHB_CUDA(cudaMemsetAsync(changed.p, 0, sizeof(int), s));
        cc_hook<<<gm_grid(n * 32, 256, sms), 256, 0, s>>>(n, d_rowptr, d_ids, label.p, changed.p);
        cc_jump<<<gm_grid(n, 256, sms), 256, 0, s>>>(n, label.p);
        count_launch(2);
        int h = 0;
        SPHB_CUDA(cudaMemcpyAsync(&h, changed.p, sizeof(int), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaStreamSynchronize(s));
        done = h == 0;
    }
    if (!done) return fail(SPHB_ERR_CUDA, "weak_components: label propagation did not converge");
    cc_flag_roots<<<gm_grid(n + 1, 256, sms), 256, 0, s>>>(n, label.p, root.p);
    count_launch();
    if (int rc = exclusive_sum<int>(root.p, rank.p, n + 1, s)) return rc;
    cc_number<<<gm_grid(n, 256, sms), 256, 0, s>>>(n, label.p, rank.p, d_labels);
    count_launch();
    int total = 0;
    SPHB_CUDA(cudaMemcpyAsync(&total, rank.p + n, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));
    *num_components = total;
    return 0;
}

}  // namespace sphb

namespace sphb {
// implemented in walks.cu
int random_walks_slab_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks, uint64_t walk_len,
                          int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize, int64_t q0,
                          int64_t q1, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s);
// implemented in prow.cu
int prow_graph_ragged_launch(const int64_t* d_rowptr, const int64_t* d_ids, const float* d_dist, int64_t n, int max_k, float perplexity, int mode,
                             uint32_t* d_cols, float* d_vals, cudaStream_t stream);
}
using namespace sphb;

// ---------------------------------------------------------------------------------------------
// host-pointer entry points
// ---------------------------------------------------------------------------------------------
namespace {
template <class T>
int up(DevBuf<T>& b, const T* h, int64_t count, cudaStream_t s) {
    SPHB_CUDA(b.alloc(static_cast<size_t>(std::max<int64_t>(count, 1)), s));
    if (count > 0) SPHB_CUDA(cudaMemcpyAsync(b.p, h, count * sizeof(T), cudaMemcpyHostToDevice, s));
    return 0;
}
// rowptr == NULL means a fixed-k graph (utils::Graph::updateFixedNumNeighbors): offsets r * k
int graph_rowptr(int64_t n, const int64_t* rowptr, int k, std::vector<int64_t>& tmp, const int64_t** out) {
    if (rowptr) {
        if (rowptr[0] != 0) return 1;
        for (int64_t r = 0; r < n; ++r)
            if (rowptr[r + 1] < rowptr[r]) return 1;
        *out = rowptr;
        return 0;
    }
    if (k <= 0) return 1;
    tmp.resize(static_cast<size_t>(n) + 1);
    for (int64_t r = 0; r <= n; ++r) tmp[r] = r * k;
    *out = tmp.data();
    return 0;
}
int graph_minmerge_host(const char* what, int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const uint64_t* parents,
                        int64_t n_out, int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (n == 0 && !parents && out_rowptr && out_nnz) {
        out_rowptr[0] = 0;
        *out_nnz = 0;
        return SPHB_OK;
    }
    std::vector<int64_t> tmp;
    const int64_t* rp = nullptr;
    if (n < 0 || n_out <= 0 || !out_rowptr || !out_nnz || cap < 0 || graph_rowptr(n, rowptr, k, tmp, &rp))
        return fail(SPHB_ERR_INVALID, std::string(what) + ": bad arguments");
    const int64_t nnz = rp[n];
    if (nnz > 0 && (!ids || !dist || !out_ids || !out_dist)) return fail(SPHB_ERR_INVALID, std::string(what) + ": bad arguments");
    for (int64_t e = 0; e < nnz; ++e)
        if (ids[e] < 0 || ids[e] >= n) return fail(SPHB_ERR_INVALID, std::string(what) + ": neighbour id out of range (missing neighbours must be repaired first)");
    if (parents)
        for (int64_t i = 0; i < n; ++i)
            if (parents[i] >= static_cast<uint64_t>(n_out)) return fail(SPHB_ERR_INVALID, std::string(what) + ": parent id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, dids, orp, oids;
    DevBuf<float> ddist, odist;
    DevBuf<uint64_t> dpar;
    if (int rc = up(drp, rp, n + 1, s)) return rc;
    if (int rc = up(dids, ids, nnz, s)) return rc;
    if (int rc = up(ddist, dist, nnz, s)) return rc;
    if (parents)
        if (int rc = up(dpar, parents, n, s)) return rc;
    SPHB_CUDA(orp.alloc(n_out + 1, s));
    SPHB_CUDA(oids.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(odist.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = graph_minmerge_dev(n, drp.p, dids.p, ddist.p, nnz, parents ? dpar.p : nullptr, n_out, orp.p, oids.p, odist.p, cap, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n_out + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_ids, oids.p, m * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_dist, odist.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}
bool bad_csr_in(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val) {
    return n < 0 || !rowptr || rowptr[0] != 0 || rowptr[n] < 0 || (rowptr[n] > 0 && (!col || !val));
}
}  // namespace

extern "C" {

SPHB_API int sphb_symmetrize_graph(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, int64_t* out_rowptr,
                                   int64_t* out_ids, float* out_dist, int64_t cap, int64_t* out_nnz) {
    return graph_minmerge_host("symmetrize_graph", n, rowptr, k, ids, dist, nullptr, std::max<int64_t>(n, 1), out_rowptr, out_ids, out_dist, cap, out_nnz);
}

SPHB_API int sphb_random_walks_slab(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, uint64_t num_walks, uint64_t walk_len,
                                    int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize,
                                    int64_t q0, int64_t q1, int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr_in(n, rowptr, col, val) || !out_rowptr || !out_nnz || cap < 0 || weighting < 0 || weighting > 4 || walk_len == 0 || q0 < 0 || q1 > n || q1 < q0)
        return fail(SPHB_ERR_INVALID, "random_walks_slab: bad arguments");
    const int64_t nnz = rowptr[n];
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, orp;
    DevBuf<uint32_t> dcol, oc;
    DevBuf<float> dval, ov;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    SPHB_CUDA(orp.alloc(q1 - q0 + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = random_walks_slab_dev(n, drp.p, dcol.p, dval.p, num_walks, walk_len, weighting, prune_value, prune_steps, seed, remove_diagonal, normalize,
                                       q0, q1, orp.p, oc.p, ov.p, cap, &m, s))
        return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (q1 - q0 + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_prow_graph_ragged(int64_t n, const int64_t* rowptr, const int64_t* ids, const float* dist, float perplexity, int linear,
                                    int64_t* out_rowptr, uint32_t* out_cols, float* out_vals) {
    if (int rc = ensure_init()) return rc;
    if (n < 0 || !rowptr || rowptr[0] != 0 || !out_rowptr) return fail(SPHB_ERR_INVALID, "prow_graph_ragged: bad arguments");
    int64_t max_k = 0;
    for (int64_t p = 0; p < n; ++p) {
        const int64_t kp = rowptr[p + 1] - rowptr[p];
        if (kp < 1) return fail(SPHB_ERR_INVALID, "prow_graph_ragged: every row must hold at least the point itself");
        max_k = std::max(max_k, kp);
        out_rowptr[p] = rowptr[p] - p;
    }
    out_rowptr[n] = rowptr[n] - n;
    const int64_t nnz = rowptr[n], onnz = nnz - n;
    if (nnz > 0 && (!ids || !dist)) return fail(SPHB_ERR_INVALID, "prow_graph_ragged: bad arguments");
    if (onnz <= 0) return SPHB_OK;
    if (!out_cols || !out_vals) return fail(SPHB_ERR_INVALID, "prow_graph_ragged: bad arguments");
    if (max_k > 1024) return fail(SPHB_ERR_UNSUPPORTED, "prow_graph_ragged: rows with more than 1024 neighbours unsupported");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, dids;
    DevBuf<float> ddist, ov;
    DevBuf<uint32_t> oc;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dids, ids, nnz, s)) return rc;
    if (int rc = up(ddist, dist, nnz, s)) return rc;
    SPHB_CUDA(oc.alloc(onnz, s));
    SPHB_CUDA(ov.alloc(onnz, s));
    if (int rc = prow_graph_ragged_launch(drp.p, dids.p, ddist.p, n, static_cast<int>(max_k), perplexity, linear ? 1 : 0, oc.p, ov.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_cols, oc.p, onnz * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_vals, ov.p, onnz * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_weak_components(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, int64_t* labels, int64_t* num_components) {
    if (int rc = ensure_init()) return rc;
    std::vector<int64_t> tmp;
    const int64_t* rp = nullptr;
    if (n < 0 || !num_components || (n > 0 && !labels) || graph_rowptr(n, rowptr, k, tmp, &rp)) return fail(SPHB_ERR_INVALID, "weak_components: bad arguments");
    if (n == 0) {
        *num_components = 0;
        return SPHB_OK;
    }
    const int64_t nnz = rp[n];
    if (nnz > 0 && !ids) return fail(SPHB_ERR_INVALID, "weak_components: bad arguments");
    for (int64_t e = 0; e < nnz; ++e)
        if (ids[e] < 0 || ids[e] >= n) return fail(SPHB_ERR_INVALID, "weak_components: neighbour id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, dids, dl;
    if (int rc = up(drp, rp, n + 1, s)) return rc;
    if (int rc = up(dids, ids, nnz, s)) return rc;
    SPHB_CUDA(dl.alloc(n, s));
    int64_t nc = 0;
    if (int rc = weak_components_dev(n, drp.p, dids.p, dl.p, &nc, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(labels, dl.p, n * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    *num_components = nc;
    return SPHB_OK;
}

// Device-pointer variant for pipelines that keep the graph in HBM (sphb_knn_exact_dev -> here -> weak components / P rows).
// d_rowptr may be NULL for a fixed-k graph.  *out_nnz is a HOST value (the call synchronises the stream once to size the output).
SPHB_API int sphb_symmetrize_graph_dev(int64_t n, const int64_t* d_rowptr, int k, const int64_t* d_ids, const float* d_dist, int64_t nnz,
                                       int64_t* d_out_rowptr, int64_t* d_out_ids, float* d_out_dist, int64_t cap, int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (n <= 0 || nnz < 0 || !d_out_rowptr || !out_nnz || cap < 0 || (!d_rowptr && (k <= 0 || nnz != n * static_cast<int64_t>(k))) ||
        (nnz > 0 && (!d_ids || !d_dist || !d_out_ids || !d_out_dist)))
        return fail(SPHB_ERR_INVALID, "symmetrize_graph_dev: bad arguments");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DevBuf<int64_t> rp;
    if (!d_rowptr) {
        SPHB_CUDA(rp.alloc(n + 1, s));
        gm_fixed_rowptr<<<gm_grid(n + 1, 256, ctx().num_sms), 256, 0, s>>>(n, k, rp.p);
        count_launch();
        d_rowptr = rp.p;
    }
    return graph_minmerge_dev(n, d_rowptr, d_ids, d_dist, nnz, nullptr, n, d_out_rowptr, d_out_ids, d_out_dist, cap, out_nnz, s);
}

SPHB_API int sphb_merge_graph_nodes(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const uint64_t* parents,
                                    int64_t n_merged, int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap, int64_t* out_nnz) {
    if (!parents) return fail(SPHB_ERR_INVALID, "merge_graph_nodes: bad arguments");
    return graph_minmerge_host("merge_graph_nodes", n, rowptr, k, ids, dist, parents, n_merged, out_rowptr, out_ids, out_dist, cap, out_nnz);
}

SPHB_API int sphb_merge_data_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, const uint64_t* parents,
                                       int64_t n_merged, int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr_in(n, rowptr, col, val) || !parents || n_merged <= 0 || !out_rowptr || !out_nnz || cap < 0)
        return fail(SPHB_ERR_INVALID, "merge_data_distances: bad arguments");
    const int64_t nnz = rowptr[n];
    for (int64_t i = 0; i < n; ++i)
        if (parents[i] >= static_cast<uint64_t>(n_merged)) return fail(SPHB_ERR_INVALID, "merge_data_distances: parent id out of range");
    for (int64_t e = 0; e < nnz; ++e)
        if (col[e] >= static_cast<uint64_t>(n)) return fail(SPHB_ERR_INVALID, "merge_data_distances: column id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, orp;
    DevBuf<uint32_t> dcol, oc;
    DevBuf<float> dval, ov;
    DevBuf<uint64_t> dpar;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    if (int rc = up(dpar, parents, n, s)) return rc;
    SPHB_CUDA(orp.alloc(n_merged + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(cap, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(cap, 1), s));
    int64_t m = 0;
    if (int rc = merge_data_distances_dev(n, drp.p, dcol.p, dval.p, nnz, dpar.p, n_merged, orp.p, oc.p, ov.p, cap, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n_merged + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_rw_as_knn_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t* out_rowptr, int64_t* out_ids,
                                      float* out_dist, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr_in(n, rowptr, col, val) || !out_rowptr || !out_nnz) return fail(SPHB_ERR_INVALID, "rw_as_knn_distances: bad arguments");
    const int64_t nnz = rowptr[n];
    if (nnz > 0 && (!out_ids || !out_dist)) return fail(SPHB_ERR_INVALID, "rw_as_knn_distances: bad arguments");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, orp, oids;
    DevBuf<uint32_t> dcol;
    DevBuf<float> dval, odist;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oids.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(odist.alloc(std::max<int64_t>(nnz, 1), s));
    int64_t m = 0;
    if (int rc = rw_as_knn_distances_dev(n, drp.p, dcol.p, dval.p, nnz, orp.p, oids.p, odist.p, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_ids, oids.p, m * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_dist, odist.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_remove_zeros(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t* out_rowptr, uint32_t* out_col,
                               float* out_val, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr_in(n, rowptr, col, val) || !out_rowptr || !out_nnz) return fail(SPHB_ERR_INVALID, "remove_zeros: bad arguments");
    const int64_t nnz = rowptr[n];
    if (nnz > 0 && (!out_col || !out_val)) return fail(SPHB_ERR_INVALID, "remove_zeros: bad arguments");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, orp;
    DevBuf<uint32_t> dcol, oc;
    DevBuf<float> dval, ov;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(nnz, 1), s));
    int64_t m = 0;
    if (int rc = remove_zeros_dev(n, drp.p, dcol.p, dval.p, orp.p, oc.p, ov.p, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_rw_post_merge(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int remove_diagonal, int norm,
                                int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t* out_nnz) {
    if (int rc = ensure_init()) return rc;
    if (!rowptr || !out_rowptr || !out_nnz || n < 0 || norm < 0 || norm > 2) return fail(SPHB_ERR_INVALID, "rw_post_merge: bad arguments");
    *out_nnz = 0;
    if (n == 0) return SPHB_OK;
    cudaStream_t s = ctx().stream;
    const int64_t nnz = rowptr[n];
    DevBuf<int64_t> drp, orp;
    DevBuf<uint32_t> dcol, oc;
    DevBuf<float> dval, ov;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    SPHB_CUDA(orp.alloc(n + 1, s));
    SPHB_CUDA(oc.alloc(std::max<int64_t>(nnz, 1), s));
    SPHB_CUDA(ov.alloc(std::max<int64_t>(nnz, 1), s));
    int64_t m = 0;
    if (int rc = rw_post_merge_dev(n, drp.p, dcol.p, dval.p, remove_diagonal, norm, orp.p, oc.p, ov.p, &m, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_rowptr, orp.p, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    if (m > 0) {
        SPHB_CUDA(cudaMemcpyAsync(out_col, oc.p, m * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaMemcpyAsync(out_val, ov.p, m * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    SPHB_CUDA(cudaStreamSynchronize(s));
    *out_nnz = m;
    return SPHB_OK;
}

SPHB_API int sphb_rw_post_merge_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int remove_diagonal, int norm,
                                    int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t* out_nnz, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_rowptr || !d_out_rowptr || !out_nnz || n < 0 || norm < 0 || norm > 2) return fail(SPHB_ERR_INVALID, "rw_post_merge_dev: bad arguments");
    *out_nnz = 0;
    if (n == 0) return SPHB_OK;
    return rw_post_merge_dev(n, d_rowptr, d_col, d_val, remove_diagonal, norm, d_out_rowptr, d_out_col, d_out_val, out_nnz, static_cast<cudaStream_t>(stream));
}

SPHB_API int sphb_level_knn_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int mode, int64_t k, int64_t* out_ids,
                                  float* out_dist) {
    if (int rc = ensure_init()) return rc;
    if (bad_csr_in(n, rowptr, col, val) || n == 0 || k <= 0 || !out_ids || !out_dist || (mode != 0 && mode != 1))
        return fail(SPHB_ERR_INVALID, "level_knn_walks: bad arguments");
    const int64_t nnz = rowptr[n];
    for (int64_t e = 0; e < nnz; ++e)
        if (col[e] >= static_cast<uint64_t>(n)) return fail(SPHB_ERR_INVALID, "level_knn_walks: column id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> drp, oids;
    DevBuf<uint32_t> dcol;
    DevBuf<float> dval, odist;
    if (int rc = up(drp, rowptr, n + 1, s)) return rc;
    if (int rc = up(dcol, col, nnz, s)) return rc;
    if (int rc = up(dval, val, nnz, s)) return rc;
    const size_t nk = static_cast<size_t>(n) * static_cast<size_t>(k);
    SPHB_CUDA(oids.alloc(nk, s));
    SPHB_CUDA(odist.alloc(nk, s));
    if (int rc = level_knn_walks_dev(n, drp.p, dcol.p, dval.p, nnz, mode, k, oids.p, odist.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_ids, oids.p, nk * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_dist, odist.p, nk * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

SPHB_API int sphb_euclid_hausdorff(const float* x, int64_t n, int d, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix,
                                   const int64_t* pa, const int64_t* pb, int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!x || n <= 0 || d <= 0 || ncomp <= 0 || !comp_ptr || comp_ptr[0] != 0 || npairs < 0 || (npairs > 0 && (!pa || !pb || !out)))
        return fail(SPHB_ERR_INVALID, "euclid_hausdorff: bad arguments");
    const int64_t npix = comp_ptr[ncomp];
    if (npix < 0 || (npix > 0 && !comp_pix)) return fail(SPHB_ERR_INVALID, "euclid_hausdorff: bad arguments");
    for (int64_t c = 0; c < ncomp; ++c)
        if (comp_ptr[c + 1] < comp_ptr[c]) return fail(SPHB_ERR_INVALID, "euclid_hausdorff: comp_ptr not monotone");
    for (int64_t i = 0; i < npix; ++i)
        if (comp_pix[i] < 0 || comp_pix[i] >= n) return fail(SPHB_ERR_INVALID, "euclid_hausdorff: pixel id out of range");
    for (int64_t p = 0; p < npairs; ++p)
        if (pa[p] < 0 || pa[p] >= ncomp || pb[p] < 0 || pb[p] >= ncomp) return fail(SPHB_ERR_INVALID, "euclid_hausdorff: component id out of range");
    if (npairs == 0) return SPHB_OK;
    cudaStream_t s = ctx().stream;
    DevBuf<float> dx, dout;
    DevBuf<int64_t> dptr, dpix, da, db;
    if (int rc = up(dx, x, n * d, s)) return rc;
    if (int rc = up(dptr, comp_ptr, ncomp + 1, s)) return rc;
    if (int rc = up(dpix, comp_pix, npix, s)) return rc;
    if (int rc = up(da, pa, npairs, s)) return rc;
    if (int rc = up(db, pb, npairs, s)) return rc;
    SPHB_CUDA(dout.alloc(npairs, s));
    if (int rc = euclid_hausdorff_dev(dx.p, d, dptr.p, dpix.p, da.p, db.p, npairs, dout.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out, dout.p, npairs * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}


SPHB_API int sphb_level_knn_euclid(const float* x, int64_t n, int d, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, int64_t k,
                                   int64_t* out_ids, float* out_dist) {
    if (int rc = ensure_init()) return rc;
    if (!x || n <= 0 || d <= 0 || ncomp <= 0 || !comp_ptr || comp_ptr[0] != 0 || k <= 0 || k > ncomp || !out_ids || !out_dist)
        return fail(SPHB_ERR_INVALID, "level_knn_euclid: bad arguments");
    if (k > 992) return fail(SPHB_ERR_UNSUPPORTED, "level_knn_euclid: k > 992 unsupported");
    const int64_t npix = comp_ptr[ncomp];
    if (npix < 0 || (npix > 0 && !comp_pix)) return fail(SPHB_ERR_INVALID, "level_knn_euclid: bad arguments");
    for (int64_t c = 0; c < ncomp; ++c)
        if (comp_ptr[c + 1] < comp_ptr[c]) return fail(SPHB_ERR_INVALID, "level_knn_euclid: comp_ptr not monotone");
    for (int64_t i = 0; i < npix; ++i)
        if (comp_pix[i] < 0 || comp_pix[i] >= n) return fail(SPHB_ERR_INVALID, "level_knn_euclid: pixel id out of range");
    cudaStream_t s = ctx().stream;
    DevBuf<float> dx, dd;
    DevBuf<int64_t> dptr, dpix, di;
    if (int rc = up(dx, x, n * d, s)) return rc;
    if (int rc = up(dptr, comp_ptr, ncomp + 1, s)) return rc;
    if (int rc = up(dpix, comp_pix, npix, s)) return rc;
    SPHB_CUDA(di.alloc(ncomp * k, s));
    SPHB_CUDA(dd.alloc(ncomp * k, s));
    if (int rc = level_knn_euclid_dev(dx.p, d, ncomp, dptr.p, dpix.p, static_cast<int>(k), di.p, dd.p, s)) return rc;
    SPHB_CUDA(cudaMemcpyAsync(out_ids, di.p, ncomp * k * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(out_dist, dd.p, ncomp * k * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    return SPHB_OK;
}

}  // extern "C"
