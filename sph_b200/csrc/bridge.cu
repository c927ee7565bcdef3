// component bridging — NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861).
//   centres of mass per component          :655-680  (utils::computeMean, sph/utils/Math.hpp:29-48: sequential fp64 sums)
//   per MST edge (compA, compB): for every point of B its nearest point of A, then the closest such pair  :592-653
// The reference runs one faiss search per MST edge (IndexFlat for npA <= 3900, else IndexIVFFlat nlist 100 / nprobe 10, an
// approximate search) and keeps only the closest pair.  Here all edges of the tree go through ONE launch of a tiled pair
// kernel that looks for the closest pair directly and exactly:
//   * 64 x 64 (B x A) tiles, fp32 sum of squared differences (relative error <= (d + 4) 2^-24, all terms >= 0),
//   * a running upper bound per edge (atomicMin over the fp32 tile minima); a pair whose fp32 distance lies within the
//     error band of the bound is re-evaluated with the exact key of Stage A (fp64 FMA chain rounded once, knn_common.cuh)
//     and enters a 64-bit atomicMin on (exact distance, position in B) — the pair(s) with the exact minimum always lie
//     inside the band, whatever the order of the tiles,
//   * a second small kernel picks the lowest position in A among the exact ties of the winning B point.
// The result is the lexicographic minimum over (distance, position in B, position in A): what the reference's loop
// (:628-636, strict <) returns when the per-point search is exact.  Distances are squared L2 as faiss returns them.
#include "../../include/sphb.h"
#include "knn_common.cuh"

#include <cub/cub.cuh>

#include <algorithm>
#include <vector>

namespace sphb {

constexpr int BR_T = 64;   // tile edge (points of B x points of A)
constexpr int BR_DC = 32;  // channels staged per step

struct BridgeEdges {
    const int64_t* comp_a;    // [m] component whose points are searched ("A", the MST edge's source)
    const int64_t* comp_b;    // [m] component whose points query ("B", the MST edge's target)
    const int64_t* tile_off;  // [m + 1] first tile of every edge
    int m;
};

__device__ __forceinline__ int edge_of_tile(const int64_t* __restrict__ tile_off, int m, int64_t t) {
    int lo = 0, hi = m;  // largest e with tile_off[e] <= t
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (tile_off[mid] <= t) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(256) bridge_pairs_kernel(const float* __restrict__ X, int d, const int64_t* __restrict__ perm,
                                                           const int64_t* __restrict__ off, BridgeEdges E, int64_t total_tiles, float band_mul,
                                                           unsigned int* __restrict__ ubits, unsigned long long* __restrict__ best) {
    __shared__ __align__(16) float sA[BR_DC][BR_T + 4];
    __shared__ __align__(16) float sB[BR_DC][BR_T + 4];
    __shared__ int64_t rowA[BR_T], rowB[BR_T];
    __shared__ float s_red[8];
    __shared__ float s_u;
    const int t = threadIdx.x, ty = t >> 4, tx = t & 15;
    const bool vec = (d & 3) == 0;
    for (int64_t tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int e = edge_of_tile(E.tile_off, E.m, tile);
        const int64_t a0 = off[E.comp_a[e]], na = off[E.comp_a[e] + 1] - a0;
        const int64_t b0 = off[E.comp_b[e]], nb = off[E.comp_b[e] + 1] - b0;
        const int64_t tiles_a = (na + BR_T - 1) / BR_T;
        const int64_t lt = tile - E.tile_off[e];
        const int64_t tb = lt / tiles_a, ta = lt % tiles_a;
        __syncthreads();  // previous tile done with the shared arrays
        if (t < BR_T) {
            const int64_t p = ta * BR_T + t;
            rowA[t] = p < na ? perm[a0 + p] : -1;
        } else if (t < 2 * BR_T) {
            const int64_t p = tb * BR_T + (t - BR_T);
            rowB[t - BR_T] = p < nb ? perm[b0 + p] : -1;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        for (int c0 = 0; c0 < d; c0 += BR_DC) {
            const int row = t & 63, q = t >> 6;  // q = 0..3
            const int64_t ga = rowA[row], gb = rowB[row];
            if (vec) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int c = c0 + (q + 4 * h) * 4;
                    float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
                    if (c < d) {
                        if (ga >= 0) va = __ldg(reinterpret_cast<const float4*>(X + ga * d + c));
                        if (gb >= 0) vb = __ldg(reinterpret_cast<const float4*>(X + gb * d + c));
                    }
                    const int cc = (q + 4 * h) * 4;
                    sA[cc][row] = va.x; sA[cc + 1][row] = va.y; sA[cc + 2][row] = va.z; sA[cc + 3][row] = va.w;
                    sB[cc][row] = vb.x; sB[cc + 1][row] = vb.y; sB[cc + 2][row] = vb.z; sB[cc + 3][row] = vb.w;
                }
            } else {
#pragma unroll
                for (int h = 0; h < BR_DC / 4; ++h) {
                    const int cc = q + 4 * h, c = c0 + cc;
                    sA[cc][row] = (c < d && ga >= 0) ? __ldg(X + ga * d + c) : 0.f;
                    sB[cc][row] = (c < d && gb >= 0) ? __ldg(X + gb * d + c) : 0.f;
                }
            }
            __syncthreads();
#pragma unroll
            for (int c = 0; c < BR_DC; ++c) {
                const float4 bv = *reinterpret_cast<const float4*>(&sB[c][ty * 4]);
                const float4 av = *reinterpret_cast<const float4*>(&sA[c][tx * 4]);
                const float b[4] = {bv.x, bv.y, bv.z, bv.w}, a[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float df = b[i] - a[j];
                        acc[i][j] = fmaf(df, df, acc[i][j]);
                    }
            }
            __syncthreads();
        }
        // tile minimum over the valid pairs -> running bound of the edge
        float mn = INFINITY;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (rowB[ty * 4 + i] >= 0 && rowA[tx * 4 + j] >= 0) mn = fminf(mn, acc[i][j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        if ((t & 31) == 0) s_red[t >> 5] = mn;
        __syncthreads();
        if (t == 0) {
            float m8 = s_red[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) m8 = fminf(m8, s_red[w]);
            const unsigned int old = atomicMin(&ubits[e], __float_as_uint(m8));  // non-negative floats order like their bit patterns
            s_u = fminf(m8, __uint_as_float(old));
        }
        __syncthreads();
        const float lim = s_u * band_mul + 1e-37f;
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t gb = rowB[ty * 4 + i], ga = rowA[tx * 4 + j];
                if (gb >= 0 && ga >= 0 && acc[i][j] <= lim) {
                    const float ex = exact_key(X, gb, ga, d, METRIC_L2);
                    const unsigned long long key = (static_cast<unsigned long long>(__float_as_uint(ex)) << 32) |
                                                   static_cast<unsigned long long>(tb * BR_T + ty * 4 + i);
                    atomicMin(&best[e], key);
                }
            }
    }
}

// Lowest position in A among the points at exactly the winning distance from the winning B point.
__global__ void __launch_bounds__(256) bridge_pick_a_kernel(const float* __restrict__ X, int d, const int64_t* __restrict__ perm,
                                                            const int64_t* __restrict__ off, BridgeEdges E,
                                                            const unsigned long long* __restrict__ best, unsigned long long* __restrict__ apos) {
    const int e = blockIdx.y;
    const unsigned long long key = best[e];
    if (key == ~0ull) return;
    const int64_t a0 = off[E.comp_a[e]], na = off[E.comp_a[e] + 1] - a0;
    const int64_t gb = perm[off[E.comp_b[e]] + static_cast<int64_t>(key & 0xffffffffull)];
    const unsigned int dbits = static_cast<unsigned int>(key >> 32);
    for (int64_t p = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < na; p += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        if (__float_as_uint(exact_key(X, gb, perm[a0 + p], d, METRIC_L2)) == dbits) atomicMin(&apos[e], static_cast<unsigned long long>(p));
    }
}

__global__ void bridge_finish_kernel(const int64_t* __restrict__ perm, const int64_t* __restrict__ off, BridgeEdges E,
                                     const unsigned long long* __restrict__ best, const unsigned long long* __restrict__ apos,
                                     int64_t* __restrict__ out_a, int64_t* __restrict__ out_b, float* __restrict__ out_dist) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= E.m) return;
    const unsigned long long key = best[e];
    if (key == ~0ull || apos[e] == ~0ull) {  // an empty component: the reference would index with -1 here
        out_a[e] = -1; out_b[e] = -1; out_dist[e] = FLT_MAX;
        return;
    }
    out_a[e] = perm[off[E.comp_a[e]] + static_cast<int64_t>(apos[e])];
    out_b[e] = perm[off[E.comp_b[e]] + static_cast<int64_t>(key & 0xffffffffull)];
    out_dist[e] = __uint_as_float(static_cast<unsigned int>(key >> 32));
}

// Centre of mass of every component: fp64 sums in ascending point order (the order of getComponentValues :574-590 and
// utils::computeMean), divided by the count, rounded to fp32.  One CTA per component; a batch of member rows is staged in
// shared memory by all threads (coalesced), then one thread per channel adds the batch in order.
constexpr int CC_THREADS = 256;
constexpr int CC_TILE = 8192;  // floats of shared memory per batch

__global__ void __launch_bounds__(CC_THREADS) component_centers_kernel(const float* __restrict__ X, int d, const int64_t* __restrict__ perm,
                                                                       const int64_t* __restrict__ off, float* __restrict__ centers) {
    __shared__ float tile[CC_TILE];
    const int comp = blockIdx.x;
    const int64_t p0 = off[comp], cnt = off[comp + 1] - p0;
    for (int c0 = 0; c0 < d; c0 += CC_THREADS) {
        const int dc = min(d - c0, CC_THREADS);
        const int batch = min(CC_TILE / dc, 1024);
        double acc = 0.0;
        for (int64_t i0 = 0; i0 < cnt; i0 += batch) {
            const int nb = static_cast<int>(cnt - i0 < batch ? cnt - i0 : batch);
            __syncthreads();
            for (int idx = threadIdx.x; idx < nb * dc; idx += CC_THREADS) {
                const int r = idx / dc, c = idx - r * dc;
                tile[idx] = __ldg(X + perm[p0 + i0 + r] * d + c0 + c);
            }
            __syncthreads();
            if (threadIdx.x < dc)
                for (int r = 0; r < nb; ++r) acc += static_cast<double>(tile[r * dc + threadIdx.x]);
        }
        if (threadIdx.x < dc) centers[static_cast<int64_t>(comp) * d + c0 + threadIdx.x] = cnt > 0 ? static_cast<float>(acc / static_cast<double>(cnt)) : 0.f;
    }
}

__global__ void bridge_label_keys_kernel(const int64_t* __restrict__ labels, int64_t n, int64_t ncomp, int64_t* __restrict__ keys,
                                         int64_t* __restrict__ idx, int* __restrict__ bad) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t l = labels[i];
    if (l < 0 || l >= ncomp) atomicExch(bad, 1);
    keys[i] = l;
    idx[i] = i;
}

__global__ void bridge_offsets_kernel(const int64_t* __restrict__ sorted_labels, int64_t n, int64_t ncomp, int64_t* __restrict__ off) {
    // off[c] = first position whose label is >= c (labels sorted ascending), off[ncomp] = n
    const int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (c > ncomp) return;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (sorted_labels[mid] < c) lo = mid + 1; else hi = mid;
    }
    off[c] = lo;
}

// Members of every component in ascending point order: perm[off[c] .. off[c+1]).
struct ComponentLists {
    DevBuf<int64_t> perm, off;
    std::vector<int64_t> h_off;
};

static int build_component_lists(const int64_t* d_labels, int64_t n, int64_t ncomp, ComponentLists& L, cudaStream_t s) {
    DevBuf<int64_t> k0, k1, i0;
    DevBuf<int> bad;
    SPHB_CUDA(k0.alloc(n, s));
    SPHB_CUDA(k1.alloc(n, s));
    SPHB_CUDA(i0.alloc(n, s));
    SPHB_CUDA(L.perm.alloc(n, s));
    SPHB_CUDA(L.off.alloc(ncomp + 1, s));
    SPHB_CUDA(bad.alloc(1, s));
    SPHB_CUDA(cudaMemsetAsync(bad.p, 0, sizeof(int), s));
    bridge_label_keys_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, s>>>(d_labels, n, ncomp, k0.p, i0.p, bad.p);
    count_launch();
    int end_bit = 1;
    while ((int64_t{1} << end_bit) < ncomp && end_bit < 62) ++end_bit;
    size_t bytes = 0;
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes, k0.p, k1.p, i0.p, L.perm.p, n, 0, end_bit, s));
    DevBuf<char> tmp;
    SPHB_CUDA(tmp.alloc(bytes, s));
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, k0.p, k1.p, i0.p, L.perm.p, n, 0, end_bit, s));  // stable: ascending point id inside a component
    count_launch();
    bridge_offsets_kernel<<<static_cast<unsigned>((ncomp + 1 + 255) / 256), 256, 0, s>>>(k1.p, n, ncomp, L.off.p);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    L.h_off.resize(static_cast<size_t>(ncomp) + 1);
    int hbad = 0;
    SPHB_CUDA(cudaMemcpyAsync(L.h_off.data(), L.off.p, (ncomp + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(&hbad, bad.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    if (hbad) return fail(SPHB_ERR_INVALID, "component labels must lie in [0, num_components)");
    return 0;
}

int component_centers_dev(const float* d_x, int64_t n, int d, const int64_t* d_labels, int64_t ncomp, float* d_centers, cudaStream_t s) {
    if (ncomp <= 0) return 0;
    ComponentLists L;
    if (int rc = build_component_lists(d_labels, n, ncomp, L, s)) return rc;
    component_centers_kernel<<<static_cast<unsigned>(ncomp), CC_THREADS, 0, s>>>(d_x, d, L.perm.p, L.off.p, d_centers);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int bridge_components_dev(const float* d_x, int64_t n, int d, const int64_t* d_labels, int64_t ncomp, const int64_t* h_edges, int64_t m,
                          int64_t* d_out_a, int64_t* d_out_b, float* d_out_dist, cudaStream_t s) {
    if (m <= 0) return 0;
    if (m > (int64_t{1} << 30)) return fail(SPHB_ERR_UNSUPPORTED, "bridge_components: too many edges");
    ComponentLists L;
    if (int rc = build_component_lists(d_labels, n, ncomp, L, s)) return rc;
    std::vector<int64_t> ha(m), hb(m), toff(m + 1);
    toff[0] = 0;
    int64_t max_na = 1;
    for (int64_t e = 0; e < m; ++e) {
        const int64_t ca = h_edges[2 * e], cb = h_edges[2 * e + 1];
        if (ca < 0 || ca >= ncomp || cb < 0 || cb >= ncomp) return fail(SPHB_ERR_INVALID, "bridge_components: edge names a component out of range");
        ha[e] = ca;
        hb[e] = cb;
        const int64_t na = L.h_off[ca + 1] - L.h_off[ca], nb = L.h_off[cb + 1] - L.h_off[cb];
        if (nb > (int64_t{1} << 32) - 1) return fail(SPHB_ERR_UNSUPPORTED, "bridge_components: component with more than 2^32 - 1 points");
        toff[e + 1] = toff[e] + ((na + BR_T - 1) / BR_T) * ((nb + BR_T - 1) / BR_T);
        max_na = std::max(max_na, na);
    }
    DevBuf<int64_t> da, db, dt;
    DevBuf<unsigned int> ub;
    DevBuf<unsigned long long> best, apos;
    SPHB_CUDA(da.alloc(m, s));
    SPHB_CUDA(db.alloc(m, s));
    SPHB_CUDA(dt.alloc(m + 1, s));
    SPHB_CUDA(ub.alloc(m, s));
    SPHB_CUDA(best.alloc(m, s));
    SPHB_CUDA(apos.alloc(m, s));
    SPHB_CUDA(cudaMemcpyAsync(da.p, ha.data(), m * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(db.p, hb.data(), m * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemcpyAsync(dt.p, toff.data(), (m + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    SPHB_CUDA(cudaMemsetAsync(ub.p, 0x7f, m * sizeof(unsigned int), s));  // 0x7f7f7f7f: a huge finite float
    SPHB_CUDA(cudaMemsetAsync(best.p, 0xff, m * sizeof(unsigned long long), s));
    SPHB_CUDA(cudaMemsetAsync(apos.p, 0xff, m * sizeof(unsigned long long), s));
    BridgeEdges E{da.p, db.p, dt.p, static_cast<int>(m)};
    const int64_t total = toff[m];
    if (total > 0) {
        const double e32 = (d + 4) * 5.9604644775390625e-08 * 1.01;  // (d + 4) 2^-24
        const float band = static_cast<float>((1.0 + e32) / (1.0 - e32) * (1.0 + 1e-6));
        const int64_t grid = std::min<int64_t>(total, static_cast<int64_t>(ctx().num_sms) * 8);
        bridge_pairs_kernel<<<static_cast<unsigned>(grid), 256, 0, s>>>(d_x, d, L.perm.p, L.off.p, E, total, band, ub.p, best.p);
        count_launch();
        const unsigned gx = static_cast<unsigned>(std::min<int64_t>((max_na + 255) / 256, 1024));
        for (int64_t e0 = 0; e0 < m; e0 += 65535) {  // gridDim.y limit
            const int64_t me = std::min<int64_t>(65535, m - e0);
            BridgeEdges Ee{da.p + e0, db.p + e0, dt.p + e0, static_cast<int>(me)};
            bridge_pick_a_kernel<<<dim3(gx, static_cast<unsigned>(me)), 256, 0, s>>>(d_x, d, L.perm.p, L.off.p, Ee, best.p + e0, apos.p + e0);
            count_launch();
        }
    }
    bridge_finish_kernel<<<static_cast<unsigned>((m + 255) / 256), 256, 0, s>>>(L.perm.p, L.off.p, E, best.p, apos.p, d_out_a, d_out_b, d_out_dist);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    SPHB_CUDA(cudaStreamSynchronize(s));  // the host edge arrays above must outlive the copies
    return 0;
}

}  // namespace sphb
