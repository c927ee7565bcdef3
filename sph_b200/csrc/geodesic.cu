// geodesic component distances — utils::geodesicDistance (sph/utils/Similarities.cpp:234-309) over
// utils::computeShortestPath (sph/utils/ShortestPath.cpp:100-171), the metric of GEO_CENTROID / GEO_WALKS
// (sph/ImageHierarchy.cpp:192-249, sph/utils/GeodesicPathSpace.hpp).
//
// The reference answers every (start, end) pair with its own search on the host (Boost A* with the L2 heuristic, Dijkstra when
// A* fails, sph/utils/AStarBoost.cpp:157-257) behind a mutex-guarded cache.  Here the pairs of a call are grouped by their
// first component; every point of that component becomes ONE single-source search whose targets are the points of all the
// components it is paired with, so a source shared by several pairs is searched once.  One CTA per search:
//
//   dist[n]   tentative distances as fp32 bit patterns (non-negative floats order like unsigned integers): atomicMin relaxes
//   bm[n/32]  one bit per vertex whose distance changed since it was last expanded
//   loop      scan the bitmap; a flagged vertex u is   dropped   if dist[u] >= bound (the largest tentative distance of a
//                                                                 target: nothing behind it can improve a target),
//                                                      deferred  if dist[u] >  thr   (bucket limit, delta-stepping style),
//                                                      expanded  otherwise: a warp relaxes its out-edges, 32 per step;
//             when a pass relaxes nothing up to thr, thr moves to (smallest deferred distance + delta), which always admits
//             that vertex; the search ends when nothing is flagged any more.
//
// The fixed point is d[v] = min over paths of the left-to-right fp32 sum of the edge weights — fp32 addition is monotone, so
// this is bit for bit what a sequential Dijkstra (boost::dijkstra_shortest_paths, float distances, combine a + b) settles,
// whatever the order of the relaxations; dropping at `bound` only skips vertices that cannot lie on a shortest path to a
// target.  With true L2 edge weights (L2squared = false, the reference default) the reference's A* heuristic is admissible
// and consistent and returns the same distance; parity against Boost itself is unpinned.
//
// Around the search, as computeShortestPath does: points in different connected components (labels) have no path; if the end
// is a stored neighbour of the start, the stored distance of the first such slot is the answer and no search is made (a
// search runs only for sources with at least one target that is neither).  "No path" is -1 for sphb_geodesic_distances and
// FLT_MAX inside the Hausdorff matrix (:278-281, :245-252).  The sub-sampled branch (:288-303, unseeded random_device) is the
// caller's: it passes the sampled point lists.
#include "../../include/sphb.h"
#include "common.cuh"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <numeric>
#include <vector>

namespace sphb {

constexpr int GEO_THREADS = 256;
constexpr int GEO_WARPS = GEO_THREADS / 32;
constexpr int GEO_ROWCAP = 1024;            // neighbour ids of the source staged in shared memory up to this row length
constexpr unsigned GEO_INF = 0x7f800000u;   // +inf: unreached
constexpr unsigned GEO_MAXF = 0x7f7fffffu;  // FLT_MAX: "no path" inside the Hausdorff matrix
#ifndef SPHB_GEO_DELTA_MUL
#define SPHB_GEO_DELTA_MUL 1.0f  // bucket width in mean edge weights
#endif

struct GeoGraph {
    int64_t n;
    const int64_t* rowptr;  // [n + 1]
    const int64_t* ids;
    const float* dist;
    const int64_t* labels;  // [n] or null
};
struct GeoComps {
    const int64_t* ptr;  // null: component c is the single point c
    const int64_t* pix;
};
struct GeoTasks {
    const int64_t* group_comp;   // [G] first component of the group's pairs
    const int64_t* group_pairs;  // [G + 1] range into pair_id
    const int64_t* pair_id;      // pair indices, grouped
    const int64_t* pb;           // [npairs] second component of every pair
    const int64_t* row_off;      // [npairs] offset of the pair's row minima (one per point of its first component)
    const int64_t* col_off;      // [npairs] offset of the pair's column minima (one per point of its second component)
    const int* task_group;       // [T]
    const int* task_local;       // [T] index of the source inside its component
    int64_t T;
};

__device__ __forceinline__ void geo_range(const GeoComps& C, int64_t c, int64_t& b, int64_t& e) {
    if (C.ptr) {
        b = C.ptr[c];
        e = C.ptr[c + 1];
    } else {
        b = c;
        e = c + 1;
    }
}
__device__ __forceinline__ int64_t geo_point(const GeoComps& C, int64_t i) { return C.pix ? C.pix[i] : i; }

// The part of computeShortestPath in front of the search (ShortestPath.cpp:113-134): 1 and *bits set when (s, t) needs none.
__device__ __forceinline__ int geo_shortcut(const GeoGraph& G, int64_t s, int64_t t, const int* s_row, int row_len, bool staged, unsigned* bits) {
    if (G.labels && G.labels[s] != G.labels[t]) {
        *bits = GEO_MAXF;
        return 1;
    }
    const int64_t r0 = G.rowptr[s];
    if (staged) {
        for (int j = 0; j < row_len; ++j)
            if (s_row[j] == static_cast<int>(t)) {
                *bits = __float_as_uint(G.dist[r0 + j]);
                return 1;
            }
    } else {
        for (int64_t e = r0; e < r0 + row_len; ++e)
            if (G.ids[e] == t) {
                *bits = __float_as_uint(G.dist[e]);
                return 1;
            }
    }
    return 0;
}

__device__ __forceinline__ unsigned geo_block_min(unsigned v, unsigned* s_red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned m = s_red[0];
#pragma unroll
    for (int w = 1; w < GEO_WARPS; ++w) m = min(m, s_red[w]);
    return m;
}

__global__ void __launch_bounds__(GEO_THREADS) geo_search_kernel(GeoGraph G, GeoComps C, GeoTasks K, float delta, unsigned* __restrict__ dist_ws,
                                                                 unsigned* __restrict__ bm_ws, unsigned* __restrict__ rowmin,
                                                                 unsigned* __restrict__ colmin, unsigned long long* __restrict__ stats) {
    __shared__ int s_row[GEO_ROWCAP];
    __shared__ unsigned s_red[GEO_WARPS];
    __shared__ unsigned s_changed, s_minpend, s_bound, s_unresolved;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int64_t n = G.n, nwords = (n + 31) >> 5;
    unsigned* __restrict__ dist = dist_ws + static_cast<size_t>(blockIdx.x) * static_cast<size_t>(n);
    unsigned* __restrict__ bm = bm_ws + static_cast<size_t>(blockIdx.x) * static_cast<size_t>(nwords);
    for (int64_t task = blockIdx.x; task < K.T; task += gridDim.x) {
        const int g = K.task_group[task];
        int64_t a0, a1;
        geo_range(C, K.group_comp[g], a0, a1);
        const int64_t s = geo_point(C, a0 + K.task_local[task]);
        const int64_t p0 = K.group_pairs[g], p1 = K.group_pairs[g + 1];
        const int64_t r0 = G.rowptr[s];
        const int row_len = static_cast<int>(G.rowptr[s + 1] - r0);
        const bool staged = row_len <= GEO_ROWCAP;
        __syncthreads();  // the previous task is done with s_row / the flags
        if (staged)
            for (int j = tid; j < row_len; j += GEO_THREADS) s_row[j] = static_cast<int>(G.ids[r0 + j]);
        if (tid == 0) s_unresolved = 0;
        __syncthreads();
        // ---- which targets need a search at all
        unsigned mine = 0;
        for (int64_t q = p0; q < p1; ++q) {
            int64_t b0, b1;
            geo_range(C, K.pb[K.pair_id[q]], b0, b1);
            for (int64_t j = b0 + tid; j < b1; j += GEO_THREADS) {
                unsigned bits;
                mine += geo_shortcut(G, s, geo_point(C, j), s_row, row_len, staged, &bits) ? 0u : 1u;
            }
        }
        if (mine) atomicAdd(&s_unresolved, mine);
        __syncthreads();
        const bool search = s_unresolved != 0;
        if (search) {
            // ---- single-source search from s
            for (int64_t i = tid; i < n; i += GEO_THREADS) dist[i] = GEO_INF;
            for (int64_t i = tid; i < nwords; i += GEO_THREADS) bm[i] = 0u;
            __syncthreads();
            if (tid == 0) {
                dist[s] = 0u;
                bm[s >> 5] = 1u << (s & 31);
                __threadfence();
            }
            float thr = delta;
            unsigned long long expanded = 0, passes = 0;
            for (;;) {
                __syncthreads();
                if (tid == 0) {
                    s_changed = 0;
                    s_minpend = GEO_INF;
                    s_bound = 0;
                }
                __syncthreads();
                // bound: the largest tentative distance of a target that can be reached at all
                unsigned bmax = 0;
                for (int64_t q = p0; q < p1; ++q) {
                    int64_t b0, b1;
                    geo_range(C, K.pb[K.pair_id[q]], b0, b1);
                    for (int64_t j = b0 + tid; j < b1; j += GEO_THREADS) {
                        const int64_t t = geo_point(C, j);
                        if (G.labels && G.labels[s] != G.labels[t]) continue;
                        bmax = max(bmax, __ldcg(dist + t));
                    }
                }
                if (bmax) atomicMax(&s_bound, bmax);
                __syncthreads();
                const unsigned bound_bits = s_bound;
                const float bound = __uint_as_float(bound_bits);
                ++passes;
                for (int64_t base = static_cast<int64_t>(warp) * 32; base < nwords; base += GEO_WARPS * 32) {
                    const int64_t wi = base + lane;
                    const unsigned word = wi < nwords ? __ldcg(bm + wi) : 0u;
                    unsigned nz = __ballot_sync(0xffffffffu, word != 0u);
                    while (nz) {
                        const int b = __ffs(nz) - 1;
                        nz &= nz - 1;
                        const int64_t w = base + b;
                        unsigned bits = 0;
                        if (lane == 0) {
                            bits = atomicExch(bm + w, 0u);
                            __threadfence();
                        }
                        bits = __shfl_sync(0xffffffffu, bits, 0);
                        unsigned keep = 0;
                        while (bits) {
                            const int bit = __ffs(bits) - 1;
                            bits &= bits - 1;
                            const int64_t u = (w << 5) + bit;
                            unsigned du_bits = 0;
                            if (lane == 0) du_bits = __ldcg(dist + u);
                            du_bits = __shfl_sync(0xffffffffu, du_bits, 0);
                            const float du = __uint_as_float(du_bits);
                            if (du_bits >= bound_bits) continue;  // cannot improve a target
                            if (!(du <= thr)) {                   // a later bucket
                                keep |= 1u << bit;
                                if (lane == 0) atomicMin(&s_minpend, du_bits);
                                continue;
                            }
                            ++expanded;
                            const int64_t e0 = G.rowptr[u], e1 = G.rowptr[u + 1];
                            for (int64_t e = e0 + lane; e < e1; e += 32) {
                                const int64_t v = G.ids[e];
                                if (v == u || v < 0 || v >= n) continue;  // createBoostGraph: point != neighbour (AStarBoost.cpp:116)
                                const float nd = __fadd_rn(du, G.dist[e]);
                                const unsigned nd_bits = __float_as_uint(nd);
                                // two filters in front of the L2 atomic (most relaxations fail): a distance at or above the bound can
                                // improve no target, and a plain load that already shows a smaller or equal distance settles it —
                                // distances only decrease, so a stale load errs towards doing the atomic
                                if (nd_bits >= bound_bits || nd_bits >= __ldcg(dist + v)) continue;
                                const unsigned old = atomicMin(dist + v, nd_bits);
                                if (nd_bits < old) {
                                    __threadfence();
                                    atomicOr(bm + (v >> 5), 1u << (v & 31));
                                    if (nd <= thr) s_changed = 1u;
                                    else atomicMin(&s_minpend, nd_bits);
                                }
                            }
                        }
                        if (keep && lane == 0) atomicOr(bm + w, keep);
                    }
                }
                __syncthreads();
                if (s_changed == 0u) {
                    if (s_minpend == GEO_INF) break;
                    thr = __fadd_rn(__uint_as_float(s_minpend), delta);
                }
            }
            if (stats && lane == 0) {
                atomicAdd(stats + 0, expanded);
                if (tid == 0) {
                    atomicAdd(stats + 1, passes);
                    atomicAdd(stats + 2, 1ull);
                }
            }
        }
        __syncthreads();
        // ---- the source's row of every pair matrix: row minimum, and its contribution to the column minima
        for (int64_t q = p0; q < p1; ++q) {
            const int64_t p = K.pair_id[q];
            int64_t b0, b1;
            geo_range(C, K.pb[p], b0, b1);
            unsigned rmin = 0xffffffffu;
            for (int64_t j = b0 + tid; j < b1; j += GEO_THREADS) {
                const int64_t t = geo_point(C, j);
                unsigned bits;
                if (!geo_shortcut(G, s, t, s_row, row_len, staged, &bits)) {
                    bits = __ldcg(dist + t);
                    if (bits >= GEO_INF) bits = GEO_MAXF;
                }
                rmin = min(rmin, bits);
                atomicMin(colmin + K.col_off[p] + (j - b0), bits);
            }
            rmin = geo_block_min(rmin, s_red);
            if (tid == 0) rowmin[K.row_off[p] + K.task_local[task]] = rmin;
        }
    }
}

// symmetricHausdorffDistance (sph/utils/Math.cpp:167-172): max(max over rows of the row minimum, max over columns of the
// column minimum); one warp per pair.  raw: a 1 x 1 "no path" entry goes back as -1 (computeShortestPath's own value).
__global__ void __launch_bounds__(256) geo_finish_kernel(GeoComps C, const int64_t* __restrict__ pa, const int64_t* __restrict__ pb,
                                                         const int64_t* __restrict__ row_off, const int64_t* __restrict__ col_off,
                                                         const unsigned* __restrict__ rowmin, const unsigned* __restrict__ colmin, int64_t npairs,
                                                         int raw, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t p = warp; p < npairs; p += nwarps) {
        int64_t a0, a1, b0, b1;
        geo_range(C, pa[p], a0, a1);
        geo_range(C, pb[p], b0, b1);
        unsigned m = 0;
        for (int64_t i = lane; i < a1 - a0; i += 32) m = max(m, rowmin[row_off[p] + i]);
        for (int64_t j = lane; j < b1 - b0; j += 32) m = max(m, colmin[col_off[p] + j]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) {
            float v = (a1 > a0 && b1 > b0) ? __uint_as_float(m) : 0.f;
            if (raw && m == GEO_MAXF) v = -1.f;
            out[p] = v;
        }
    }
}

__global__ void geo_fill_kernel(unsigned* __restrict__ p, int64_t count, unsigned v) {
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += static_cast<int64_t>(gridDim.x) * blockDim.x) p[i] = v;
}

namespace {
template <class T>
int geo_up(DevBuf<T>& b, const T* h, int64_t count, cudaStream_t s) {
    SPHB_CUDA(b.alloc(static_cast<size_t>(std::max<int64_t>(count, 1)), s));
    if (count > 0) SPHB_CUDA(cudaMemcpyAsync(b.p, h, count * sizeof(T), cudaMemcpyHostToDevice, s));
    return 0;
}
thread_local int64_t g_geo_stats[4] = {0, 0, 0, 0};  // vertices expanded, passes, searches, sources
}  // namespace

// The kNN graph resident on the device: uploaded and validated once, used by any number of calls (a hierarchy asks for the
// distances of every level on the same graph; the upload is 121 MB at C2 size, about half of a level-0 call).
struct GeoGraphDev {
    int64_t n = 0, nnz = 0;
    float delta = 1.f;  // bucket width of the searches: the mean positive edge weight
    bool has_labels = false;
    DevBuf<int64_t> rowptr, ids, labels;
    DevBuf<float> dist;
};

static int geo_graph_upload(GeoGraphDev& g, const char* who, int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist,
                            const int64_t* labels) {
    if (n <= 0 || n >= (int64_t(1) << 31) || !ids || !dist || (!rowptr && k <= 0)) return fail(SPHB_ERR_INVALID, std::string(who) + ": bad arguments");
    std::vector<int64_t> rp_own;
    if (!rowptr) {
        rp_own.resize(static_cast<size_t>(n) + 1);
        for (int64_t r = 0; r <= n; ++r) rp_own[r] = r * k;
        rowptr = rp_own.data();
    } else {
        if (rowptr[0] != 0) return fail(SPHB_ERR_INVALID, std::string(who) + ": rowptr[0] != 0");
        for (int64_t r = 0; r < n; ++r)
            if (rowptr[r + 1] < rowptr[r]) return fail(SPHB_ERR_INVALID, std::string(who) + ": rowptr not monotone");
    }
    const int64_t nnz = rowptr[n];
    double wsum = 0.0;
    int64_t wcount = 0;
    for (int64_t e = 0; e < nnz; ++e) {
        if (!(dist[e] >= 0.f)) return fail(SPHB_ERR_INVALID, std::string(who) + ": edge weights must be non-negative numbers");
        if (dist[e] > 0.f && std::isfinite(dist[e])) {
            wsum += dist[e];
            ++wcount;
        }
    }
    cudaStream_t s = ctx().stream;
    g.n = n;
    g.nnz = nnz;
    g.delta = wcount ? SPHB_GEO_DELTA_MUL * static_cast<float>(wsum / static_cast<double>(wcount)) : 1.f;
    g.has_labels = labels != nullptr;
    if (int rc = geo_up(g.rowptr, rowptr, n + 1, s)) return rc;
    if (int rc = geo_up(g.ids, ids, nnz, s)) return rc;
    if (int rc = geo_up(g.dist, dist, nnz, s)) return rc;
    if (labels)
        if (int rc = geo_up(g.labels, labels, n, s)) return rc;
    SPHB_CUDA(cudaStreamSynchronize(s));  // rp_own (and the caller's arrays) may go away after the return
    return SPHB_OK;
}

// Host driver on a resident graph.  comp_ptr == NULL: components are single points (pa / pb are point ids).
static int geodesic_pairs(const GeoGraphDev& gr, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, int64_t npairs, const int64_t* pa,
                          const int64_t* pb, int raw, float* out) {
    const char* who = raw ? "geodesic_distances" : "geodesic_hausdorff";
    const int64_t n = gr.n;
    if (npairs < 0 || (npairs > 0 && (!pa || !pb || !out))) return fail(SPHB_ERR_INVALID, std::string(who) + ": bad arguments");
    if (comp_ptr) {
        if (ncomp <= 0 || comp_ptr[0] != 0) return fail(SPHB_ERR_INVALID, std::string(who) + ": bad component lists");
        for (int64_t c = 0; c < ncomp; ++c)
            if (comp_ptr[c + 1] < comp_ptr[c]) return fail(SPHB_ERR_INVALID, std::string(who) + ": comp_ptr not monotone");
        if (comp_ptr[ncomp] > 0 && !comp_pix) return fail(SPHB_ERR_INVALID, std::string(who) + ": bad component lists");
        for (int64_t i = 0; i < comp_ptr[ncomp]; ++i)
            if (comp_pix[i] < 0 || comp_pix[i] >= n) return fail(SPHB_ERR_INVALID, std::string(who) + ": point id out of range");
    } else {
        ncomp = n;
    }
    for (int64_t p = 0; p < npairs; ++p)
        if (pa[p] < 0 || pa[p] >= ncomp || pb[p] < 0 || pb[p] >= ncomp) return fail(SPHB_ERR_INVALID, std::string(who) + ": id out of range");
    if (npairs == 0) return SPHB_OK;
    auto csize = [&](int64_t c) -> int64_t { return comp_ptr ? comp_ptr[c + 1] - comp_ptr[c] : 1; };

    // ---- pairs grouped by their first component; one search task per point of that component
    std::vector<int64_t> order(static_cast<size_t>(npairs));
    std::iota(order.begin(), order.end(), int64_t(0));
    std::stable_sort(order.begin(), order.end(), [&](int64_t x, int64_t y) { return pa[x] < pa[y]; });
    std::vector<int64_t> group_comp, group_pairs, row_off(static_cast<size_t>(npairs)), col_off(static_cast<size_t>(npairs));
    std::vector<int> task_group, task_local;
    int64_t rows = 0, cols = 0;
    for (int64_t q = 0; q < npairs; ++q) {
        const int64_t p = order[q];
        if (q == 0 || pa[p] != pa[order[q - 1]]) {
            if (group_comp.size() >= size_t(0x7fffffff)) return fail(SPHB_ERR_UNSUPPORTED, std::string(who) + ": too many groups");
            group_comp.push_back(pa[p]);
            group_pairs.push_back(q);
            const int64_t sz = csize(pa[p]);
            if (sz >= (int64_t(1) << 31)) return fail(SPHB_ERR_UNSUPPORTED, std::string(who) + ": component too large");
            for (int64_t i = 0; i < sz; ++i) {
                task_group.push_back(static_cast<int>(group_comp.size() - 1));
                task_local.push_back(static_cast<int>(i));
            }
        }
        row_off[p] = rows;
        col_off[p] = cols;
        rows += csize(pa[p]);
        cols += csize(pb[p]);
    }
    group_pairs.push_back(npairs);
    const int64_t T = static_cast<int64_t>(task_group.size());
    const float delta = gr.delta;

    cudaStream_t s = ctx().stream;
    DevBuf<int64_t> d_cptr, d_cpix, d_pa, d_pb, d_gc, d_gp, d_order, d_roff, d_coff;
    DevBuf<float> d_out;
    DevBuf<int> d_tg, d_tl;
    DevBuf<unsigned> d_rowmin, d_colmin, d_dw, d_bm;
    DevBuf<unsigned long long> d_stats;
    if (comp_ptr) {
        if (int rc = geo_up(d_cptr, comp_ptr, ncomp + 1, s)) return rc;
        if (int rc = geo_up(d_cpix, comp_pix, comp_ptr[ncomp], s)) return rc;
    }
    if (int rc = geo_up(d_pa, pa, npairs, s)) return rc;
    if (int rc = geo_up(d_pb, pb, npairs, s)) return rc;
    if (int rc = geo_up(d_gc, group_comp.data(), static_cast<int64_t>(group_comp.size()), s)) return rc;
    if (int rc = geo_up(d_gp, group_pairs.data(), static_cast<int64_t>(group_pairs.size()), s)) return rc;
    if (int rc = geo_up(d_order, order.data(), npairs, s)) return rc;
    if (int rc = geo_up(d_roff, row_off.data(), npairs, s)) return rc;
    if (int rc = geo_up(d_coff, col_off.data(), npairs, s)) return rc;
    if (int rc = geo_up(d_tg, task_group.data(), T, s)) return rc;
    if (int rc = geo_up(d_tl, task_local.data(), T, s)) return rc;
    SPHB_CUDA(d_rowmin.alloc(static_cast<size_t>(std::max<int64_t>(rows, 1)), s));
    SPHB_CUDA(d_colmin.alloc(static_cast<size_t>(std::max<int64_t>(cols, 1)), s));
    SPHB_CUDA(d_out.alloc(static_cast<size_t>(npairs), s));
    SPHB_CUDA(d_stats.alloc(4, s));
    SPHB_CUDA(cudaMemsetAsync(d_stats.p, 0, 4 * sizeof(unsigned long long), s));
    // one search slot per resident CTA: n distances + n / 32 flag words each, at most ~4 GB in all
    const int64_t nwords = (n + 31) >> 5;
    const int64_t slot_bytes = (n + nwords) * 4;
    int64_t grid = std::min<int64_t>(T, static_cast<int64_t>(ctx().num_sms) * 8);
    grid = std::max<int64_t>(1, std::min<int64_t>(grid, (int64_t(4) << 30) / slot_bytes));
    SPHB_CUDA(d_dw.alloc(static_cast<size_t>(grid * n), s));
    SPHB_CUDA(d_bm.alloc(static_cast<size_t>(grid * nwords), s));
    const int sms = ctx().num_sms;
    geo_fill_kernel<<<static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((cols + 255) / 256, static_cast<int64_t>(sms) * 8))), 256, 0, s>>>(d_colmin.p, cols, 0xffffffffu);
    count_launch();
    GeoGraph G{n, gr.rowptr.p, gr.ids.p, gr.dist.p, gr.has_labels ? gr.labels.p : nullptr};
    GeoComps C{comp_ptr ? d_cptr.p : nullptr, comp_ptr ? d_cpix.p : nullptr};
    GeoTasks K{d_gc.p, d_gp.p, d_order.p, d_pb.p, d_roff.p, d_coff.p, d_tg.p, d_tl.p, T};
    geo_search_kernel<<<static_cast<unsigned>(grid), GEO_THREADS, 0, s>>>(G, C, K, delta, d_dw.p, d_bm.p, d_rowmin.p, d_colmin.p, d_stats.p);
    count_launch();
    geo_finish_kernel<<<static_cast<unsigned>(std::min<int64_t>((npairs + 7) / 8, static_cast<int64_t>(sms) * 8)), 256, 0, s>>>(
        C, d_pa.p, d_pb.p, d_roff.p, d_coff.p, d_rowmin.p, d_colmin.p, npairs, raw, d_out.p);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    unsigned long long st[4] = {0, 0, 0, 0};
    SPHB_CUDA(cudaMemcpyAsync(out, d_out.p, npairs * sizeof(float), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaMemcpyAsync(st, d_stats.p, sizeof(st), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    g_geo_stats[0] = static_cast<int64_t>(st[0]);
    g_geo_stats[1] = static_cast<int64_t>(st[1]);
    g_geo_stats[2] = static_cast<int64_t>(st[2]);
    g_geo_stats[3] = T;
    return SPHB_OK;
}

}  // namespace sphb
using namespace sphb;

extern "C" {
SPHB_API int sphb_geodesic_hausdorff(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels, int64_t ncomp,
                                     const int64_t* comp_ptr, const int64_t* comp_pix, const int64_t* pa, const int64_t* pb, int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!comp_ptr) return fail(SPHB_ERR_INVALID, "geodesic_hausdorff: bad component lists");
    GeoGraphDev g;
    if (int rc = geo_graph_upload(g, "geodesic_hausdorff", n, rowptr, k, ids, dist, labels)) return rc;
    return geodesic_pairs(g, ncomp, comp_ptr, comp_pix, npairs, pa, pb, 0, out);
}
SPHB_API int sphb_geodesic_distances(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels,
                                     const int64_t* start, const int64_t* end, int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    GeoGraphDev g;
    if (int rc = geo_graph_upload(g, "geodesic_distances", n, rowptr, k, ids, dist, labels)) return rc;
    return geodesic_pairs(g, n, nullptr, nullptr, npairs, start, end, 1, out);
}
// The same two calls on a graph kept on the device between calls (ImageHierarchy asks for the distances of every level on the
// same data graph).  A handle belongs to the thread (device context) that created it.
SPHB_API int sphb_geodesic_graph_create(int64_t n, const int64_t* rowptr, int k, const int64_t* ids, const float* dist, const int64_t* labels,
                                        void** handle) {
    if (int rc = ensure_init()) return rc;
    if (!handle) return fail(SPHB_ERR_INVALID, "geodesic_graph_create: null handle");
    *handle = nullptr;
    GeoGraphDev* g = new GeoGraphDev();
    if (int rc = geo_graph_upload(*g, "geodesic_graph_create", n, rowptr, k, ids, dist, labels)) {
        delete g;
        return rc;
    }
    *handle = g;
    return SPHB_OK;
}
SPHB_API int sphb_geodesic_graph_destroy(void* handle) {
    delete static_cast<GeoGraphDev*>(handle);
    return SPHB_OK;
}
SPHB_API int sphb_geodesic_hausdorff_g(void* handle, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, const int64_t* pa, const int64_t* pb,
                                       int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!handle || !comp_ptr) return fail(SPHB_ERR_INVALID, "geodesic_hausdorff_g: bad arguments");
    return geodesic_pairs(*static_cast<GeoGraphDev*>(handle), ncomp, comp_ptr, comp_pix, npairs, pa, pb, 0, out);
}
SPHB_API int sphb_geodesic_distances_g(void* handle, const int64_t* start, const int64_t* end, int64_t npairs, float* out) {
    if (int rc = ensure_init()) return rc;
    if (!handle) return fail(SPHB_ERR_INVALID, "geodesic_distances_g: bad arguments");
    const GeoGraphDev& g = *static_cast<GeoGraphDev*>(handle);
    return geodesic_pairs(g, g.n, nullptr, nullptr, npairs, start, end, 1, out);
}
SPHB_API int sphb_geodesic_last_stats(int64_t* out4) {
    if (!out4) return fail(SPHB_ERR_INVALID, "geodesic_last_stats: null output");
    for (int i = 0; i < 4; ++i) out4[i] = g_geo_stats[i];
    return SPHB_OK;
}
}
