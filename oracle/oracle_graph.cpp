// TEST INFRASTRUCTURE ONLY — see oracle_common.hpp.
// Oracle for the ragged-graph passes either side of the kNN search and the level graphs:
//   F1  sph/utils/GraphUtils.cpp:197-317      symmetrizeGraph (symmetrizeGraphImpl, min-merge) + :23-95 ensureClosestPointIsSelf
//   RW3 sph/utils/SparseMatrixAlgorithms.cpp:486-561 mergeGraphNodes, :443-484 mergeNodesDataDistances
//   D9  sph/LevelSimilarities.cpp:346-389     useRandomWalksAsKnnDistances
//   B7  sph/LevelSimilarities.cpp:566-581     zero removal
//   D6  sph/utils/Similarities.cpp:414-483    euclidDistance (all-pairs branch) + sph/utils/Math.cpp:167-172 symmetricHausdorffDistance
//   D7  sph/LevelSimilarities.cpp:211-252     computeExactKnn over a component metric
// Graphs cross as (rowptr int64[n+1], ids int64[nnz], dist float[nnz]) — utils::Graph's knnIndices / knnDistances / offsets
// (sph/utils/Graph.hpp:155-273).  Pinned by tests/UtilsTests.cpp:562-660 (symmetrize) and tests/HierarchyTest.cpp:902-979 (merge).
#include "oracle_common.hpp"

#include <algorithm>
#include <cmath>
#include <limits>
#include <map>
#include <unordered_map>

namespace {

using Pair = std::pair<int64_t, float>;
using Rows = std::vector<std::vector<Pair>>;

inline bool pair_less(const Pair& a, const Pair& b) {  // the comparator of GraphUtils.cpp:283-287 / SparseMatrixAlgorithms.cpp:529-534
    if (a.second == b.second) return a.first < b.first;
    return a.second < b.second;
}

// GraphUtils.cpp:23-95 on one ragged row.
void ensure_self_first(std::vector<Pair>& row, int64_t p) {
    if (row.empty() || row[0].first == p) return;
    size_t swap_pos = row.size() - 1;
    size_t found = row.size();
    for (size_t t = 0; t < row.size(); ++t)
        if (row[t].first == p) {
            found = t;
            break;
        }
    if (found < row.size()) swap_pos = found;
    else if (std::all_of(row.begin(), row.end(), [](const Pair& q) { return q.second == 0; })) row[swap_pos].first = p;
    else {
        for (size_t i = row.size() - 1; i > 0; --i) row[i] = row[i - 1];
        swap_pos = 0;
        row[0] = {p, 0.f};
    }
    std::swap(row[0], row[swap_pos]);
}

int64_t emit(const Rows& rows, int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap) {
    int64_t e = 0;
    out_rowptr[0] = 0;
    for (size_t i = 0; i < rows.size(); ++i) {
        for (auto& q : rows[i]) {
            if (e >= cap) return -1;
            out_ids[e] = q.first;
            out_dist[e] = q.second;
            ++e;
        }
        out_rowptr[i + 1] = e;
    }
    return e;
}

inline int64_t is_direct_neighbor(const int64_t* rowptr, const int64_t* ids, int64_t a, int64_t b) {  // Graph.cpp:126-136
    for (int64_t e = rowptr[a]; e < rowptr[a + 1]; ++e)
        if (ids[e] == b) return e - rowptr[a];
    return -1;
}

inline float fp64_l2(const float* a, const float* b, int d) {  // the Stage-A contract distance (oracle_knn.cpp pair_key) + sqrt
    double acc = 0.0;
    for (int c = 0; c < d; ++c) {
        const double diff = static_cast<double>(a[c]) - static_cast<double>(b[c]);
        acc = std::fma(diff, diff, acc);
    }
    return std::sqrt(static_cast<float>(acc));
}

}  // namespace

// F1.  Literal symmetrizeGraphImpl: per (i, j) the first stored d_ij, min-merged with the first stored d_ji; emplace never
// overwrites; rows sorted by (distance, index); ensureClosestPointIsSelf.  Returns nnz or -1 (capacity).
ORC_API int64_t orc_symmetrize_graph(int64_t n, const int64_t* rowptr, const int64_t* ids, const float* dist, int64_t* out_rowptr,
                                     int64_t* out_ids, float* out_dist, int64_t cap) {
    std::vector<std::map<int64_t, float>> sym(static_cast<size_t>(n));  // iteration order is irrelevant: rows are sorted afterwards
    std::vector<std::vector<Pair>> empty(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e) {
            const int64_t j = ids[e];
            float value = dist[e];
            const int64_t dn = is_direct_neighbor(rowptr, ids, j, i);
            if (dn == -1) empty[i].push_back({j, value});
            else value = std::min(value, dist[rowptr[j] + dn]);
            sym[i].emplace(j, value);
        }
    }
    for (int64_t i = 0; i < n; ++i)
        for (auto& q : empty[i]) sym[q.first].emplace(i, q.second);
    Rows rows(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        rows[i].assign(sym[i].begin(), sym[i].end());
        std::sort(rows[i].begin(), rows[i].end(), pair_less);
        ensure_self_first(rows[i], i);
    }
    return emit(rows, out_rowptr, out_ids, out_dist, cap);
}

// RW3, graph form (mergeGraphNodes): min over all child entries that map to (parent(r), parent(c)).
ORC_API int64_t orc_merge_graph_nodes(int64_t n, const int64_t* rowptr, const int64_t* ids, const float* dist, const uint64_t* parents,
                                      int64_t n_merged, int64_t* out_rowptr, int64_t* out_ids, float* out_dist, int64_t cap) {
    std::vector<std::unordered_map<int64_t, float>> merged(static_cast<size_t>(n_merged));
    for (int64_t c = 0; c < n; ++c) {
        auto& into = merged[parents[c]];
        for (int64_t e = rowptr[c]; e < rowptr[c + 1]; ++e) {
            const int64_t col = static_cast<int64_t>(parents[ids[e]]);
            auto it = into.find(col);
            if (it == into.end()) into[col] = dist[e];
            else if (it->second > dist[e]) it->second = dist[e];
        }
    }
    Rows rows(static_cast<size_t>(n_merged));
    for (int64_t m = 0; m < n_merged; ++m) {
        rows[m].assign(merged[m].begin(), merged[m].end());
        std::sort(rows[m].begin(), rows[m].end(), pair_less);
        ensure_self_first(rows[m], m);
    }
    return emit(rows, out_rowptr, out_ids, out_dist, cap);
}

// RW3, sparse form (mergeNodesDataDistances): Eigen rows, a stored or missing 0 counts as "empty" and is overwritten by
// whatever comes next (`coeff == 0`), otherwise keep the smaller.  Entries are visited row by row in column order.
ORC_API int64_t orc_merge_data_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, const uint64_t* parents,
                                         int64_t n_merged, int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap) {
    std::vector<std::map<uint32_t, float>> merged(static_cast<size_t>(n_merged));
    for (int64_t c = 0; c < n; ++c) {
        auto& into = merged[parents[c]];
        for (int64_t e = rowptr[c]; e < rowptr[c + 1]; ++e) {
            const uint32_t mc = static_cast<uint32_t>(parents[col[e]]);
            auto it = into.find(mc);
            const float cur = it == into.end() ? 0.f : it->second;
            if (cur == 0.f) into[mc] = val[e];
            else if (cur > val[e]) it->second = val[e];
        }
    }
    int64_t e = 0;
    out_rowptr[0] = 0;
    for (int64_t m = 0; m < n_merged; ++m) {
        for (auto& kv : merged[m]) {
            if (e >= cap) return -1;
            out_col[e] = kv.first;
            out_val[e] = kv.second;
            ++e;
        }
        out_rowptr[m + 1] = e;
    }
    return e;
}

// D9.  Non-zero entries become (column, 1 - value), stable-sorted ascending by distance (ties keep column order).
ORC_API int64_t orc_rw_as_knn_distances(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t* out_rowptr,
                                        int64_t* out_ids, float* out_dist, int64_t cap) {
    Rows rows(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e)
            if (val[e] != 0.f) rows[i].push_back({static_cast<int64_t>(col[e]), 1.f - val[e]});
        std::stable_sort(rows[i].begin(), rows[i].end(), [](const Pair& a, const Pair& b) { return a.second < b.second; });
    }
    return emit(rows, out_rowptr, out_ids, out_dist, cap);
}

// B7.  Erase exact zeros from every row.
ORC_API int64_t orc_remove_zeros(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int64_t* out_rowptr, uint32_t* out_col,
                                 float* out_val) {
    int64_t o = 0;
    out_rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e)
            if (val[e] != 0.f) {
                out_col[o] = col[e];
                out_val[o] = val[e];
                ++o;
            }
        out_rowptr[i + 1] = o;
    }
    return o;
}

// D6.  Symmetric Hausdorff distance between the pixel sets of two components (all-pairs branch; the sub-sampled branch of
// the reference draws from an unseeded random_device and is reproduced by the caller passing the sampled lists).
// A pair of single pixels reduces to the plain L2 distance, which is the level-0 branch (:425-430).
ORC_API void orc_euclid_hausdorff(const float* x, int d, const int64_t* comp_ptr, const int64_t* comp_pix, const int64_t* pa, const int64_t* pb,
                                  int64_t npairs, float* out) {
#pragma omp parallel for schedule(dynamic, 8)
    for (int64_t p = 0; p < npairs; ++p) {
        const int64_t a0 = comp_ptr[pa[p]], a1 = comp_ptr[pa[p] + 1], b0 = comp_ptr[pb[p]], b1 = comp_ptr[pb[p] + 1];
        if (a1 == a0 || b1 == b0) {
            out[p] = 0.f;
            continue;
        }
        std::vector<float> colmin(static_cast<size_t>(b1 - b0), std::numeric_limits<float>::infinity());
        float rowmax = -std::numeric_limits<float>::infinity();
        for (int64_t i = a0; i < a1; ++i) {
            float rmin = std::numeric_limits<float>::infinity();
            for (int64_t j = b0; j < b1; ++j) {
                const float v = fp64_l2(x + comp_pix[i] * d, x + comp_pix[j] * d, d);
                rmin = std::min(rmin, v);
                colmin[j - b0] = std::min(colmin[j - b0], v);
            }
            rowmax = std::max(rowmax, rmin);
        }
        const float colmax = *std::max_element(colmin.begin(), colmin.end());
        out[p] = std::max(rowmax, colmax);
    }
}

// D7 for the random-walk metrics.  The hnswlib spaces the reference searches in do NOT call componentDistance: the
// distance of (i, j), i == j included, is 1 - randomWalksBhattacharyya(i, j) (mode 0,
// sph/utils/NeighborWalksBhattacharyyaSpace.hpp:28-41) or 1 - randomWalksSingleOverlap(i, j) = 1 - A[i][j], row i only
// (mode 1, sph/utils/NeighborWalksSingleOverlapSpace.hpp:26-36); per row the k smallest.  The reference sorts with an
// unstable std::sort on the distance alone (ties unordered); the contract here is (distance, index).
ORC_API void orc_level_knn_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int mode, int64_t k, int64_t* out_ids,
                                 float* out_dist) {
    auto coeff = [&](int64_t r, uint32_t c) -> float {
        const uint32_t* b = col + rowptr[r];
        const uint32_t* e = col + rowptr[r + 1];
        const uint32_t* it = std::lower_bound(b, e, c);
        return (it != e && *it == c) ? val[it - col] : 0.f;
    };
#pragma omp parallel
    {
        std::vector<Pair> all(static_cast<size_t>(n));
#pragma omp for schedule(dynamic, 4)
        for (int64_t i = 0; i < n; ++i) {
            for (int64_t j = 0; j < n; ++j) {
                float sim;
                if (mode == 1) sim = coeff(i, static_cast<uint32_t>(j));
                else {
                    sim = 0.f;
                    for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e) {
                        const float v2 = coeff(j, col[e]);
                        if (v2 > 0.0f) sim += std::sqrt(val[e] * v2);
                    }
                }
                all[j] = {j, 1 - sim};
            }
            const int64_t kk = std::min(k, n);
            std::partial_sort(all.begin(), all.begin() + kk, all.end(), pair_less);
            for (int64_t t = 0; t < k; ++t) {
                out_ids[i * k + t] = t < kk ? all[t].first : -1;
                out_dist[i * k + t] = t < kk ? all[t].second : -1.f;
            }
        }
    }
}

// D7 in the neighbour-overlap metric (NeighborOverlapSpace, sph/utils/NeighborOverlapSpace.hpp:31-43): distance of (i, j),
// i == j included, = 1 - simNeighborOverlap (sph/utils/Similarities.cpp:174-228): U_c = sorted-unique union of the kNN
// lists of the pixels of component c, sim = card(U_i ∩ U_j) / min(card U_i, card U_j) (0 when a union is empty).
ORC_API void orc_level_knn_overlap(const int64_t* knn_ids, int k_data, int64_t ncomp, const int64_t* comp_ptr, const int64_t* comp_pix, int64_t k,
                                   int64_t* out_ids, float* out_dist) {
    std::vector<std::vector<uint64_t>> uni(static_cast<size_t>(ncomp));
    for (int64_t c = 0; c < ncomp; ++c) {
        auto& u = uni[c];
        for (int64_t e = comp_ptr[c]; e < comp_ptr[c + 1]; ++e)
            for (int t = 0; t < k_data; ++t) u.push_back(static_cast<uint64_t>(knn_ids[comp_pix[e] * k_data + t]));
        std::sort(u.begin(), u.end());
        u.erase(std::unique(u.begin(), u.end()), u.end());
    }
#pragma omp parallel
    {
        std::vector<Pair> all(static_cast<size_t>(ncomp));
#pragma omp for schedule(dynamic, 4)
        for (int64_t i = 0; i < ncomp; ++i) {
            for (int64_t j = 0; j < ncomp; ++j) {
                const auto &a = uni[i], &b = uni[j];
                uint64_t cnt = 0;
                for (size_t x = 0, y = 0; x < a.size() && y < b.size();) {
                    if (a[x] < b[y]) ++x;
                    else if (b[y] < a[x]) ++y;
                    else { ++cnt; ++x; ++y; }
                }
                const size_t mn = a.size() <= b.size() ? a.size() : b.size();
                const float sim = mn > 0 ? static_cast<float>(cnt) / mn : 0;
                all[j] = {j, 1.f - sim};
            }
            const int64_t kk = std::min(k, ncomp);
            std::partial_sort(all.begin(), all.begin() + kk, all.end(), pair_less);
            for (int64_t t = 0; t < k; ++t) {
                out_ids[i * k + t] = t < kk ? all[t].first : -1;
                out_dist[i * k + t] = t < kk ? all[t].second : -1.f;
            }
        }
    }
}

// F2.  Connected components with edges taken as undirected (utils::labelGraphWeakComponents, sph/utils/GraphUtils.cpp:371-407):
// the reference symmetrises and runs boost::strong_components, whose DFS starts at vertex 0, 1, 2, ... and therefore numbers
// the components in ascending order of their lowest vertex.  Returns the number of components.
ORC_API int64_t orc_weak_components(int64_t n, const int64_t* rowptr, const int64_t* ids, int64_t* labels) {
    std::vector<int64_t> parent(static_cast<size_t>(n));
    for (int64_t v = 0; v < n; ++v) parent[v] = v;
    auto find = [&](int64_t a) {
        while (parent[a] != a) {
            parent[a] = parent[parent[a]];
            a = parent[a];
        }
        return a;
    };
    for (int64_t v = 0; v < n; ++v)
        for (int64_t e = rowptr[v]; e < rowptr[v + 1]; ++e) {
            const int64_t a = find(v), b = find(ids[e]);
            if (a != b) parent[std::max(a, b)] = std::min(a, b);
        }
    int64_t next = 0;
    std::vector<int64_t> number(static_cast<size_t>(n), -1);
    for (int64_t v = 0; v < n; ++v) {
        const int64_t r = find(v);
        if (number[r] < 0) number[r] = next++;
        labels[v] = number[r];
    }
    return next;
}

// F3.  Component bridging, NearestNeighbors::connectComponents (sph/NearestNeighbors.cpp:494-861) — the two loops the device
// replaces, restated literally (sequential, brute force).
// computeCentersOfMass (:655-680) with utils::computeMean (sph/utils/Math.hpp:29-48): fp64 sums over the component's points
// in ascending id order (getComponentValues :574-590), / count, cast to float.
ORC_API void orc_component_centers(const float* x, int64_t n, int d, const int64_t* labels, int64_t ncomp, float* centers) {
    std::vector<double> sum(static_cast<size_t>(ncomp) * d, 0.0);
    std::vector<int64_t> cnt(static_cast<size_t>(ncomp), 0);
    for (int64_t i = 0; i < n; ++i) {
        const int64_t c = labels[i];
        ++cnt[c];
        for (int t = 0; t < d; ++t) sum[c * d + t] += static_cast<double>(x[i * d + t]);
    }
    for (int64_t c = 0; c < ncomp; ++c)
        for (int t = 0; t < d; ++t) centers[c * d + t] = cnt[c] ? static_cast<float>(sum[c * d + t] / static_cast<double>(cnt[c])) : 0.f;
}

// insertConnectionsBetweenComponents (:592-653) with an exact per-point search (faiss::IndexFlat, the npA <= 3900 branch :607-616;
// the distance of a pair is the Stage-A contract value: fp64 FMA chain rounded once to fp32, squared L2 as faiss returns it):
// for every point of B (ascending id) its nearest point of A (ties: lowest id), then the first B point with the strictly
// smallest distance (:628-636).
ORC_API void orc_bridge_components(const float* x, int64_t n, int d, const int64_t* labels, const int64_t* edges, int64_t m, int64_t* ids_a,
                                   int64_t* ids_b, float* dist) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t e = 0; e < m; ++e) {
        const int64_t ca = edges[2 * e], cb = edges[2 * e + 1];
        std::vector<int64_t> A, B;
        for (int64_t i = 0; i < n; ++i) {
            if (labels[i] == ca) A.push_back(i);
            if (labels[i] == cb) B.push_back(i);
        }
        float best = std::numeric_limits<float>::max();
        int64_t ba = -1, bb = -1;
        for (int64_t b : B) {
            float nearest = std::numeric_limits<float>::infinity();
            int64_t na = -1;
            for (int64_t a : A) {
                double acc = 0.0;
                for (int c = 0; c < d; ++c) {
                    const double diff = static_cast<double>(x[b * d + c]) - static_cast<double>(x[a * d + c]);
                    acc = std::fma(diff, diff, acc);
                }
                const float v = static_cast<float>(acc);
                if (v < nearest) { nearest = v; na = a; }
            }
            if (nearest < best) { best = nearest; ba = na; bb = b; }
        }
        ids_a[e] = ba;
        ids_b[e] = bb;
        dist[e] = best;
    }
}
