// TEST INFRASTRUCTURE ONLY — see oracle_common.hpp.
// Stage D oracle: per-level component similarities.
//   D1 sph/utils/SparseMatrixAlgorithms.cpp:963-995 createSimilarities dispatcher
//      -> createSimilaritiesHDI  (:1173-1378, n > 5000)
//      -> createSimilaritiesEigen (:998-1171) + convertEigenSparseVecToHDILibSparseVec (:1490-1526)
//   D2 createSimilaritiesSPH (:1380-1488)
//   D3 findTopK / findBottomK (:720-776) with MaxSizeDeque.hpp:97-200 semantics
//   D4 sph/utils/Similarities.cpp:374-396 Bhattacharyya / single overlap
//   D5 sph/utils/Similarities.cpp:174-228 neighbour overlap
// Eigen 5.0.1 (external) semantics restated here:
//   * X.pruned(ref) keeps |v| > |ref| * dummy_precision<float>() (= 1e-5f), i.e.
//     pruned(1e-4) keeps |v| > 1e-9 — NOT v >= 1e-4.  D2 uses a plain `< pruneVal` test.
//   * sparse products / dots accumulate fp32 in ascending inner-index order.
#include "oracle_common.hpp"

#include <algorithm>
#include <cmath>
#include <deque>
#include <limits>
#include <map>

namespace {
using namespace orc;

inline bool eigen_pruned_keep(float v, float ref) { return !(std::fabs(v) <= std::fabs(ref) * 1e-5f); }

// SortedMinPairSizeDequeR / SortedMaxPairSizeDequeR (MaxSizeDeque.hpp:97-200), literal.
struct MinDeque {
    std::deque<std::pair<uint32_t, float>> data;
    size_t maxSize;
    static constexpr float DEFAULT = std::numeric_limits<float>::max();
    explicit MinDeque(size_t ms) : data(ms, {0u, DEFAULT}), maxSize(ms) {}
    void insert(uint32_t index, float value) {
        if (maxSize == 0 || data.back().second <= value) return;
        if (data.size() >= maxSize) data.pop_back();
        auto it = std::lower_bound(data.begin(), data.end(), value, [](const std::pair<uint32_t, float>& p, float v) { return p.second < v; });
        data.insert(it, {index, value});
    }
    void prune() {
        data.erase(std::remove_if(data.begin(), data.end(), [](const std::pair<uint32_t, float>& e) { return e.second == DEFAULT; }), data.end());
    }
};
struct MaxDeque {
    std::deque<std::pair<uint32_t, float>> data;
    size_t maxSize;
    static constexpr float DEFAULT = std::numeric_limits<float>::min();
    explicit MaxDeque(size_t ms) : data(ms, {0u, DEFAULT}), maxSize(ms) {}
    void insert(uint32_t index, float value) {
        if (maxSize == 0 || data.back().second >= value) return;
        if (data.size() >= maxSize) data.pop_back();
        auto it = std::lower_bound(data.begin(), data.end(), value, [](const std::pair<uint32_t, float>& p, float v) { return p.second > v; });
        data.insert(it, {index, value});
    }
    void prune() {
        data.erase(std::remove_if(data.begin(), data.end(), [](const std::pair<uint32_t, float>& e) { return e.second == DEFAULT; }), data.end());
    }
};

template <class DQ>
SparseRow select_k(const SparseRow& row, size_t k) {
    DQ dq(k);
    for (auto& p : row) dq.insert(p.first, p.second);
    dq.prune();
    SparseRow out(dq.data.begin(), dq.data.end());
    std::sort(out.begin(), out.end(), [](auto& a, auto& b) { return a.first < b.first; });
    return out;
}

// fp32 merge dot of two column-sorted rows (Eigen sparse dot / SpGEMM inner order).
inline float sparse_dot(const SparseRow& a, const SparseRow& b, bool& any) {
    float res = 0.f;
    any = false;
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size()) {
        if (a[i].first == b[j].first) { res += a[i].second * b[j].second; any = true; ++i; ++j; }
        else if (a[i].first < b[j].first) ++i;
        else ++j;
    }
    return res;
}

// S = sqrt(prune(A)) [* sqrt(component size)]
SparseMat sqrt_rows(const SparseMat& in, float pruneVal, bool eigen_prune, const int64_t* comp_sizes) {
    SparseMat s(in.size());
    for (size_t i = 0; i < in.size(); ++i) {
        for (auto& p : in[i]) {
            if (eigen_prune && !eigen_pruned_keep(p.second, pruneVal)) continue;
            float v = std::sqrt(p.second);
            if (comp_sizes) v *= static_cast<float>(std::sqrt(static_cast<double>(comp_sizes[i])));
            s[i].emplace_back(p.first, v);
        }
    }
    return s;
}
}  // namespace

// D1, first restatement (explicit merge dot per candidate pair; kept as the cross-check of the row-wise form below).
// variant: 0 = auto (n > 5000 -> HDI else Eigen), 1 = force HDI, 2 = force Eigen.
// Output: fixed stride k per row (cols padded 0xFFFFFFFF), row_nnz[n].
ORC_API void orc_create_similarities_mergedot(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                     int64_t k, float pruneVal, int variant, const int64_t* comp_sizes,
                                     uint32_t* out_cols, float* out_vals, int32_t* row_nnz) {
    const SparseMat in = from_csr(n, rowptr, col, val);
    const SparseMat s = sqrt_rows(in, pruneVal, true, comp_sizes);
    const bool hdi = variant == 1 || (variant == 0 && n > 5000);
    // transpose pattern for candidate generation (which rows share a column)
    std::vector<std::vector<uint32_t>> col_rows(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        for (auto& p : s[i]) col_rows[p.first].push_back(static_cast<uint32_t>(i));

#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; ++i) {
        // candidate columns j: rows sharing at least one column with row i
        std::vector<uint32_t> cand;
        for (auto& p : s[i]) cand.insert(cand.end(), col_rows[p.first].begin(), col_rows[p.first].end());
        std::sort(cand.begin(), cand.end());
        cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
        SparseRow drow;  // ascending column
        for (uint32_t j : cand) {
            bool any;
            const float c = sparse_dot(s[i], s[j], any);
            if (!any) continue;
            if (!eigen_pruned_keep(c, pruneVal)) continue;
            if (hdi) {
                if (static_cast<int64_t>(j) == i) continue;              // :1250-1252
                drow.emplace_back(j, c == 1.f ? c : -1.f * std::log(c));  // :1290-1292
            } else {
                drow.emplace_back(j, static_cast<int64_t>(j) == i ? c : -1.f * std::log(c));  // :1143-1153
            }
        }
        if (!hdi) {
            // removeDiagonalElements(keepSingleEntry=true) (:704-718)
            if (drow.size() > 1) {
                bool had = false;
                for (auto& p : drow) if (static_cast<int64_t>(p.first) == i) { p.second = 0.f; had = true; }
                (void)had;
                drow.erase(std::remove_if(drow.begin(), drow.end(), [](auto& p) { return p.second == 0.f; }), drow.end());
            }
        }
        SparseRow sel = select_k<MinDeque>(drow, static_cast<size_t>(k));
        double sum = 0;
        for (auto& e : sel) sum += e.second;
        if (hdi || sum > 0)
            for (auto& e : sel) e.second /= static_cast<float>(sum);
        row_nnz[i] = static_cast<int32_t>(sel.size());
        for (int64_t t = 0; t < k; ++t) {
            out_cols[i * k + t] = t < static_cast<int64_t>(sel.size()) ? sel[t].first : 0xFFFFFFFFu;
            out_vals[i * k + t] = t < static_cast<int64_t>(sel.size()) ? sel[t].second : 0.f;
        }
    }
}

// D1, row-wise (Gustavson) form for the rows [row0, row1): C[i][j] accumulated over the columns c of row i in ascending
// order through the transposed matrix — the products and their fp32 addition order are exactly those of the merge dot
// above (ascending shared column), so both forms are bit-identical; this one costs one multiply-add per expanded product,
// as the reference's blocked Eigen SpGEMM does (sph/utils/SparseMatrixAlgorithms.cpp:1214-1282), which makes it the fair
// CPU baseline of bench.py.  out_*: (row1 - row0) rows of stride k.
ORC_API void orc_create_similarities_rows(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                          int64_t k, float pruneVal, int variant, const int64_t* comp_sizes, int64_t row0, int64_t row1,
                                          uint32_t* out_cols, float* out_vals, int32_t* row_nnz) {
    const SparseMat in = from_csr(n, rowptr, col, val);
    const SparseMat s = sqrt_rows(in, pruneVal, true, comp_sizes);
    const bool hdi = variant == 1 || (variant == 0 && n > 5000);
    std::vector<std::vector<std::pair<uint32_t, float>>> st(static_cast<size_t>(n));  // S transposed: column -> (row, value), rows ascending
    for (int64_t i = 0; i < n; ++i)
        for (auto& p : s[i]) st[p.first].emplace_back(static_cast<uint32_t>(i), p.second);
#pragma omp parallel
    {
        std::vector<float> acc(static_cast<size_t>(n), 0.f);
        std::vector<uint8_t> seen(static_cast<size_t>(n), 0);
        std::vector<uint32_t> touched;
#pragma omp for schedule(dynamic, 16)
        for (int64_t i = row0; i < row1; ++i) {
            touched.clear();
            for (auto& p : s[i]) {
                const float sic = p.second;
                for (auto& q : st[p.first]) {
                    if (!seen[q.first]) { seen[q.first] = 1; touched.push_back(q.first); }
                    acc[q.first] += sic * q.second;
                }
            }
            std::sort(touched.begin(), touched.end());
            SparseRow drow;  // ascending column
            for (uint32_t j : touched) {
                const float c = acc[j];
                acc[j] = 0.f;
                seen[j] = 0;
                if (!eigen_pruned_keep(c, pruneVal)) continue;
                if (hdi) {
                    if (static_cast<int64_t>(j) == i) continue;              // :1250-1252
                    drow.emplace_back(j, c == 1.f ? c : -1.f * std::log(c));  // :1290-1292
                } else {
                    drow.emplace_back(j, static_cast<int64_t>(j) == i ? c : -1.f * std::log(c));  // :1143-1153
                }
            }
            if (!hdi && drow.size() > 1) {  // removeDiagonalElements(keepSingleEntry=true) (:704-718)
                for (auto& p : drow) if (static_cast<int64_t>(p.first) == i) p.second = 0.f;
                drow.erase(std::remove_if(drow.begin(), drow.end(), [](auto& p) { return p.second == 0.f; }), drow.end());
            }
            SparseRow sel = select_k<MinDeque>(drow, static_cast<size_t>(k));
            double sum = 0;
            for (auto& e : sel) sum += e.second;
            if (hdi || sum > 0)
                for (auto& e : sel) e.second /= static_cast<float>(sum);
            const int64_t o = i - row0;
            row_nnz[o] = static_cast<int32_t>(sel.size());
            for (int64_t t = 0; t < k; ++t) {
                out_cols[o * k + t] = t < static_cast<int64_t>(sel.size()) ? sel[t].first : 0xFFFFFFFFu;
                out_vals[o * k + t] = t < static_cast<int64_t>(sel.size()) ? sel[t].second : 0.f;
            }
        }
    }
}

// D1.  variant: 0 = auto (n > 5000 -> HDI else Eigen), 1 = force HDI, 2 = force Eigen.
// Output: fixed stride k per row (cols padded 0xFFFFFFFF), row_nnz[n].
ORC_API void orc_create_similarities(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                     int64_t k, float pruneVal, int variant, const int64_t* comp_sizes,
                                     uint32_t* out_cols, float* out_vals, int32_t* row_nnz) {
    orc_create_similarities_rows(n, rowptr, col, val, k, pruneVal, variant, comp_sizes, 0, n, out_cols, out_vals, row_nnz);
}

// D2: createSimilaritiesSPH — full (unpruned-to-k) symmetric -log(sqrt(a_i)·sqrt(a_j)) matrix, CSR out.
ORC_API int64_t orc_create_similarities_sph(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, float pruneVal,
                                            int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap) {
    const SparseMat in = from_csr(n, rowptr, col, val);
    const SparseMat s = sqrt_rows(in, 0.f, false, nullptr);
    std::vector<std::map<uint32_t, float>> out(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        for (int64_t j = i + 1; j < n; ++j) {
            bool any;
            const float dp = sparse_dot(s[i], s[j], any);
            if (dp < pruneVal) continue;
            const float ov = -1.f * std::log(dp);
            out[i][static_cast<uint32_t>(j)] = ov;
            out[j][static_cast<uint32_t>(i)] = ov;
        }
    int64_t e = 0;
    out_rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        for (auto& kv : out[i]) {
            if (e >= cap) return -1;
            out_col[e] = kv.first; out_val[e] = kv.second; ++e;
        }
        out_rowptr[i + 1] = e;
    }
    return e;
}

// D3: findTopK (top=1) / findBottomK (top=0) of one column-sorted row; returns count.
ORC_API int orc_find_k(const uint32_t* col, const float* val, int nnz, int k, int top, uint32_t* out_col, float* out_val) {
    SparseRow row;
    for (int e = 0; e < nnz; ++e) row.emplace_back(col[e], val[e]);
    SparseRow sel = top ? select_k<MaxDeque>(row, k) : select_k<MinDeque>(row, k);
    for (size_t t = 0; t < sel.size(); ++t) { out_col[t] = sel[t].first; out_val[t] = sel[t].second; }
    return static_cast<int>(sel.size());
}

// D4 batched: for each pair (a[p], b[p]): mode 0 = Bhattacharyya coefficient
// sum_c sqrt(A[a][c]*A[b][c]) over shared columns with A[b][c] > 0; mode 1 = single
// overlap max(A[a][b], A[b][a]).  (componentDistance = 1 - sim, Similarities.cpp:123-156.)
ORC_API void orc_pair_sims(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                           const int64_t* pa, const int64_t* pb, int64_t npairs, int mode, float* out) {
    (void)n;
    auto coeff = [&](int64_t r, uint32_t c) -> float {
        const uint32_t* b = col + rowptr[r];
        const uint32_t* e = col + rowptr[r + 1];
        const uint32_t* it = std::lower_bound(b, e, c);
        return (it != e && *it == c) ? val[it - col] : 0.f;
    };
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < npairs; ++p) {
        const int64_t a = pa[p], b = pb[p];
        if (mode == 1) {
            out[p] = std::max(coeff(a, static_cast<uint32_t>(b)), coeff(b, static_cast<uint32_t>(a)));
        } else {
            float bc = 0.f;
            for (int64_t e = rowptr[a]; e < rowptr[a + 1]; ++e) {
                const float v2 = coeff(b, col[e]);
                if (v2 > 0.0f) bc += std::sqrt(val[e] * v2);
            }
            out[p] = bc;
        }
    }
}

// D5 batched: neighbour overlap card(U1 ∩ U2) / min(card U1, card U2), where U_c is the
// sorted-unique union of the data-level kNN lists of the pixels of component c.
// comp_ptr[ncomp+1] / comp_pix: component -> pixel lists; knn ids [N*k].
ORC_API void orc_neighbor_overlap(const int64_t* knn_ids, int k, const int64_t* comp_ptr, const int64_t* comp_pix,
                                  const int64_t* pa, const int64_t* pb, int64_t npairs, float* out) {
    auto uni = [&](int64_t c) {
        std::vector<int64_t> u;
        for (int64_t e = comp_ptr[c]; e < comp_ptr[c + 1]; ++e) {
            const int64_t* nb = knn_ids + comp_pix[e] * k;
            u.insert(u.end(), nb, nb + k);
        }
        std::sort(u.begin(), u.end());
        u.erase(std::unique(u.begin(), u.end()), u.end());
        return u;
    };
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t p = 0; p < npairs; ++p) {
        const auto u1 = uni(pa[p]), u2 = uni(pb[p]);
        size_t i = 0, j = 0, cnt = 0;
        while (i < u1.size() && j < u2.size()) {
            if (u1[i] == u2[j]) { ++cnt; ++i; ++j; }
            else if (u1[i] < u2[j]) ++i;
            else ++j;
        }
        const size_t mn = std::min(u1.size(), u2.size());
        out[p] = mn > 0 ? static_cast<float>(cnt) / mn : 0.f;
    }
}
