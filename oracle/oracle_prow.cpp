// TEST INFRASTRUCTURE ONLY — see oracle_common.hpp.
// Stage B oracle: perplexity-calibrated probability rows (the t-SNE P matrix).
//   B1 sph/utils/HDILibHelper.hpp:23-109   computeGaussianDistributionWithFixedPerplexity
//   B2 sph/utils/GraphNormalization.cpp:191-338  graph -> rows
//   B3 sph/utils/GraphNormalization.cpp:38-189   in-place on sparse rows
//   B4 sph/utils/HDILibHelper.hpp:112-140        flat fixed-k variant
//   B5 sph/utils/GraphNormalization.cpp:340-411  linear rows
//   B6 sph/utils/HDILibHelper.hpp:260-302        symmetrizeTSNE / symmetrizeUMAP
//   B8 sph/LevelSimilarities.cpp:83-115          k / perplexity schedule
//   §8(f)-3 sph/utils/GraphNormalization.cpp:413-593  exponential (UMAP) rows — parity unpinned, see below
#include "oracle_common.hpp"

#include <algorithm>
#include <cmath>
#include <limits>
#include <map>
#include <numeric>

namespace orc {

// B1.  dist/prob are fp32 arrays of length size; returns sigma (0 on failure).
double gaussian_fixed_perplexity(const float* dist, float* prob, int size, double perplexity, int max_iterations, double tol, int ignore) {
    bool found = false;
    double beta = 1.;
    double sigma = std::sqrt(1 / (2 * beta));
    double min_beta = -std::numeric_limits<double>::max();
    double max_beta = std::numeric_limits<double>::max();
    constexpr double double_max = std::numeric_limits<double>::max();

    int iter = 0;
    double sum_distribution = std::numeric_limits<double>::min();
    while (!found && iter < max_iterations) {
        sum_distribution = std::numeric_limits<double>::min();
        for (int idx = 0; idx < size; ++idx) {
            if (idx == ignore) { prob[idx] = 0; continue; }
            const double v = std::exp(-beta * static_cast<double>(dist[idx]));
            sigma = std::sqrt(1 / (2 * beta));
            prob[idx] = static_cast<float>(v);  // fp32 round trip before the entropy sum (:55,68)
            sum_distribution += v;
        }
        double H = .0;
        for (int idx = 0; idx < size; ++idx) {
            if (idx == ignore) continue;
            H += beta * (dist[idx] * prob[idx]);  // float * float product, promoted to double afterwards
        }
        H = (H / sum_distribution) + std::log(sum_distribution);

        const double Hdiff = H - std::log(perplexity);
        if (Hdiff < tol && -Hdiff < tol) {
            found = true;
        } else {
            if (Hdiff > 0) {
                min_beta = beta;
                if (max_beta == double_max || max_beta == -double_max) beta *= 2.0;
                else beta = (beta + max_beta) / 2.0;
            } else {
                max_beta = beta;
                if (min_beta == -double_max || min_beta == double_max) beta /= 2.0;
                else beta = (beta + min_beta) / 2.0;
            }
        }
        iter++;
    }
    if (!found) {
        const float v = static_cast<float>(1. / (size + ((ignore < 0 || ignore >= size) ? 0 : -1)));
        for (int idx = 0; idx < size; ++idx) prob[idx] = v;
        return 0;
    }
    for (int idx = 0; idx < size; ++idx) prob[idx] = static_cast<float>(prob[idx] / sum_distribution);
    return sigma;
}

// utils::normalizeUnit (sph/utils/Math.hpp:126-137): double sum, value / sum cast back.
static void normalize_unit(float* v, int n) {
    double sum = 0.0;
    for (int i = 0; i < n; ++i) sum += v[i];
    for (int i = 0; i < n; ++i) v[i] = static_cast<float>(v[i] / sum);
}

// min-sigma fallback shared by B2 (:249-282) and B3 (:96-130).
static void min_sigma_fallback(const float* input, float* condProb, int nn, bool zero_first) {
    std::copy(input, input + nn, condProb);
    auto allZero = [&]() { return std::all_of(condProb, condProb + nn, [](float v) { return v == 0; }); };
    auto assignSame = [&]() { std::fill(condProb, condProb + nn, 1.f / (nn - 1)); };
    if (allZero()) { assignSame(); return; }
    normalize_unit(condProb, nn);
    for (int i = 0; i < nn; ++i) condProb[i] = 1.0f - condProb[i];
    if (allZero()) { assignSame(); return; }
    if (zero_first) condProb[0] = 0.f;
    normalize_unit(condProb, nn);
}

}  // namespace orc

using namespace orc;

ORC_API double orc_gaussian_row(const float* dist, float* prob, int size, double perplexity, int max_iterations, double tol, int ignore) {
    return gaussian_fixed_perplexity(dist, prob, size, perplexity, max_iterations, tol, ignore);
}

// B2: graph (fixed k, self first) -> rows with k-1 entries each, column sorted.
// out_cols/out_vals are [n*(k-1)]; sigma_out (optional) [n].
ORC_API void orc_prow_gaussian_graph(const int64_t* ids, const float* dist, int64_t n, int k, float fixed_perplexity, uint32_t* out_cols, float* out_vals, double* sigma_out) {
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p) {
        const int nn = k, nn_eff = k - 1;
        std::vector<float> condProb(nn, -1.f);
        const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : nn_eff / 3.f;
        const float* dp = dist + p * k;
        const double sigma = gaussian_fixed_perplexity(dp, condProb.data(), nn, perplexity, 200, 1e-6, 0);
        if (sigma < 0.001) min_sigma_fallback(dp, condProb.data(), nn, true);
        if (sigma_out) sigma_out[p] = sigma;
        // map insertion semantics: sorted by column, later duplicates overwrite
        std::map<uint32_t, float> row;
        for (int t = 1; t < nn; ++t) row[static_cast<uint32_t>(ids[p * k + t])] = condProb[t];
        int e = 0;
        for (auto& kv : row) { out_cols[p * nn_eff + e] = kv.first; out_vals[p * nn_eff + e] = kv.second; ++e; }
        for (; e < nn_eff; ++e) { out_cols[p * nn_eff + e] = 0xFFFFFFFFu; out_vals[p * nn_eff + e] = 0.f; }
    }
}

// B3: in place on CSR rows (values replaced; pattern unchanged), including the
// "p < 1e-10 stops the write-back cursor" quirk (:132-153).
ORC_API void orc_prow_gaussian_inplace(int64_t n, const int64_t* rowptr, float* val, float fixed_perplexity, int ignore) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t p = 0; p < n; ++p) {
        const int nn = static_cast<int>(rowptr[p + 1] - rowptr[p]);
        if (nn <= 1) continue;
        float* v = val + rowptr[p];
        std::vector<float> input(v, v + nn), condProb(nn, -1.f);
        const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : nn / 3.f;
        const double sigma = gaussian_fixed_perplexity(input.data(), condProb.data(), nn, perplexity, 200, 1e-6, ignore);
        if (sigma < 0.001) min_sigma_fallback(input.data(), condProb.data(), nn, ignore >= 0);
        int kk = 0;
        for (int e = 0; e < nn; ++e) {
            if (condProb[kk] < 1.0e-10f) continue;
            v[e] = condProb[kk];
            kk++;
        }
    }
}

// B4: flat variant, tol 1e-5, ignore 0; writes the temp vector [n*k]
// (entry 0 of each row is the ignored self slot).  The caller scatters
// entries 1..k-1 skipping ids == -1 (HDILibHelper.hpp:131-139).
ORC_API void orc_prow_gaussian_flat(const float* dist, int64_t n, int k, double perplexity, float* temp) {
#pragma omp parallel for schedule(static)
    for (int64_t j = 0; j < n; ++j) gaussian_fixed_perplexity(dist + j * k, temp + j * k, k, perplexity, 200, 1e-5, 0);
}

// B5: linear rows: p = normalise(1/(1+d)) over neighbours 1..k-1.
// B2 / B5 on a ragged graph (any GraphBaseInterface, GraphNormalization.cpp:191-411): row p = [rowptr[p], rowptr[p+1]), the
// point itself first; k_p - 1 column-sorted output entries per row starting at rowptr[p] - p.
ORC_API void orc_prow_graph_ragged(int64_t n, const int64_t* rowptr, const int64_t* ids, const float* dist, float fixed_perplexity, int linear,
                                   int64_t* out_rowptr, uint32_t* out_cols, float* out_vals) {
    for (int64_t p = 0; p <= n; ++p) out_rowptr[p] = rowptr[p] - p;
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p) {
        const int64_t b = rowptr[p];
        const int nn = static_cast<int>(rowptr[p + 1] - b), nn_eff = nn - 1;
        if (nn_eff <= 0) continue;
        std::vector<float> condProb(nn, -1.f);
        if (linear) {
            for (int t = 1; t < nn; ++t) condProb[t] = 1.0f / (1.0f + dist[b + t]);
            normalize_unit(condProb.data() + 1, nn_eff);
        } else {
            const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : nn_eff / 3.f;
            const double sigma = gaussian_fixed_perplexity(dist + b, condProb.data(), nn, perplexity, 200, 1e-6, 0);
            if (sigma < 0.001) min_sigma_fallback(dist + b, condProb.data(), nn, true);
        }
        std::map<uint32_t, float> row;
        for (int t = 1; t < nn; ++t) row[static_cast<uint32_t>(ids[b + t])] = condProb[t];
        int64_t e = out_rowptr[p];
        for (auto& kv : row) { out_cols[e] = kv.first; out_vals[e] = kv.second; ++e; }
    }
}

ORC_API void orc_prow_linear_graph(const int64_t* ids, const float* dist, int64_t n, int k, uint32_t* out_cols, float* out_vals) {
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p) {
        const int ne = k - 1;
        std::vector<float> condProb(dist + p * k + 1, dist + p * k + k);
        for (auto& v : condProb) v = 1.0f / (1.0f + v);
        normalize_unit(condProb.data(), ne);
        std::map<uint32_t, float> row;
        for (int t = 1; t < k; ++t) row[static_cast<uint32_t>(ids[p * k + t])] = condProb[t - 1];
        int e = 0;
        for (auto& kv : row) { out_cols[p * ne + e] = kv.first; out_vals[p * ne + e] = kv.second; ++e; }
        for (; e < ne; ++e) { out_cols[p * ne + e] = 0xFFFFFFFFu; out_vals[p * ne + e] = 0.f; }
    }
}

// B6: symmetrizeTSNE (umap=0) / symmetrizeUMAP (umap=1), literal restatement on
// std::map rows (operator[] inserts a zero entry, as MapMemEff does).
// Output CSR capacity must be >= 2*nnz_in.  Returns nnz_out or -1.
ORC_API int64_t orc_symmetrize(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val, int umap, int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap) {
    std::vector<std::map<uint32_t, float>> dist(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i)
        for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e) dist[i][col[e]] = val[e];
    for (uint32_t j = 0; j < static_cast<uint32_t>(n); ++j) {
        // iterate over a snapshot of the keys: inserting into row i != j never
        // touches row j, and for i == j the entry already exists.
        for (auto& e : dist[j]) {
            const uint32_t i = e.first;
            float& val_ji = e.second;
            float& val_ij = dist[i][j];
            if (j > i && val_ij > 0.f) continue;
            const float nv = umap ? (val_ij + val_ji) - (val_ij * val_ji) : (val_ji + val_ij) * 0.5f;
            val_ji = nv;
            val_ij = nv;
        }
    }
    int64_t e = 0;
    out_rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        for (auto& kv : dist[i]) {
            if (e >= cap) return -1;
            out_col[e] = kv.first;
            out_val[e] = kv.second;
            ++e;
        }
        out_rowptr[i + 1] = e;
    }
    return e;
}

// B8: LevelSimilarities::updateNumberOfNeighbors (sph/LevelSimilarities.cpp:83-115).
// level_sizes[L]: number of components per level; k0 = data-level k (incl. self).
// perplexity_0 = clamp((k0-1)/3, 10, 100), ks[0] = k0 (unchanged);
// level >= 1: perplexity_l = min((k0-1)/3 [UNCLAMPED], clamp(n_l/100, 10, 100));
//             k_l = min(int64(perplexity_l)*3+1, n_l).
ORC_API void orc_level_schedule(const int64_t* level_sizes, int num_levels, int64_t k0, float* perplexities, int64_t* ks) {
    auto clampf = [](float v, float lo, float hi) { return std::min(std::max(v, lo), hi); };
    if (num_levels <= 0) return;
    const float dataPerplexity = static_cast<float>(k0 - 1) / 3.f;
    perplexities[0] = clampf(dataPerplexity, 10.f, 100.f);
    ks[0] = k0;
    for (int l = 1; l < num_levels; ++l) {
        float pl = clampf(static_cast<float>(level_sizes[l]) / 100.f, 10.f, 100.f);
        pl = std::min(dataPerplexity, pl);
        perplexities[l] = pl;
        ks[l] = std::min<int64_t>(static_cast<int64_t>(pl) * 3 + 1, level_sizes[l]);
    }
}

// ---- §8(f)-3: UMAP-flavoured rows -------------------------------------------------------------------------------------
// utils::computeExponentialDistributions (sph/utils/GraphNormalization.cpp:413-593) hands each row to
// umappp::neighbor_similarities (libscran-umappp 3.2.0, vcpkg.json:146-147; NOT vendored under /root/reference and absent
// from this image) with default options: local_connectivity 1, bandwidth 1, Float_ = float (NeighborList<int64_t, float>,
// :418,514).  PARITY UNPINNED for the arithmetic below: the reference holds no golden vector for these rows and the
// library source is not available here, so this is a restatement of the published algorithm the library implements
// (UMAP's smooth_knn_dist, McInnes et al. 2018 Alg. 3 + the reference Python implementation's constants):
//   rho    = distance of the first neighbour with a non-zero distance (local_connectivity = 1: index 1, no interpolation),
//            0 when all distances are zero;
//   sigma  = root of  sum_j [ d_j > rho ? exp(-(d_j - rho) / sigma) : 1 ] = log2(n_neighbours + 1) * bandwidth
//            (the +1 is the point itself, which the neighbour list does not hold), found from sigma = 1 by doubling while no
//            upper bound is known and bisection afterwards, at most 64 evaluations, stop at |sum - target| < 1e-5;
//   sigma  = max(sigma, 1e-3 * mean of the row's distances)        (min_k_dist_scale);
//   w_j    = d_j > rho ? exp(-(d_j - rho) / sigma) : 1.
// All arithmetic in fp32 like the library's Float_.  Any conforming implementation (umappp refines the same root with
// Newton steps inside the same bracket) agrees with this to the search tolerance, not bit for bit; tests compare the CUDA
// kernel with THIS function at rtol 1e-4 and check the defining equation sum_j w_j = log2(n + 1) directly.
static void umap_row_weights(const float* dist, float* out, int nn) {
    if (nn <= 0) return;
    float rho = 0.f;
    for (int j = 0; j < nn; ++j)
        if (dist[j] != 0.f) { rho = dist[j]; break; }
    int num_le = 0;
    float total = 0.f;
    for (int j = 0; j < nn; ++j) { num_le += !(dist[j] > rho); total += dist[j]; }
    if (num_le == nn) {
        for (int j = 0; j < nn; ++j) out[j] = 1.f;
        return;
    }
    const float target = std::log2(static_cast<float>(nn + 1));
    constexpr float fmax = std::numeric_limits<float>::max();
    float sigma = 1.f, lo = 0.f, hi = fmax;
    for (int iter = 0; iter < 64; ++iter) {
        float val = static_cast<float>(num_le);
        for (int j = 0; j < nn; ++j)
            if (dist[j] > rho) val += std::exp(-(dist[j] - rho) / sigma);
        if (std::fabs(val - target) < 1e-5f) break;
        if (val > target) { hi = sigma; sigma = (lo + hi) / 2.f; }
        else { lo = sigma; sigma = (hi == fmax) ? sigma * 2.f : (lo + hi) / 2.f; }
    }
    sigma = std::max(sigma, 1e-3f * (total / static_cast<float>(nn)));
    for (int j = 0; j < nn; ++j) out[j] = dist[j] > rho ? std::exp(-(dist[j] - rho) / sigma) : 1.f;
}

// computeExponentialDistributions(const GraphBaseInterface&, vector<T>&) (GraphNormalization.cpp:508-593): neighbours 1..k-1
// of every point (the point itself, neighbour 0, is skipped :523), weights stored in NEIGHBOUR order for T = MapMemEff
// (storage.emplace_back :566-569; sort_cols = 0) and in column order for T = SparseVecSPH (insert :559-562; sort_cols = 1).
ORC_API void orc_prow_exponential_graph(const int64_t* ids, const float* dist, int64_t n, int k, int sort_cols, uint32_t* out_cols,
                                        float* out_vals) {
    const int ne = k - 1;
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p) {
        std::vector<float> w(static_cast<size_t>(std::max(ne, 0)));
        umap_row_weights(dist + p * k + 1, w.data(), ne);
        std::vector<std::pair<uint32_t, float>> row(w.size());
        for (int t = 0; t < ne; ++t) row[t] = {static_cast<uint32_t>(ids[p * k + 1 + t]), w[t]};
        if (sort_cols) std::stable_sort(row.begin(), row.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        for (int t = 0; t < ne; ++t) { out_cols[p * ne + t] = row[t].first; out_vals[p * ne + t] = row[t].second; }
    }
}

// computeExponentialDistributions(vector<T>&, int ignore) in place on sparse rows (GraphNormalization.cpp:413-506): the
// entries of a row in STORED (column) order are the "neighbours" (:430-447; they are not sorted by value, so rho is the first
// non-zero stored value), an entry whose COLUMN equals `ignore` is left out and keeps its value (SparseVecSPH write-back by
// id :472-475).  Called with ignore = -1 (sph/LevelSimilarities.cpp:502).
ORC_API void orc_prow_exponential_inplace(int64_t n, const int64_t* rowptr, const uint32_t* col, float* val, int ignore) {
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < n; ++p) {
        std::vector<float> d, w;
        std::vector<int64_t> pos;
        for (int64_t e = rowptr[p]; e < rowptr[p + 1]; ++e) {
            if (ignore >= 0 && col && col[e] == static_cast<uint32_t>(ignore)) continue;
            d.push_back(val[e]);
            pos.push_back(e);
        }
        w.resize(d.size());
        umap_row_weights(d.data(), w.data(), static_cast<int>(d.size()));
        for (size_t t = 0; t < d.size(); ++t) val[pos[t]] = w[t];
    }
}
