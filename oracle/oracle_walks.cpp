// TEST INFRASTRUCTURE ONLY — see oracle_common.hpp.
// Stage C oracle: random-walk transition sums and their merges.
//   RW1 sph/utils/SparseMatrixAlgorithms.cpp:34-290   doRandomWalks
//   RW2 sph/utils/SparseMatrixAlgorithms.cpp:304-352  mergeNodesRandomWalksSingleThread
//   RW4 sph/utils/SparseMatrixAlgorithms.cpp:617-718  normalisers / removeDiagonalElements
// Rows are std::map<uint32,float> while being built (Eigen coeffRef semantics:
// sorted by index, zero-initialised on first touch), then emitted column-sorted.
#include "oracle_common.hpp"

#include <algorithm>
#include <cmath>
#include <map>
#include <random>
#include <set>

namespace {

enum Weighting { W_CONSTANT = 0, W_LINEAR = 1, W_NORMAL = 2, W_ONLYLAST = 3, W_FIRST_VISIT = 4 };

inline float stepLinear(uint64_t step, uint64_t L) { return static_cast<float>(1.0 - (static_cast<double>(step) / L)); }
inline float normWeight(float index) { return std::exp(-0.5f * std::pow(index, 2.f)); }
inline float stepNormal(uint64_t step, uint64_t L) { return normWeight(step * 3.f / L); }

using MapRow = std::map<uint32_t, float>;

// Eigen: sparseVec.sum() accumulates in float, in storage order.
inline float row_sum_f(const MapRow& r) {
    float s = 0.f;
    for (auto& kv : r) s += kv.second;
    return s;
}
inline void normalize_unit_row(MapRow& r) {
    const float nv = row_sum_f(r);
    for (auto& kv : r) kv.second /= nv;
}
// Eigen prune(0, dummy_precision): keeps |v| > 0
inline void prune_zeros(MapRow& r) {
    for (auto it = r.begin(); it != r.end();)
        if (it->second == 0.f) it = r.erase(it);
        else ++it;
}

}  // namespace

// RW1.  in: row-stochastic CSR (column sorted).  threads: 1 = the reproducible
// serial mode (settings.parallel=false); >1 emulates the reference's OpenMP
// layout: T identically-seeded mt19937_64 streams, static contiguous chunks of
// start points per thread (SparseMatrixAlgorithms.cpp:80-101).
// Output CSR capacity `cap`; returns nnz or -1.
ORC_API int64_t orc_random_walks(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                                 uint64_t num_walks, uint64_t walk_len, int weighting, float prune_value, uint64_t prune_steps,
                                 uint64_t seed, int threads, int remove_diagonal, int normalize,
                                 int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap) {
    float pruneValue = prune_value;
    if (prune_steps > 0) {
        const uint64_t pruneStep = std::min(walk_len - 1, prune_steps);
        const uint64_t pruneStepInv = walk_len - pruneStep;
        if (weighting == W_LINEAR) pruneValue = stepLinear(pruneStepInv, walk_len);
        else if (weighting == W_NORMAL) pruneValue = stepNormal(pruneStepInv, walk_len);
        else if (weighting == W_CONSTANT) pruneValue = static_cast<float>(pruneStepInv);
        else if (weighting == W_FIRST_VISIT) pruneValue = static_cast<float>(pruneStep) / walk_len;
        if (pruneValue > 0.5f) pruneValue = 0.5f;
    }

    std::vector<MapRow> walks(static_cast<size_t>(n));
    const int T = std::max(1, threads);
    std::seed_seq seeds{seed};
    std::vector<std::mt19937_64> generators(T);
    for (int t = 0; t < T; ++t) {
        std::seed_seq s2{seed};
        generators[t] = std::mt19937_64(s2);
    }
    std::vector<std::uniform_real_distribution<double>> dists(T, std::uniform_real_distribution<double>(0.0, 1.0));

    // OpenMP static schedule: thread t gets a contiguous chunk; the first (n % T) threads get one extra.
    const int64_t base = n / T, rem = n % T;
    int64_t begin = 0;
    for (int t = 0; t < T; ++t) {
        const int64_t len = base + (t < rem ? 1 : 0);
        for (int64_t start = begin; start < begin + len; ++start) {
            MapRow& hits = walks[start];
            MapRow hitCount;
            for (uint64_t w = 0; w < num_walks; ++w) {
                float stepWeight = 1.f;
                int64_t cur = start;
                std::set<int64_t> wasVisited;
                wasVisited.insert(cur);
                for (uint64_t step = 0; step < walk_len; ++step) {
                    const double u = dists[t](generators[t]);
                    int64_t next = cur;
                    double inc = 0.;
                    for (int64_t e = rowptr[cur]; e < rowptr[cur + 1]; ++e) {
                        inc += val[e];
                        if (u < inc) { next = col[e]; break; }
                    }
                    switch (weighting) {
                        case W_LINEAR: stepWeight = stepLinear(step, walk_len); break;
                        case W_NORMAL: stepWeight = stepNormal(step, walk_len); break;
                        case W_ONLYLAST: stepWeight = (step < walk_len - 1) ? 0.f : 1.f; break;
                        case W_CONSTANT: break;
                        case W_FIRST_VISIT:
                            if (wasVisited.count(next)) stepWeight = 0;
                            else {
                                wasVisited.insert(next);
                                hitCount[static_cast<uint32_t>(next)] += 1.f;
                                stepWeight = static_cast<float>(step + 1);
                            }
                            break;
                    }
                    if (stepWeight > 0.f) hits[static_cast<uint32_t>(next)] += stepWeight;
                    cur = next;
                }
            }
            if (weighting == W_FIRST_VISIT) {
                for (auto& kv : hits) kv.second /= hitCount[kv.first];
                const float m = -1.f / (walk_len - 1.f);
                const float c = walk_len / (walk_len - 1.f);
                for (auto& kv : hits) kv.second = std::max(0.f, m * kv.second + c);
            }
            if (pruneValue > 0.f) {
                for (auto& kv : hits)
                    if (kv.second <= pruneValue) kv.second = 0.f;
                prune_zeros(hits);
            }
        }
        begin += len;
    }

    // removeDiagonalElements(keepSingleEntry=true) (:704-718): coeffRef(i)=0 then prune zeros
    if (remove_diagonal) {
        for (int64_t i = 0; i < n; ++i) {
            if (walks[i].size() <= 1) continue;
            walks[i][static_cast<uint32_t>(i)] = 0.0f;
            prune_zeros(walks[i]);
        }
    }
    if (normalize)
        for (int64_t i = 0; i < n; ++i) normalize_unit_row(walks[i]);

    int64_t e = 0;
    out_rowptr[0] = 0;
    for (int64_t i = 0; i < n; ++i) {
        for (auto& kv : walks[i]) {
            if (e >= cap) return -1;
            out_col[e] = kv.first;
            out_val[e] = kv.second;
            ++e;
        }
        out_rowptr[i + 1] = e;
    }
    return e;
}

// RW2: M[parent(r)][parent(c)] += A[r][c] * w_r, w_r = nnz(row r) or 1;
// rows / sum(w) if weightBySize; optional L1 normalise.  Accumulation order is
// row-major over the input, as in the single-thread reference.
ORC_API int64_t orc_rw_merge(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val,
                             const uint64_t* parents, int64_t n_merged, int norm, int weight_by_size,
                             int64_t* out_rowptr, uint32_t* out_col, float* out_val, int64_t cap) {
    std::vector<MapRow> merged(static_cast<size_t>(n_merged));
    std::vector<uint64_t> rowWeights(static_cast<size_t>(n_merged), 0);
    for (int64_t r = 0; r < n; ++r) {
        const uint64_t into = parents[r];
        const int64_t nnz = rowptr[r + 1] - rowptr[r];
        const long long rowWeight = weight_by_size ? static_cast<long long>(nnz) : 1ll;
        rowWeights[into] += rowWeight;
        for (int64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
            const uint32_t intoCol = static_cast<uint32_t>(parents[col[e]]);
            // it.value() * rowWeight: float * long long -> float
            merged[into][intoCol] += val[e] * rowWeight;
        }
    }
    if (weight_by_size)
        for (int64_t i = 0; i < n_merged; ++i)
            for (auto& kv : merged[i]) kv.second /= rowWeights[i];  // float /= uint64 -> float division
    if (norm)
        for (int64_t i = 0; i < n_merged; ++i) normalize_unit_row(merged[i]);

    int64_t e = 0;
    out_rowptr[0] = 0;
    for (int64_t i = 0; i < n_merged; ++i) {
        for (auto& kv : merged[i]) {
            if (e >= cap) return -1;
            out_col[e] = kv.first;
            out_val[e] = kv.second;
            ++e;
        }
        out_rowptr[i + 1] = e;
    }
    return e;
}
