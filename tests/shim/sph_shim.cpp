// The INTEGRATION.md shims as ONE compiled C++ translation unit: it includes include/sphb.h, links libsphb200.so (nothing
// else) and drives the C ABI through the reference's own container shapes —
//   utils::Graph  (sph/utils/Graph.hpp:155-273: knnIndices int64, knnDistances float, fixed k),
//   SparseMatHDI = std::vector<hdi::data::MapMemEff<uint32_t, float>> (sph/utils/CommonDefinitions.hpp:87-89), stood in by a
//   class with the same storage (a column-sorted std::vector<std::pair<uint32_t, float>> reachable through memory()) and the
//   same initialize(begin, end) the reference's tests use (tests/UtilsTests.cpp:1143-1161).
// Built and run by tests/test_shim.py.  Exit code 0 + "shim ok" on success; 77 when no CUDA device is present.
#include "../../include/sphb.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <random>
#include <utility>
#include <vector>

namespace shim {

// stand-in of hdi::data::MapMemEff<uint32_t, float>: sorted pair storage
class MapMemEff {
public:
    using storage_type = std::vector<std::pair<uint32_t, float>>;
    template <class It>
    void initialize(It begin, It end) { _mem.assign(begin, end); }
    storage_type& memory() { return _mem; }
    const storage_type& memory() const { return _mem; }
    size_t size() const { return _mem.size(); }
private:
    storage_type _mem;
};
using SparseMatHDI = std::vector<MapMemEff>;

struct Graph {  // the fields of utils::Graph the hot path touches
    std::vector<int64_t> knnIndices;
    std::vector<float> knnDistances;
    int64_t numPoints = 0, k = 0;
};

#define SPHB_CHECK(call)                                                                      \
    do {                                                                                      \
        if (int rc_ = (call); rc_ != SPHB_OK) {                                               \
            std::fprintf(stderr, "%s: %s (%d)\n", #call, sphb_last_error(), rc_);             \
            return rc_;                                                                       \
        }                                                                                     \
    } while (0)

// sph/utils/Knn.cpp:110-136 computeIndexFlat (+ the fused post-passes of NearestNeighbors::compute)
int computeIndexFlat(int64_t nn, int faissMetric, const float* data, int64_t numPoints, int numDims, bool l2squared, Graph& g) {
    g.numPoints = numPoints;
    g.k = nn;
    g.knnIndices.assign(static_cast<size_t>(numPoints * nn), -1);      // sph/NearestNeighbors.cpp:113-118
    g.knnDistances.assign(static_cast<size_t>(numPoints * nn), -1.0f);
    SPHB_CHECK(sphb_knn_exact(data, numPoints, numDims, static_cast<int>(nn), faissMetric, l2squared ? 1 : 0, g.knnIndices.data(), g.knnDistances.data()));
    return SPHB_OK;
}

// sph/utils/GraphNormalization.cpp:191-338 computeGaussianDistributions(graph, rows, perplexity)
int computeGaussianDistributions(const Graph& g, SparseMatHDI& rows, float perplexity) {
    const int64_t n = g.numPoints;
    const int k = static_cast<int>(g.k);
    std::vector<uint32_t> cols(static_cast<size_t>(n) * (k - 1));
    std::vector<float> vals(static_cast<size_t>(n) * (k - 1));
    SPHB_CHECK(sphb_prow_gaussian_graph(g.knnIndices.data(), g.knnDistances.data(), n, k, perplexity, cols.data(), vals.data(), nullptr));
    rows.resize(static_cast<size_t>(n));
    std::vector<std::pair<uint32_t, float>> tmp(static_cast<size_t>(k - 1));
    for (int64_t i = 0; i < n; ++i) {
        for (int t = 0; t < k - 1; ++t) tmp[t] = {cols[i * (k - 1) + t], vals[i * (k - 1) + t]};
        rows[i].initialize(tmp.begin(), tmp.end());  // rows arrive column-sorted
    }
    return SPHB_OK;
}

struct Csr {
    std::vector<int64_t> rowptr;
    std::vector<uint32_t> col;
    std::vector<float> val;
};
Csr to_csr(const SparseMatHDI& m) {
    Csr c;
    c.rowptr.assign(1, 0);
    for (const auto& r : m) {
        for (const auto& p : r.memory()) { c.col.push_back(p.first); c.val.push_back(p.second); }
        c.rowptr.push_back(static_cast<int64_t>(c.col.size()));
    }
    return c;
}
void from_csr(const Csr& c, int64_t n, SparseMatHDI& m) {
    m.assign(static_cast<size_t>(n), MapMemEff());
    std::vector<std::pair<uint32_t, float>> tmp;
    for (int64_t i = 0; i < n; ++i) {
        tmp.clear();
        for (int64_t e = c.rowptr[i]; e < c.rowptr[i + 1]; ++e) tmp.emplace_back(c.col[e], c.val[e]);
        m[i].initialize(tmp.begin(), tmp.end());
    }
}

// sph/utils/HDILibHelper.hpp:260-280 symmetrizeTSNE(std::vector<map_type>&)
int symmetrizeTSNE(SparseMatHDI& distribution) {
    const Csr in = to_csr(distribution);
    const int64_t n = static_cast<int64_t>(distribution.size());
    Csr out;
    const int64_t cap = std::max<int64_t>(1, 2 * static_cast<int64_t>(in.col.size()));
    out.rowptr.resize(n + 1);
    out.col.resize(cap);
    out.val.resize(cap);
    int64_t m = 0;
    SPHB_CHECK(sphb_symmetrize(n, in.rowptr.data(), in.col.data(), in.val.data(), 0, out.rowptr.data(), out.col.data(), out.val.data(), cap, &m));
    from_csr(out, n, distribution);
    return SPHB_OK;
}

// LevelSimilarities::computeProbDistOnLevel (useRandomWalks) + symmetrizeOutput for one level, sph/LevelSimilarities.cpp:460-508, 589-623:
// what getProbDist(level) hands to ComputeEmbedding::computeTSNE
int levelProbDist(const SparseMatHDI& walks, int64_t k, float perplexity, SparseMatHDI& probDist) {
    const Csr in = to_csr(walks);
    const int64_t n = static_cast<int64_t>(walks.size());
    Csr out;
    const int64_t cap = std::max<int64_t>(1, 2 * n * k);
    out.rowptr.resize(n + 1);
    out.col.resize(cap);
    out.val.resize(cap);
    int64_t m = 0;
    SPHB_CHECK(sphb_level_prob_dist(n, in.rowptr.data(), in.col.data(), in.val.data(), k, 1e-4f, 0, nullptr, 1, perplexity, 1, 1, out.rowptr.data(),
                                    out.col.data(), out.val.data(), cap, &m));
    from_csr(out, n, probDist);
    return SPHB_OK;
}

}  // namespace shim

int main() {
    using namespace shim;
    if (sphb_init(0) != SPHB_OK) {
        std::printf("no device: %s\n", sphb_last_error());
        return 77;
    }
    // 2000 points on a noisy circle in 16-D
    const int64_t n = 2000;
    const int d = 16, k = 31;
    std::mt19937 rng(5);
    std::normal_distribution<float> nd(0.f, 0.02f);
    std::vector<float> x(static_cast<size_t>(n) * d);
    for (int64_t i = 0; i < n; ++i)
        for (int c = 0; c < d; ++c) x[i * d + c] = std::sin(6.2831853f * i / n + 0.4f * c) + nd(rng);
    Graph g;
    if (computeIndexFlat(k, SPHB_METRIC_L2, x.data(), n, d, false, g)) return 1;
    for (int64_t i = 0; i < n; ++i) {
        if (g.knnIndices[i * k] != i || g.knnDistances[i * k] != 0.f) { std::printf("row %lld: self not first\n", static_cast<long long>(i)); return 2; }
        for (int t = 1; t < k; ++t)
            if (g.knnDistances[i * k + t] < g.knnDistances[i * k + t - 1] || g.knnIndices[i * k + t] < 0) { std::printf("row %lld not monotone\n", static_cast<long long>(i)); return 3; }
    }
    SparseMatHDI P;
    if (computeGaussianDistributions(g, P, 0.f)) return 1;
    for (int64_t i = 0; i < n; ++i) {
        double s = 0;
        const auto& mem = P[i].memory();
        if (static_cast<int>(mem.size()) != k - 1) return 4;
        for (size_t t = 0; t < mem.size(); ++t) {
            s += mem[t].second;
            if (t && mem[t].first <= mem[t - 1].first) return 5;  // MapMemEff order: ascending column
        }
        if (std::fabs(s - 1.0) > 1e-4) { std::printf("row %lld sums to %f\n", static_cast<long long>(i), s); return 6; }
    }
    SparseMatHDI L1;
    if (levelProbDist(P, 16, 5.f, L1)) return 1;     // the level-0 rows serve as a stand-in "walk" matrix
    if (symmetrizeTSNE(P)) return 1;
    // symmetric now: P[i][j] == P[j][i]
    for (int64_t i = 0; i < n; i += 97)
        for (const auto& e : P[i].memory()) {
            const auto& other = P[e.first].memory();
            auto it = std::lower_bound(other.begin(), other.end(), std::make_pair(static_cast<uint32_t>(i), -1.f));
            if (it == other.end() || it->first != static_cast<uint32_t>(i) || it->second != e.second) return 7;
        }
    size_t nnz1 = 0;
    for (const auto& r : L1) nnz1 += r.size();
    std::printf("shim ok: n=%lld k=%d level-0 P nnz=%zu (symmetrised), level rows nnz=%zu, launches=%lld\n", static_cast<long long>(n), k,
                static_cast<size_t>(to_csr(P).col.size()), nnz1, static_cast<long long>(sphb_launch_count()));
    sphb_shutdown();
    return 0;
}
