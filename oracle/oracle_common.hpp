// TEST INFRASTRUCTURE ONLY — CPU oracle for the SPH hot path.
//
// This directory is a CPU restatement of the reference algorithms (alxvth/SPH
// @ 8d54e737) for the hot path named in BASELINE.json.  It is the *checker*:
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.  The product (sph_b200/, include/) never
// links, imports or calls anything in here.
//
// Parity status:
//   Stage A (kNN):    parity UNPINNED against the reference — the reference
//                     holds no golden index/distance vectors for the faiss call
//                     (SURVEY §8c); the contract is fp64 direct distances,
//                     ordered by (fp32 distance, index).
//   Stage B search:   parity UNPINNED (no reference test of the Gaussian rows);
//                     symmetrizeTSNE pinned by tests/UtilsTests.cpp:1143-1221.
//   Stage C / D:      pinned by the reference goldens listed in
//                     tests/golden/README.md.
//   Geodesic (8f-4):  parity UNPINNED against Boost.Graph (absent); pinned to the
//                     reference's own A* test bar and to scipy's Dijkstra, see
//                     oracle_geodesic.cpp.
//   UMAP rows (8f-3): parity UNPINNED (umappp absent), see oracle_prow.cpp.
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>
#include <utility>

#if defined(_OPENMP)
#include <omp.h>
#endif

#define ORC_API extern "C" __attribute__((visibility("default")))

namespace orc {

// One sparse row: column-sorted (col, value) pairs.  Stands in for both
// Eigen::SparseVector<float,0,int32> (SparseVecSPH) and
// hdi::data::MapMemEff<uint32,float> (SparseVecHDI); both iterate in ascending
// column order (sph/utils/CommonDefinitions.hpp:87-93).
using SparseRow = std::vector<std::pair<uint32_t, float>>;
using SparseMat = std::vector<SparseRow>;

// CSR <-> rows helpers (rowptr int64[n+1], col uint32[nnz], val float[nnz]).
inline SparseMat from_csr(int64_t n, const int64_t* rowptr, const uint32_t* col, const float* val) {
    SparseMat m(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        m[i].reserve(static_cast<size_t>(rowptr[i + 1] - rowptr[i]));
        for (int64_t e = rowptr[i]; e < rowptr[i + 1]; ++e) m[i].emplace_back(col[e], val[e]);
    }
    return m;
}

inline int64_t nnz_of(const SparseMat& m) {
    int64_t s = 0;
    for (auto& r : m) s += static_cast<int64_t>(r.size());
    return s;
}

// Writes rows into caller-provided CSR arrays; returns nnz or -1 if capacity too small.
inline int64_t to_csr(const SparseMat& m, int64_t* rowptr, uint32_t* col, float* val, int64_t cap) {
    int64_t e = 0;
    rowptr[0] = 0;
    for (size_t i = 0; i < m.size(); ++i) {
        for (auto& p : m[i]) {
            if (e >= cap) return -1;
            col[e] = p.first;
            val[e] = p.second;
            ++e;
        }
        rowptr[i + 1] = e;
    }
    return e;
}

}  // namespace orc
