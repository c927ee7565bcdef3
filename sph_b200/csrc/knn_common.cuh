// Device routines shared by the tensor-core sweep and the SIMT exact kernel:
// the exact pair key, candidate compaction (exact re-rank + sort + truncate) and
// the fused write-out with the reference's post-passes.
#pragma once
#include "common.cuh"
#include <cfloat>

namespace sphb {

constexpr int METRIC_IP = 0;  // faiss::METRIC_INNER_PRODUCT
constexpr int METRIC_L2 = 1;  // faiss::METRIC_L2

// Exact pair key, bit-identical to the oracle (oracle/oracle_knn.cpp pair_key):
// fp64 sequential FMA chain over c = 0..d-1, rounded once to fp32.
// L2: squared distance.  IP: NEGATED dot (ascending order = descending similarity).
__device__ __forceinline__ float exact_key(const float* __restrict__ X, int64_t i, int64_t j, int d, int metric) {
    const float* xi = X + i * d;
    const float* xj = X + j * d;
    double acc = 0.0;
    if ((d & 3) == 0) {
        const float4* a4 = reinterpret_cast<const float4*>(xi);
        const float4* b4 = reinterpret_cast<const float4*>(xj);
        if (metric == METRIC_L2) {
            for (int c = 0; c < (d >> 2); ++c) {
                const float4 a = __ldg(a4 + c), b = __ldg(b4 + c);
                double t;
                t = (double)a.x - (double)b.x; acc = fma(t, t, acc);
                t = (double)a.y - (double)b.y; acc = fma(t, t, acc);
                t = (double)a.z - (double)b.z; acc = fma(t, t, acc);
                t = (double)a.w - (double)b.w; acc = fma(t, t, acc);
            }
        } else {
            for (int c = 0; c < (d >> 2); ++c) {
                const float4 a = __ldg(a4 + c), b = __ldg(b4 + c);
                acc = fma((double)a.x, (double)b.x, acc);
                acc = fma((double)a.y, (double)b.y, acc);
                acc = fma((double)a.z, (double)b.z, acc);
                acc = fma((double)a.w, (double)b.w, acc);
            }
        }
    } else {
        if (metric == METRIC_L2) {
            for (int c = 0; c < d; ++c) {
                const double t = (double)__ldg(xi + c) - (double)__ldg(xj + c);
                acc = fma(t, t, acc);
            }
        } else {
            for (int c = 0; c < d; ++c) acc = fma((double)__ldg(xi + c), (double)__ldg(xj + c), acc);
        }
    }
    const float v = (float)acc;
    return metric == METRIC_L2 ? v : -v;
}

// Fused write-out of one row (executed by one warp).
//   sm[0..nvalid): keys sorted ascending by (value, index); p = global row id.
//   Applies, in the reference's order (sph/NearestNeighbors.cpp:152-167):
//   A6 `d <= FLT_EPSILON -> 0`, A7 sqrt unless l2_squared, A8 monotone re-sort, A9 self first
//   (sph/utils/GraphUtils.cpp:23-95: swap / overwrite-last-then-swap / shift).
//   A8 (checkDistancesAreNonDecreasing + stable synchronizedSort, NearestNeighbors.cpp:244-269) never
//   fires for L2.  For the inner product faiss returns DESCENDING similarities, so the reference re-sorts
//   every row that is not constant: a stable ascending sort of a non-increasing sequence reverses the
//   order of the runs of equal values and keeps the order inside each run — output slot s takes the
//   element gs + ge - m, m = nvalid-1-s, [gs, ge] = the run of equal values around m.
//   Slots t >= nvalid keep the caller's -1 / -1.0f convention.
__device__ __forceinline__ void write_row(const uint64_t* sm, int nvalid, int k, int64_t p, int metric, int l2_squared,
                                          int64_t* __restrict__ ids_out, float* __restrict__ dist_out, int lane) {
    auto dist_of = [&](uint64_t key) -> float {
        float v = key_val(key);
        if (metric != METRIC_L2) v = -v;
        if (v <= FLT_EPSILON) v = 0.f;
        if (!l2_squared) v = sqrtf(v);
        return v;
    };
    // A8 for the inner product: does the row need the re-sort (any strict decrease)?
    bool resort = false;
    if (metric != METRIC_L2 && nvalid > 1) {
        bool dec = false;
        for (int t = lane + 1; t < nvalid; t += 32) dec |= dist_of(sm[t]) < dist_of(sm[t - 1]);
        resort = __any_sync(0xffffffffu, dec);
    }
    auto at = [&](int s) -> uint64_t {  // key at position s of the A8-ordered row
        if (!resort) return sm[s];
        const int m = nvalid - 1 - s;
        const float v = dist_of(sm[m]);
        int gs = m, ge = m;
        while (gs > 0 && dist_of(sm[gs - 1]) == v) --gs;
        while (ge + 1 < nvalid && dist_of(sm[ge + 1]) == v) ++ge;
        return sm[gs + ge - m];
    };
    // locate self and test "all distances zero"
    int self_pos = 0x7fffffff;
    bool all_zero = true;
    for (int t = lane; t < nvalid; t += 32) {
        const uint64_t key = at(t);
        if (static_cast<int64_t>(key_idx(key)) == p) self_pos = min(self_pos, t);
        if (dist_of(key) != 0.f) all_zero = false;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) self_pos = min(self_pos, __shfl_xor_sync(0xffffffffu, self_pos, o));
    all_zero = __all_sync(0xffffffffu, all_zero);
    const bool found = self_pos != 0x7fffffff;
    // mode 0: identity, 1: swap(0, self_pos), 2: overwrite last with p then swap, 3: shift right
    int mode = 0;
    if (nvalid > 0 && self_pos != 0) mode = found ? 1 : (all_zero ? 2 : 3);

    for (int t = lane; t < k; t += 32) {
        int64_t id = -1;
        float dv = -1.0f;
        if (t < nvalid) {
            int src = t;
            bool is_self = false;
            if (mode == 1) {
                src = (t == 0) ? self_pos : (t == self_pos ? 0 : t);
            } else if (mode == 2) {
                if (t == 0) { src = nvalid - 1; is_self = true; }
                else if (t == nvalid - 1) src = 0;
            } else if (mode == 3) {
                if (t == 0) is_self = true;
                else src = t - 1;
            }
            const uint64_t key = at(src);
            id = is_self ? p : static_cast<int64_t>(key_idx(key));
            dv = (mode == 3 && t == 0) ? 0.f : dist_of(key);
        }
        ids_out[t] = id;
        dist_out[t] = dv;
    }
}

}  // namespace sphb
