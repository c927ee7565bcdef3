// Stage B: perplexity-calibrated probability rows (t-SNE P matrix), one warp per row.
//   B1 sph/utils/HDILibHelper.hpp:23-109   binary search on beta (fp64, fp32 round trip of p)
//   B2 sph/utils/GraphNormalization.cpp:191-338  graph -> column-sorted rows (+ min-sigma fallback)
//   B3 sph/utils/GraphNormalization.cpp:38-189   in place on sparse rows (+ the 1e-10 cursor quirk)
//   B4 sph/utils/HDILibHelper.hpp:112-140        flat variant, tol 1e-5
//   B5 sph/utils/GraphNormalization.cpp:340-411  linear rows
//   §8(f)-3 sph/utils/GraphNormalization.cpp:413-593  exponential (UMAP) rows, at the end of this file
// Distances live in registers (R values per lane), all reductions are warp shuffles,
// the beta loop never touches memory.  HBM traffic is the one read of (ids, dist)
// and the one write of (cols, vals).
#include "common.cuh"
#include <cfloat>

namespace sphb {

// Returns sigma (0 when not found).  On return p[] holds the fp32 probabilities
// exactly as the reference leaves them (normalised if found, uniform if not).
//
// The iterate sequence is the reference's (sph/utils/HDILibHelper.hpp:40-93: fp64 beta, doubling / halving until bracketed,
// then bisection), and every decision that can depend on fp64 detail is taken by the reference's own fp64 evaluation.
// What changed against round 1 (ncu: issue slots 79 % busy, fp64 pipe 41 % — instruction bound, ~25 fp64 instructions
// per exp): an iteration first evaluates H in fp32 on SHIFTED exponents x_j = beta (d_j - d_min) >= 0 (H is invariant
// under the shift, and the shifted sum lies in [1, n], so nothing over- or underflows), together with a rigorous bound
// on |H_fp32 - H_reference|:  dH = sum_j p_j e_j (x_j - E + 1) for relative errors e_j of the terms,
// |e_j| <= 2e-7 x_j + 5e-7 (fp32 product, ex2.approx, the reference's own fp32 round trip of p), hence
// |dH| <= 2e-7 (E[x²] + E (E + 1)) + 5e-7 (2 E + 1) with E = E_p[x]; E[x²] is accumulated in the same pass.  If
// |H_fp32 - ln(perplexity)| exceeds four times that bound plus the tolerance, the branch the reference takes is
// certain and the fp64 evaluation is skipped; near convergence (the last ~6 of ~30 steps) every step is fp64 as before.
template <int R>
__device__ __forceinline__ double perplexity_search(const float (&dist)[R], float (&p)[R], int nn, int lane, double perplexity, double tol,
                                                    int ignore) {
    bool found = false;
    double beta = 1.0;
    double min_beta = -DBL_MAX, max_beta = DBL_MAX;
    double sum = DBL_MIN;
    const double logp = log(perplexity);
    // fp32 screen: shifted distances
    float dd[R];
    float dmin = INFINITY, dmax = -INFINITY;
    int nvalid = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int idx = lane + 32 * r;
        if (idx < nn && idx != ignore) {
            dmin = fminf(dmin, dist[r]);
            dmax = fmaxf(dmax, dist[r]);
            ++nvalid;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        dmin = fminf(dmin, __shfl_xor_sync(0xffffffffu, dmin, o));
        dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
        nvalid += __shfl_xor_sync(0xffffffffu, nvalid, o);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int idx = lane + 32 * r;
        dd[r] = (idx < nn && idx != ignore) ? dist[r] - dmin : INFINITY;  // excluded slots contribute exp(-inf) = 0
    }
    const bool screen_ok = nvalid >= 2 && dmin >= 0.f && dmax < 1.0e30f;
    const float logp_f = static_cast<float>(logp);
    const float tol_f = static_cast<float>(tol) * 1.01f + 1.0e-6f;  // + the fp32 rounding of log(s), E and ln(perplexity)
    int iter = 0;
    while (!found && iter < 200) {
        bool decided = false, up = false;  // up: H > ln(perplexity), beta must grow
        // beta * d_min <= 60: the reference's UNSHIFTED exp(-beta d) then stays >= 1e-26 for the nearest entry, so every term that its
        // fp32 round trip of p flushes to a denormal or zero weighs < 1e-12 of the sum and the error model above holds; beyond that
        // (p underflows for the whole row and the reference's H collapses to log(sum)) only the fp64 replica below is trusted
        if (screen_ok && beta < 1.0e30 && beta > 1.0e-30 && beta * static_cast<double>(dmin) <= 60.0) {
            const float bf = static_cast<float>(beta);
            float s = 0.f, h = 0.f, q = 0.f;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float x = bf * dd[r];              // +inf for excluded slots (bf > 0)
                const float v = __expf(-x);              // ex2.approx(-x log2 e); 0 for x = +inf
                const float xv = (v > 0.f) ? x * v : 0.f;  // inf * 0 guarded
                s += v;
                h += xv;
                q = fmaf(xv, (v > 0.f) ? x : 0.f, q);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                s += __shfl_xor_sync(0xffffffffu, s, o);
                h += __shfl_xor_sync(0xffffffffu, h, o);
                q += __shfl_xor_sync(0xffffffffu, q, o);
            }
            const float E = h / s, E2 = q / s;
            const float Hd = (E + logf(s)) - logp_f;
            const float err = 4.f * (2.0e-7f * (E2 + E * (E + 1.f)) + 5.0e-7f * (2.f * E + 1.f)) + tol_f;
            if (fabsf(Hd) > err) {  // false for NaN: falls through to the fp64 evaluation
                decided = true;
                up = Hd > 0.f;
            }
        }
        if (!decided) {
            double s = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int idx = lane + 32 * r;
                if (idx < nn) {
                    if (idx == ignore) {
                        p[r] = 0.f;
                    } else {
                        const double v = exp(-beta * static_cast<double>(dist[r]));
                        p[r] = static_cast<float>(v);
                        s += v;
                    }
                }
            }
            s = warp_sum(s);
            sum = DBL_MIN + s;
            double h = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int idx = lane + 32 * r;
                if (idx < nn && idx != ignore) h += beta * static_cast<double>(dist[r] * p[r]);
            }
            h = warp_sum(h);
            const double H = (h / sum) + log(sum);
            const double Hdiff = H - logp;
            if (Hdiff < tol && -Hdiff < tol) found = true;
            else up = Hdiff > 0;
        }
        if (!found) {
            if (up) {
                min_beta = beta;
                if (max_beta == DBL_MAX || max_beta == -DBL_MAX) beta *= 2.0;
                else beta = (beta + max_beta) / 2.0;
            } else {
                max_beta = beta;
                if (min_beta == -DBL_MAX || min_beta == DBL_MAX) beta /= 2.0;
                else beta = (beta + min_beta) / 2.0;
            }
        }
        ++iter;
    }
    if (!found) {
        const float v = static_cast<float>(1. / (nn + ((ignore < 0 || ignore >= nn) ? 0 : -1)));
#pragma unroll
        for (int r = 0; r < R; ++r) p[r] = v;
        return 0.0;
    }
#pragma unroll
    for (int r = 0; r < R; ++r) p[r] = static_cast<float>(static_cast<double>(p[r]) / sum);
    return sqrt(1.0 / (2.0 * beta));
}

template <int R>
__device__ __forceinline__ bool all_zero(const float (&p)[R], int nn, int lane) {
    bool z = true;
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (lane + 32 * r < nn && p[r] != 0.f) z = false;
    return __all_sync(0xffffffffu, z);
}
template <int R>
__device__ __forceinline__ void normalize_unit(float (&p)[R], int nn, int lane) {
    double s = 0.0;
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (lane + 32 * r < nn) s += static_cast<double>(p[r]);
    s = warp_sum(s);
#pragma unroll
    for (int r = 0; r < R; ++r) p[r] = static_cast<float>(static_cast<double>(p[r]) / s);
}
// min-sigma fallback (GraphNormalization.cpp:96-130 / :249-282)
template <int R>
__device__ __forceinline__ void min_sigma_fallback(const float (&dist)[R], float (&p)[R], int nn, int lane, bool zero_first) {
#pragma unroll
    for (int r = 0; r < R; ++r) p[r] = dist[r];
    const float same = 1.f / (nn - 1);
    if (all_zero<R>(p, nn, lane)) {
#pragma unroll
        for (int r = 0; r < R; ++r) p[r] = same;
        return;
    }
    normalize_unit<R>(p, nn, lane);
#pragma unroll
    for (int r = 0; r < R; ++r) p[r] = 1.0f - p[r];
    if (all_zero<R>(p, nn, lane)) {
#pragma unroll
        for (int r = 0; r < R; ++r) p[r] = same;
        return;
    }
    if (zero_first && lane == 0) p[0] = 0.f;
    normalize_unit<R>(p, nn, lane);
}

// mode 0: Gaussian (B2), mode 1: linear (B5)
template <int R>
__global__ void __launch_bounds__(128) prow_graph_kernel(const int64_t* __restrict__ ids, const float* __restrict__ dist, int64_t n, int k,
                                                         float fixed_perplexity, int mode, int cap, uint32_t* __restrict__ cols,
                                                         float* __restrict__ vals, double* __restrict__ sigma_out) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    const int ne = k - 1;
    for (int64_t p = gw; p < n; p += nw) {
        float dv[R], pv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            dv[r] = idx < k ? dist[p * k + idx] : 0.f;
            pv[r] = 0.f;
        }
        if (mode == 0) {
            const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : ne / 3.f;
            const double sigma = perplexity_search<R>(dv, pv, k, lane, static_cast<double>(perplexity), 1e-6, 0);
            if (sigma < 0.001) min_sigma_fallback<R>(dv, pv, k, lane, true);
            if (sigma_out && lane == 0) sigma_out[p] = sigma;
        } else {
            // invlin over neighbours 1..k-1, normalise (double sum)
            double s = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int idx = lane + 32 * r;
                pv[r] = (idx >= 1 && idx < k) ? 1.0f / (1.0f + dv[r]) : 0.f;
                s += static_cast<double>(pv[r]);
            }
            s = warp_sum(s);
#pragma unroll
            for (int r = 0; r < R; ++r) pv[r] = static_cast<float>(static_cast<double>(pv[r]) / s);
        }
        // column sort of entries 1..k-1
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < cap) {
                uint64_t key = ~0ull;
                if (idx >= 1 && idx < k) key = (static_cast<uint64_t>(static_cast<uint32_t>(ids[p * k + idx])) << 32) | __float_as_uint(pv[r]);
                sm[idx] = key;
            }
        }
        for (int t = 32 * R + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        for (int e = lane; e < ne; e += 32) {
            const uint64_t key = sm[e];
            cols[p * ne + e] = static_cast<uint32_t>(key >> 32);
            vals[p * ne + e] = __uint_as_float(static_cast<uint32_t>(key));
        }
        __syncwarp();
    }
}

// B3: in place on CSR rows.
template <int R>
__global__ void __launch_bounds__(128) prow_inplace_kernel(int64_t n, const int64_t* __restrict__ rowptr, float* __restrict__ val,
                                                           float fixed_perplexity, int ignore, int* __restrict__ err) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t p = gw; p < n; p += nw) {
        const int64_t b = rowptr[p];
        const int64_t nn64 = rowptr[p + 1] - b;
        if (nn64 <= 1) continue;
        if (nn64 > 32 * R) {
            if (lane == 0) atomicExch(err, 1);
            continue;
        }
        const int nn = static_cast<int>(nn64);
        float dv[R], pv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            dv[r] = idx < nn ? val[b + idx] : 0.f;
            pv[r] = 0.f;
        }
        const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : nn / 3.f;
        const double sigma = perplexity_search<R>(dv, pv, nn, lane, static_cast<double>(perplexity), 1e-6, ignore);
        if (sigma < 0.001) min_sigma_fallback<R>(dv, pv, nn, lane, ignore >= 0);
        // write-back cursor stops at the first p < 1e-10 (GraphNormalization.cpp:132-153)
        int z = 0x7fffffff;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < nn && pv[r] < 1.0e-10f) z = min(z, idx);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z = min(z, __shfl_xor_sync(0xffffffffu, z, o));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < nn && idx < z) val[b + idx] = pv[r];
        }
    }
}

// B4: flat variant -> temp[n*k]
template <int R>
__global__ void __launch_bounds__(128) prow_flat_kernel(const float* __restrict__ dist, int64_t n, int k, double perplexity,
                                                        float* __restrict__ temp) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t p = gw; p < n; p += nw) {
        float dv[R], pv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            dv[r] = idx < k ? dist[p * k + idx] : 0.f;
            pv[r] = 0.f;
        }
        perplexity_search<R>(dv, pv, k, lane, perplexity, 1e-5, 0);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < k) temp[p * k + idx] = pv[r];
        }
    }
}

// The same rows for a RAGGED graph (per-point neighbour counts: merged / symmetrised graphs,
// computeGaussianDistributions / computeLinearDistributions take any GraphBaseInterface): row p is
// [rowptr[p], rowptr[p+1]) with the point itself first, its k_p - 1 output entries start at rowptr[p] - p.
template <int R>
__global__ void __launch_bounds__(128) prow_graph_ragged_kernel(const int64_t* __restrict__ rowptr, const int64_t* __restrict__ ids,
                                                                const float* __restrict__ dist, int64_t n, float fixed_perplexity, int mode,
                                                                int cap, uint32_t* __restrict__ cols, float* __restrict__ vals) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t p = gw; p < n; p += nw) {
        const int64_t b = rowptr[p];
        const int k = static_cast<int>(rowptr[p + 1] - b);
        const int ne = k - 1;
        if (ne <= 0) continue;  // warp-uniform
        const int64_t ob = b - p;
        float dv[R], pv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            dv[r] = idx < k ? dist[b + idx] : 0.f;
            pv[r] = 0.f;
        }
        if (mode == 0) {
            const float perplexity = fixed_perplexity > 0 ? fixed_perplexity : ne / 3.f;
            const double sigma = perplexity_search<R>(dv, pv, k, lane, static_cast<double>(perplexity), 1e-6, 0);
            if (sigma < 0.001) min_sigma_fallback<R>(dv, pv, k, lane, true);
        } else {
            double s = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int idx = lane + 32 * r;
                pv[r] = (idx >= 1 && idx < k) ? 1.0f / (1.0f + dv[r]) : 0.f;
                s += static_cast<double>(pv[r]);
            }
            s = warp_sum(s);
#pragma unroll
            for (int r = 0; r < R; ++r) pv[r] = static_cast<float>(static_cast<double>(pv[r]) / s);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < cap) {
                uint64_t key = ~0ull;
                if (idx >= 1 && idx < k) key = (static_cast<uint64_t>(static_cast<uint32_t>(ids[b + idx])) << 32) | __float_as_uint(pv[r]);
                sm[idx] = key;
            }
        }
        for (int t = 32 * R + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        for (int e = lane; e < ne; e += 32) {
            const uint64_t key = sm[e];
            cols[ob + e] = static_cast<uint32_t>(key >> 32);
            vals[ob + e] = __uint_as_float(static_cast<uint32_t>(key));
        }
        __syncwarp();
    }
}

static int pow2_at_least(int v) {
    int p = 32;
    while (p < v) p <<= 1;
    return p;
}

template <int R>
static int launch_graph(const int64_t* ids, const float* dist, int64_t n, int k, float perplexity, int mode, uint32_t* cols, float* vals,
                        double* sigma, cudaStream_t stream) {
    const int cap = pow2_at_least(32 * R);
    const size_t smem = 4 * static_cast<size_t>(cap) * sizeof(uint64_t);
    SPHB_CUDA(cudaFuncSetAttribute(prow_graph_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    prow_graph_kernel<R><<<static_cast<unsigned>(blocks), 128, smem, stream>>>(ids, dist, n, k, perplexity, mode, cap, cols, vals, sigma);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int prow_graph_launch(const int64_t* d_ids, const float* d_dist, int64_t n, int k, float perplexity, int mode, uint32_t* d_cols,
                      float* d_vals, double* d_sigma, cudaStream_t stream) {
    if (n <= 0) return 0;
    if (k < 2) return fail(2, "prow: k must be >= 2");
    if (k <= 32) return launch_graph<1>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    if (k <= 64) return launch_graph<2>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    if (k <= 96) return launch_graph<3>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    if (k <= 128) return launch_graph<4>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    if (k <= 320) return launch_graph<10>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    if (k <= 1024) return launch_graph<32>(d_ids, d_dist, n, k, perplexity, mode, d_cols, d_vals, d_sigma, stream);
    return fail(5, "prow: k > 1024 unsupported");
}

template <int R>
static int launch_graph_ragged(const int64_t* rowptr, const int64_t* ids, const float* dist, int64_t n, float perplexity, int mode, uint32_t* cols,
                               float* vals, cudaStream_t stream) {
    const int cap = pow2_at_least(32 * R);
    const size_t smem = 4 * static_cast<size_t>(cap) * sizeof(uint64_t);
    SPHB_CUDA(cudaFuncSetAttribute(prow_graph_ragged_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    prow_graph_ragged_kernel<R><<<static_cast<unsigned>(blocks), 128, smem, stream>>>(rowptr, ids, dist, n, perplexity, mode, cap, cols, vals);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int prow_graph_ragged_launch(const int64_t* d_rowptr, const int64_t* d_ids, const float* d_dist, int64_t n, int max_k, float perplexity, int mode,
                             uint32_t* d_cols, float* d_vals, cudaStream_t stream) {
    if (n <= 0 || max_k < 2) return 0;
    if (max_k <= 32) return launch_graph_ragged<1>(d_rowptr, d_ids, d_dist, n, perplexity, mode, d_cols, d_vals, stream);
    if (max_k <= 64) return launch_graph_ragged<2>(d_rowptr, d_ids, d_dist, n, perplexity, mode, d_cols, d_vals, stream);
    if (max_k <= 128) return launch_graph_ragged<4>(d_rowptr, d_ids, d_dist, n, perplexity, mode, d_cols, d_vals, stream);
    if (max_k <= 320) return launch_graph_ragged<10>(d_rowptr, d_ids, d_dist, n, perplexity, mode, d_cols, d_vals, stream);
    if (max_k <= 1024) return launch_graph_ragged<32>(d_rowptr, d_ids, d_dist, n, perplexity, mode, d_cols, d_vals, stream);
    return fail(5, "prow: rows with more than 1024 neighbours unsupported");
}

int prow_inplace_launch(int64_t n, const int64_t* d_rowptr, float* d_val, int max_nnz, float perplexity, int ignore, cudaStream_t stream) {
    if (n <= 0) return 0;
    DevBuf<int> err;
    SPHB_CUDA(err.alloc(1, stream));
    SPHB_CUDA(cudaMemsetAsync(err.p, 0, sizeof(int), stream));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    if (max_nnz <= 128) prow_inplace_kernel<4><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(n, d_rowptr, d_val, perplexity, ignore, err.p);
    else if (max_nnz <= 320) prow_inplace_kernel<10><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(n, d_rowptr, d_val, perplexity, ignore, err.p);
    else if (max_nnz <= 1024) prow_inplace_kernel<32><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(n, d_rowptr, d_val, perplexity, ignore, err.p);
    else return fail(5, "prow_inplace: rows with more than 1024 entries unsupported");
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    int herr = 0;
    SPHB_CUDA(cudaMemcpyAsync(&herr, err.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaStreamSynchronize(stream));
    if (herr) return fail(5, "prow_inplace: row longer than the dispatched register tile");
    return 0;
}

int prow_flat_launch(const float* d_dist, int64_t n, int k, double perplexity, float* d_temp, cudaStream_t stream) {
    if (n <= 0) return 0;
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    if (k <= 128) prow_flat_kernel<4><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(d_dist, n, k, perplexity, d_temp);
    else if (k <= 320) prow_flat_kernel<10><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(d_dist, n, k, perplexity, d_temp);
    else if (k <= 1024) prow_flat_kernel<32><<<static_cast<unsigned>(blocks), 128, 0, stream>>>(d_dist, n, k, perplexity, d_temp);
    else return fail(5, "prow_flat: k > 1024 unsupported");
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}


// ---- §8(f)-3: UMAP-flavoured rows, utils::computeExponentialDistributions (sph/utils/GraphNormalization.cpp:413-593) ----
// The reference hands each row to umappp::neighbor_similarities (libscran-umappp 3.2.0, default options: local_connectivity 1,
// bandwidth 1, fp32).  Per row: rho = first non-zero distance in stored order; sigma solves
//   sum_j [ d_j > rho ? exp(-(d_j - rho) / sigma) : 1 ] = log2(n_neighbours + 1)
// (doubling from 1 until bracketed, then bisection, <= 64 evaluations, |sum - target| < 1e-5), is clamped from below by
// 1e-3 * mean(d), and the row becomes w_j = d_j > rho ? exp(-(d_j - rho) / sigma) : 1.  Same shape as the Gaussian rows: one
// warp per row, the distances in registers, warp-shuffle sums, no memory traffic inside the search.
// vm: bit r set <=> register r of this lane holds a row entry.
template <int R>
__device__ __forceinline__ void umap_row_weights(const float (&d)[R], float (&w)[R], uint32_t vm, int lane) {
    int first = 0x7fffffff, nn = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool v = (vm >> r) & 1u;
        nn += v;
        if (v && d[r] != 0.f) first = min(first, lane + 32 * r);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
        nn += __shfl_xor_sync(0xffffffffu, nn, o);
    }
    float rho = 0.f;
    if (first != 0x7fffffff) {
        float c = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (r == (first >> 5)) c = d[r];
        rho = __shfl_sync(0xffffffffu, c, first & 31);
    }
    int num_le = 0;
    float total = 0.f;
    float delta[R];  // > 0 for the entries beyond rho, 0 marks "not active"
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool v = (vm >> r) & 1u;
        const bool act = v && d[r] > rho;
        num_le += v && !act;
        total += v ? d[r] : 0.f;
        delta[r] = act ? d[r] - rho : 0.f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        num_le += __shfl_xor_sync(0xffffffffu, num_le, o);
        total += __shfl_xor_sync(0xffffffffu, total, o);
    }
    if (num_le == nn) {
#pragma unroll
        for (int r = 0; r < R; ++r) w[r] = 1.f;
        return;
    }
    const float target = log2f(static_cast<float>(nn + 1));
    float sigma = 1.f, lo = 0.f, hi = FLT_MAX;
    for (int iter = 0; iter < 64; ++iter) {
        float part = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (delta[r] > 0.f) part += expf(-delta[r] / sigma);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        const float val = static_cast<float>(num_le) + part;
        if (fabsf(val - target) < 1e-5f) break;
        if (val > target) {
            hi = sigma;
            sigma = (lo + hi) / 2.f;
        } else {
            lo = sigma;
            sigma = (hi == FLT_MAX) ? sigma * 2.f : (lo + hi) / 2.f;
        }
    }
    sigma = fmaxf(sigma, 1e-3f * (total / static_cast<float>(nn)));
#pragma unroll
    for (int r = 0; r < R; ++r) w[r] = delta[r] > 0.f ? expf(-delta[r] / sigma) : 1.f;
}

// Graph form (GraphNormalization.cpp:508-593): neighbours 1..k-1 of a fixed-k graph; sort_cols = 0 keeps the neighbour order
// (T = MapMemEff: storage.emplace_back), 1 sorts the row by column (T = SparseVecSPH: insert).
template <int R>
__global__ void __launch_bounds__(128) prow_exponential_graph_kernel(const int64_t* __restrict__ ids, const float* __restrict__ dist, int64_t n,
                                                                     int k, int sort_cols, int cap, uint32_t* __restrict__ cols,
                                                                     float* __restrict__ vals) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    const int ne = k - 1;
    for (int64_t p = gw; p < n; p += nw) {
        float dv[R], wv[R];
        uint32_t vm = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            dv[r] = idx < ne ? dist[p * k + 1 + idx] : 0.f;
            vm |= (idx < ne ? 1u : 0u) << r;
        }
        umap_row_weights<R>(dv, wv, vm, lane);
        if (!sort_cols) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int idx = lane + 32 * r;
                if (idx < ne) {
                    cols[p * ne + idx] = static_cast<uint32_t>(ids[p * k + 1 + idx]);
                    vals[p * ne + idx] = wv[r];
                }
            }
            continue;
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            if (idx < cap)
                sm[idx] = idx < ne ? (static_cast<uint64_t>(static_cast<uint32_t>(ids[p * k + 1 + idx])) << 32) | __float_as_uint(wv[r]) : ~0ull;
        }
        for (int t = 32 * R + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        for (int e = lane; e < ne; e += 32) {
            const uint64_t key = sm[e];
            cols[p * ne + e] = static_cast<uint32_t>(key >> 32);
            vals[p * ne + e] = __uint_as_float(static_cast<uint32_t>(key));
        }
        __syncwarp();
    }
}

// In-place form on CSR rows (GraphNormalization.cpp:413-506): the stored entries of a row are its "neighbours", an entry whose
// COLUMN equals `ignore` (>= 0) is left out and keeps its value.
template <int R>
__global__ void __launch_bounds__(128) prow_exponential_inplace_kernel(int64_t n, const int64_t* __restrict__ rowptr,
                                                                       const uint32_t* __restrict__ col, float* __restrict__ val, int ignore,
                                                                       int* __restrict__ err) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t p = gw; p < n; p += nw) {
        const int64_t b = rowptr[p];
        const int64_t nn64 = rowptr[p + 1] - b;
        if (nn64 <= 0) continue;
        if (nn64 > 32 * R) {
            if (lane == 0) atomicExch(err, 1);
            continue;
        }
        const int nn = static_cast<int>(nn64);
        float dv[R], wv[R];
        uint32_t vm = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int idx = lane + 32 * r;
            const bool v = idx < nn && !(ignore >= 0 && col != nullptr && col[b + (idx < nn ? idx : 0)] == static_cast<uint32_t>(ignore));
            dv[r] = v ? val[b + idx] : 0.f;
            vm |= (v ? 1u : 0u) << r;
        }
        umap_row_weights<R>(dv, wv, vm, lane);
#pragma unroll
        for (int r = 0; r < R; ++r)
            if ((vm >> r) & 1u) val[b + lane + 32 * r] = wv[r];
    }
}

template <int R>
static int launch_exponential_graph(const int64_t* ids, const float* dist, int64_t n, int k, int sort_cols, uint32_t* cols, float* vals,
                                    cudaStream_t stream) {
    const int cap = pow2_at_least(32 * R);
    const size_t smem = sort_cols ? 4 * static_cast<size_t>(cap) * sizeof(uint64_t) : 0;
    SPHB_CUDA(cudaFuncSetAttribute(prow_exponential_graph_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    prow_exponential_graph_kernel<R><<<static_cast<unsigned>(blocks), 128, smem, stream>>>(ids, dist, n, k, sort_cols, cap, cols, vals);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int prow_exponential_graph_launch(const int64_t* d_ids, const float* d_dist, int64_t n, int k, int sort_cols, uint32_t* d_cols, float* d_vals,
                                  cudaStream_t stream) {
    if (n <= 0) return 0;
    if (k < 2) return fail(2, "prow_exponential: k must be >= 2");
    const int ne = k - 1;
    if (ne <= 32) return launch_exponential_graph<1>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    if (ne <= 64) return launch_exponential_graph<2>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    if (ne <= 96) return launch_exponential_graph<3>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    if (ne <= 128) return launch_exponential_graph<4>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    if (ne <= 320) return launch_exponential_graph<10>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    if (ne <= 1024) return launch_exponential_graph<32>(d_ids, d_dist, n, k, sort_cols, d_cols, d_vals, stream);
    return fail(5, "prow_exponential: k > 1025 unsupported");
}

int prow_exponential_inplace_launch(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, float* d_val, int max_nnz, int ignore,
                                    cudaStream_t stream) {
    if (n <= 0) return 0;
    DevBuf<int> err;
    SPHB_CUDA(err.alloc(1, stream));
    SPHB_CUDA(cudaMemsetAsync(err.p, 0, sizeof(int), stream));
    const int64_t blocks = std::min<int64_t>((n + 3) / 4, static_cast<int64_t>(ctx().num_sms) * 16);
    const unsigned g = static_cast<unsigned>(blocks);
    if (max_nnz <= 128) prow_exponential_inplace_kernel<4><<<g, 128, 0, stream>>>(n, d_rowptr, d_col, d_val, ignore, err.p);
    else if (max_nnz <= 320) prow_exponential_inplace_kernel<10><<<g, 128, 0, stream>>>(n, d_rowptr, d_col, d_val, ignore, err.p);
    else if (max_nnz <= 1024) prow_exponential_inplace_kernel<32><<<g, 128, 0, stream>>>(n, d_rowptr, d_col, d_val, ignore, err.p);
    else return fail(5, "prow_exponential_inplace: rows with more than 1024 entries unsupported");
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    int herr = 0;
    SPHB_CUDA(cudaMemcpyAsync(&herr, err.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaStreamSynchronize(stream));
    if (herr) return fail(5, "prow_exponential_inplace: row longer than the dispatched register tile");
    return 0;
}

}  // namespace sphb
