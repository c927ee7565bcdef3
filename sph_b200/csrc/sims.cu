// Stage D, D1: per-level component similarities (replaces utils::createSimilarities,
// sph/utils/SparseMatrixAlgorithms.cpp:963-1378): D = -ln(sqrt(A) sqrt(A)ᵀ), pruned, per row the k
// smallest distances, column-sorted, L1-normalised (fp64 sum).
//
// Row-wise Gustavson SpGEMM that never materialises the product: S = sqrt(prune(A)) and its
// transpose (built by one radix sort on (column, row)) stay in HBM/L2; one persistent CTA per output
// row accumulates s_ic * s_jc into a dense accumulator row, then selects the bottom k of (-ln C, column)
// with per-warp sorted lists and a final block-wide bitonic merge.  Two accumulator homes:
//   DENSE (n <= what fits beside the lists in 220 KB of shared memory, ~52 k components): the row lives in
//     SHARED memory and the CTA walks the row's entries c in order, one column of Sᵀ at a time — the rows j
//     of a column are distinct, so the 256 threads update distinct accumulators with plain LDS / FADD / STS
//     (no atomics; a barrier between columns), the (row, value) pairs of the next 16 column segments are in
//     flight in registers while the current ones are added.  Every C[i][j] is therefore summed in ascending c
//     — the order of the reference's sparse dot — and the result is deterministic.  Round 1's version of this
//     kernel (float atomics with return into an L2 slab) ran 381 ms at C2 level 1 (ncu); see DESIGN.md §4.5.
//   SLAB (larger n): a private slab per CTA in HBM/L2, float atomics, first-touch list (the round-1 kernel).
// Candidates are screened with a fast fp32 log before the reference's fp64 log is evaluated (D is monotone in C,
// so anything whose approximate D exceeds the running k-th D by more than 1e-4 cannot enter the list).
// Algorithmic bytes: 8 B per expanded product (row id + value of Sᵀ) + 8 n k written.
#include "../../include/sphb.h"
#include "graph_common.cuh"

#include <cfloat>
#include <cstdlib>

namespace sphb {

// Eigen's pruned(ref): keeps |v| > |ref| * NumTraits<float>::dummy_precision() (= 1e-5f)
__device__ __forceinline__ bool eigen_keep(float v, float ref) { return !(fabsf(v) <= fabsf(ref) * 1e-5f); }

__global__ void sims_count_kept(int64_t n, const int64_t* __restrict__ rowptr, const float* __restrict__ val, float prune_val,
                                int64_t* __restrict__ cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r <= n; r += nwarps) {
        int c = 0;
        if (r < n)
            for (int64_t e = rowptr[r] + lane; e < rowptr[r + 1]; e += 32) c += eigen_keep(val[e], prune_val) ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) cnt[r] = c;
    }
}
// S rows (column order preserved) + the (column, row) keys of the transpose
__global__ void sims_build_s(int64_t n, const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ col, const float* __restrict__ val,
                             float prune_val, const int64_t* __restrict__ comp_sizes, const int64_t* __restrict__ srowptr,
                             uint32_t* __restrict__ scol, float* __restrict__ sval, uint64_t* __restrict__ tkeys) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t r = warp; r < n; r += nwarps) {
        int64_t out = srowptr[r];
        const float scale = comp_sizes ? static_cast<float>(sqrt(static_cast<double>(comp_sizes[r]))) : 1.f;
        for (int64_t base = rowptr[r]; base < rowptr[r + 1]; base += 32) {
            const int64_t e = base + lane;
            const bool keep = e < rowptr[r + 1] && eigen_keep(val[e], prune_val);
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int64_t o = out + __popc(mask & ((1u << lane) - 1u));
                float v = sqrtf(val[e]);
                if (comp_sizes) v *= scale;
                scol[o] = col[e];
                sval[o] = v;
                tkeys[o] = (static_cast<uint64_t>(col[e]) << 32) | static_cast<uint64_t>(r);
            }
            out += __popc(mask);
        }
    }
}
__global__ void colptr_from_keys(const uint64_t* __restrict__ keys, int64_t num, int64_t ncols, int64_t* __restrict__ colptr) {
    for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c <= ncols; c += static_cast<int64_t>(gridDim.x) * blockDim.x)
        colptr[c] = c == ncols ? num : lower_bound_u64(keys, num, static_cast<uint64_t>(c) << 32);
}

// block-wide ascending bitonic sort of `cap` (power of two) u64 keys in shared memory
__device__ __forceinline__ void block_bitonic_sort(uint64_t* sm, int cap) {
    for (int k2 = 2; k2 <= cap; k2 <<= 1) {
        for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
            for (int t = threadIdx.x; t < (cap >> 1); t += blockDim.x) {
                const int i = ((t & ~(j2 - 1)) << 1) | (t & (j2 - 1));
                const int l = i | j2;
                const uint64_t a = sm[i], b = sm[l];
                const bool up = (i & k2) == 0;
                if ((a > b) == up) {
                    sm[i] = b;
                    sm[l] = a;
                }
            }
            __syncthreads();
        }
    }
}

struct SimsParams {
    int64_t n;
    const int64_t* srowptr;   // S rows
    const uint32_t* scol;
    const float* sval;
    const int64_t* colptr;    // Sᵀ: for column c the entries tkeys/tval[colptr[c] .. colptr[c+1]) (row = low 32 bits of the key)
    const uint64_t* tkeys;
    const uint32_t* trow;     // low halves of tkeys (row ids), 4 B per product for the DENSE kernel
    const float* tval;
    float prune_val;
    int hdi;                  // 1: createSimilaritiesHDI semantics, 0: createSimilaritiesEigen + bottom-k conversion
    int k, cap;               // cap: per-warp list capacity, power of two >= k + 32
    float* acc;               // SLAB: [grid][n] zero-initialised accumulator slabs
    uint32_t* touched;        // SLAB: [grid][n]
    int ch;                   // DENSE: accumulator floats in shared memory (>= n)
    uint32_t* out_cols;       // [n][k], padded 0xFFFFFFFF
    float* out_vals;          // [n][k]
    int32_t* row_nnz;
};

// testing hook (sphb_debug_force_sims_slab, not in the public header): run the L2-slab kernel where the shared-memory row would fit
static int g_force_slab = 0;
void sims_force_slab(int on) { g_force_slab = on; }

constexpr int SIMS_WARPS = 16;
constexpr int SIMS_EB = 512;   // DENSE: row entries staged per batch (column start, length, s_ic)
constexpr int SIMS_D = 8;      // DENSE: columns per register group (two groups: one being added, one in flight)

struct ColGroup {
    uint32_t j[SIMS_D][2];
    float v[SIMS_D][2], s[SIMS_D];
};

__global__ void sims_split_rows(const uint64_t* __restrict__ keys, int64_t num, uint32_t* __restrict__ rows) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        rows[t] = static_cast<uint32_t>(keys[t]);
}

template <bool DENSE>
__global__ void __launch_bounds__(SIMS_WARPS * 32) sims_row_kernel(const SimsParams P) {
    extern __shared__ uint64_t sm_keys[];  // SIMS_WARPS * cap keys | DENSE: entry batch | accumulator row
    __shared__ unsigned int s_touched, s_valid, s_nsel;
    __shared__ double s_sum;
    constexpr int THREADS = SIMS_WARPS * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* acc = DENSE ? nullptr : P.acc + static_cast<size_t>(blockIdx.x) * P.n;
    uint32_t* touched = DENSE ? nullptr : P.touched + static_cast<size_t>(blockIdx.x) * P.n;
    uint64_t* mylist = sm_keys + static_cast<size_t>(warp) * P.cap;
    // DENSE shared-memory carve-up behind the lists
    int64_t* e_cb = reinterpret_cast<int64_t*>(sm_keys + static_cast<size_t>(SIMS_WARPS) * P.cap);
    int* e_len = reinterpret_cast<int*>(e_cb + SIMS_EB);
    float* e_sic = reinterpret_cast<float*>(e_len + SIMS_EB);
    float* sacc = e_sic + SIMS_EB;
    const int k = P.k, cap = P.cap;
    if constexpr (DENSE) {
        for (int t = threadIdx.x; t < P.ch; t += THREADS) sacc[t] = 0.f;
    }
    for (int64_t i = blockIdx.x; i < P.n; i += gridDim.x) {
        if (threadIdx.x == 0) { s_touched = 0; s_valid = 0; s_nsel = 0; s_sum = 0.0; }
        __syncthreads();
        const int64_t rb = P.srowptr[i], re = P.srowptr[i + 1];
        if constexpr (DENSE) {
            // ---- 1. expand, one column of Sᵀ at a time (distinct rows j within a column: no atomics)
            for (int64_t eb = rb; eb < re; eb += SIMS_EB) {
                const int ne = static_cast<int>(min(static_cast<int64_t>(SIMS_EB), re - eb));
                for (int t = threadIdx.x; t < ne; t += THREADS) {
                    const uint32_t c = P.scol[eb + t];
                    const int64_t cb = P.colptr[c];
                    e_cb[t] = cb;
                    e_len[t] = static_cast<int>(P.colptr[c + 1] - cb);
                    e_sic[t] = P.sval[eb + t];
                }
                __syncthreads();
                // register-staged software pipeline over the columns of this batch: while group A (SIMS_D columns) is
                // added, the (row, value) pairs of group B are in flight.  A thread holds the first two elements of
                // every column of a group (columns longer than 2 * THREADS finish with direct loads); the two groups
                // are separate objects so that every register is addressed statically.  (The first version of this
                // loop drove a flag-per-slot cursor: ncu counted 111 warp instructions per 32 products, issue-bound.)
                ColGroup ga, gb;
                auto load_group = [&](ColGroup& G, int c0) {
#pragma unroll
                    for (int d = 0; d < SIMS_D; ++d) {
                        const int c = c0 + d;
                        if (c < ne) {  // uniform
                            const int len = e_len[c];
                            const int64_t cb = e_cb[c];
                            G.s[d] = e_sic[c];
#pragma unroll
                            for (int e = 0; e < 2; ++e) {
                                const int off = e * THREADS + static_cast<int>(threadIdx.x);
                                if (off < len) {
                                    G.j[d][e] = __ldg(P.trow + cb + off);
                                    G.v[d][e] = __ldg(P.tval + cb + off);
                                }
                            }
                        }
                    }
                };
                auto consume_group = [&](const ColGroup& G, int c0) {
#pragma unroll
                    for (int d = 0; d < SIMS_D; ++d) {
                        const int c = c0 + d;
                        if (c >= ne) break;  // uniform
                        const int len = e_len[c];
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const int off = e * THREADS + static_cast<int>(threadIdx.x);
                            if (off < len) sacc[G.j[d][e]] = __fadd_rn(sacc[G.j[d][e]], __fmul_rn(G.s[d], G.v[d][e]));
                        }
                        if (len > 2 * THREADS) {  // uniform; the rows of one column are distinct, no barrier inside a column
                            const int64_t cb = e_cb[c];
                            for (int off = 2 * THREADS + static_cast<int>(threadIdx.x); off < len; off += THREADS) {
                                const uint32_t j = __ldg(P.trow + cb + off);
                                sacc[j] = __fadd_rn(sacc[j], __fmul_rn(G.s[d], __ldg(P.tval + cb + off)));
                            }
                        }
                        __syncthreads();  // the next column may hit the same rows from other threads
                    }
                };
                load_group(ga, 0);
                for (int c0 = 0; c0 < ne; c0 += 2 * SIMS_D) {
                    load_group(gb, c0 + SIMS_D);
                    consume_group(ga, c0);
                    load_group(ga, c0 + 2 * SIMS_D);
                    consume_group(gb, c0 + SIMS_D);
                }
                __syncthreads();  // e_* are rewritten by the next batch
            }
        } else {
            // ---- 1. expand: C[i][j] += s_ic * s_jc, one warp per entry c of row i, lanes over column c of S
            for (int64_t e = rb + warp; e < re; e += SIMS_WARPS) {
                const uint32_t c = P.scol[e];
                const float sic = P.sval[e];
                const int64_t cb = P.colptr[c], ce = P.colptr[c + 1];
                for (int64_t f = cb + lane; f < ce; f += 32) {
                    const uint32_t j = static_cast<uint32_t>(P.tkeys[f]);
                    const float old = atomicAdd(acc + j, sic * P.tval[f]);
                    if (old == 0.f) touched[atomicAdd(&s_touched, 1u)] = j;
                }
            }
            __syncthreads();
        }
        const unsigned int T = DENSE ? static_cast<unsigned int>(P.n) : s_touched;
        // ---- 2. (Eigen variant) number of entries that survive the product prune, diagonal included
        if (!P.hdi) {
            unsigned int v = 0;
            for (unsigned int t = threadIdx.x; t < T; t += blockDim.x) v += eigen_keep(DENSE ? sacc[t] : acc[touched[t]], P.prune_val) ? 1u : 0u;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && v) atomicAdd(&s_valid, v);
            __syncthreads();
        }
        const bool multi = s_valid > 1;
        // ---- 3. per-warp bottom-k of (D, column); the accumulator is reset on the way
        int cnt = 0;
        uint64_t kth = ~0ull;
        float dk_screen = INFINITY;  // running k-th D + margin: a candidate whose fast-log D exceeds it cannot enter the list
        for (unsigned int base = warp * 32; base < T; base += SIMS_WARPS * 32) {
            const unsigned int t = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (t < T) {
                const uint32_t j = DENSE ? t : touched[t];
                float c;
                if constexpr (DENSE) { c = sacc[t]; sacc[t] = 0.f; }
                else { c = acc[j]; acc[j] = 0.f; }
                const bool diag = static_cast<int64_t>(j) == i;
                if (eigen_keep(c, P.prune_val) && (diag || c == 1.f || !(-__logf(c) > dk_screen))) {
                    float dv;
                    bool use = true;
                    if (P.hdi) {
                        use = !diag;                                                                  // :1250-1252
                        dv = c == 1.f ? c : -1.f * static_cast<float>(log(static_cast<double>(c)));  // :1290-1292
                    } else {
                        dv = diag ? c : -1.f * static_cast<float>(log(static_cast<double>(c)));      // :1143-1153
                        if (multi && (diag || dv == 0.f)) use = false;                        // removeDiagonalElements(keepSingleEntry)
                    }
                    if (use) {
                        key = make_key(dv, j);
                        pass = key <= kth;
                    }
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) mylist[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int u = cnt + lane; u < cap; u += 32) mylist[u] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(mylist, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? mylist[k - 1] : ~0ull;
                    if (cnt == k) dk_screen = key_val(kth) + 1e-4f;
                }
            }
        }
        for (int u = cnt + lane; u < cap; u += 32) mylist[u] = ~0ull;
        __syncthreads();
        // ---- 4. merge the warps' lists: block-wide sort of all SIMS_WARPS * cap keys, the first k are the row
        block_bitonic_sort(sm_keys, SIMS_WARPS * cap);
        int m = 0;
        for (int t = threadIdx.x; t < k; t += blockDim.x) m += sm_keys[t] != ~0ull ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
        if (lane == 0 && m) atomicAdd(&s_nsel, static_cast<unsigned int>(m));
        __syncthreads();
        const int nsel = static_cast<int>(s_nsel);
        // fp64 sum of the selected distances (SparseMatrixAlgorithms.cpp:1349-1369)
        if (warp == 0) {
            double s = 0.0;
            for (int t = lane; t < nsel; t += 32) s += static_cast<double>(key_val(sm_keys[t]));
            s = warp_sum(s);
            if (lane == 0) s_sum = s;
        }
        __syncthreads();
        // re-key by (column, value bits) and sort the first pow2 >= k slots by column
        int cap2 = 32;
        while (cap2 < k) cap2 <<= 1;
        for (int t = threadIdx.x; t < cap2; t += blockDim.x) {
            uint64_t key = ~0ull;
            if (t < nsel) {
                const uint64_t kk = sm_keys[t];
                key = (static_cast<uint64_t>(key_idx(kk)) << 32) | __float_as_uint(key_val(kk));
            }
            sm_keys[t] = key;
        }
        __syncthreads();
        block_bitonic_sort(sm_keys, cap2);
        const double sum = s_sum;
        const bool divide = P.hdi || sum > 0.0;
        for (int t = threadIdx.x; t < k; t += blockDim.x) {
            uint32_t oc = 0xFFFFFFFFu;
            float ov = 0.f;
            if (t < nsel) {
                const uint64_t key = sm_keys[t];
                oc = static_cast<uint32_t>(key >> 32);
                ov = __uint_as_float(static_cast<uint32_t>(key));
                if (divide) ov /= static_cast<float>(sum);
            }
            P.out_cols[i * k + t] = oc;
            P.out_vals[i * k + t] = ov;
        }
        if (threadIdx.x == 0) P.row_nnz[i] = nsel;
        __syncthreads();
    }
}

int create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                            int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s) {
    if (n >= (1ll << 32)) return fail(SPHB_ERR_UNSUPPORTED, "create_similarities: n must be < 2^32");
    if (k > 992) return fail(SPHB_ERR_UNSUPPORTED, "create_similarities: k > 992 unsupported");
    const int sms = ctx().num_sms;
    const bool hdi = variant == 1 || (variant == 0 && n > 5000);
    // ---- S = sqrt(prune(A)) and its transpose
    DevBuf<int64_t> cnt, srowptr, colptr;
    SPHB_CUDA(cnt.alloc(n + 1, s));
    SPHB_CUDA(srowptr.alloc(n + 1, s));
    SPHB_CUDA(colptr.alloc(n + 1, s));
    const unsigned g = static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((n * 32 + 255) / 256, static_cast<int64_t>(sms) * 8)));
    sims_count_kept<<<g, 256, 0, s>>>(n, d_rowptr, d_val, prune_val, cnt.p);
    count_launch();
    if (int rc = exclusive_sum<int64_t>(cnt.p, srowptr.p, n + 1, s)) return rc;
    int64_t snnz = 0;
    SPHB_CUDA(cudaMemcpyAsync(&snnz, srowptr.p + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
    SPHB_CUDA(cudaStreamSynchronize(s));
    (void)nnz;
    DevBuf<uint32_t> scol;
    DevBuf<float> sval, tval;
    DevBuf<uint64_t> tk0, tk1;
    SPHB_CUDA(scol.alloc(std::max<int64_t>(snnz, 1), s));
    SPHB_CUDA(sval.alloc(std::max<int64_t>(snnz, 1), s));
    SPHB_CUDA(tval.alloc(std::max<int64_t>(snnz, 1), s));
    SPHB_CUDA(tk0.alloc(std::max<int64_t>(snnz, 1), s));
    SPHB_CUDA(tk1.alloc(std::max<int64_t>(snnz, 1), s));
    sims_build_s<<<g, 256, 0, s>>>(n, d_rowptr, d_col, d_val, prune_val, d_comp_sizes, srowptr.p, scol.p, sval.p, tk0.p);
    count_launch();
    if (int rc = sort_pairs<float>(tk0.p, tk1.p, sval.p, tval.p, snnz, 32 + bits_for(static_cast<uint64_t>(n)), s)) return rc;
    colptr_from_keys<<<static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((n + 256) / 256, static_cast<int64_t>(sms) * 8))), 256, 0, s>>>(tk1.p, snnz, n, colptr.p);
    count_launch();

    // ---- the row kernel
    int cap = 64;
    while (cap < k + 32) cap <<= 1;
    const size_t list_bytes = static_cast<size_t>(SIMS_WARPS) * cap * sizeof(uint64_t);
    SimsParams P{};
    P.n = n; P.srowptr = srowptr.p; P.scol = scol.p; P.sval = sval.p; P.colptr = colptr.p; P.tkeys = tk1.p; P.tval = tval.p;
    P.prune_val = prune_val; P.hdi = hdi ? 1 : 0; P.k = static_cast<int>(k); P.cap = cap;
    P.out_cols = d_out_cols; P.out_vals = d_out_vals; P.row_nnz = d_row_nnz;
    const size_t dense_fixed = list_bytes + static_cast<size_t>(SIMS_EB) * 16;
    const size_t dense_smem = dense_fixed + static_cast<size_t>(n) * sizeof(float);
    DevBuf<float> acc;
    DevBuf<uint32_t> touched, trow;
    if (dense_smem <= 220 * 1024 && !g_force_slab) {
        SPHB_CUDA(trow.alloc(std::max<int64_t>(snnz, 1), s));
        sims_split_rows<<<g, 256, 0, s>>>(tk1.p, snnz, trow.p);
        count_launch();
        P.trow = trow.p;
        P.ch = static_cast<int>(n);
        SPHB_CUDA(cudaFuncSetAttribute(sims_row_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(dense_smem)));
        const int64_t grid = std::min<int64_t>(n, static_cast<int64_t>(sms) * (dense_smem <= 100 * 1024 ? 2 : 1));
        sims_row_kernel<true><<<static_cast<unsigned>(grid), SIMS_WARPS * 32, dense_smem, s>>>(P);
    } else {
        SPHB_CUDA(cudaFuncSetAttribute(sims_row_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(list_bytes)));
        int64_t grid = std::min<int64_t>(n, static_cast<int64_t>(sms) * 2);
        // bound the accumulator slabs to 8 GiB
        while (grid > 1 && static_cast<double>(grid) * n * 8.0 > 8.0 * (1ull << 30)) grid /= 2;
        SPHB_CUDA(acc.alloc(static_cast<size_t>(grid) * n, s));
        SPHB_CUDA(touched.alloc(static_cast<size_t>(grid) * n, s));
        SPHB_CUDA(cudaMemsetAsync(acc.p, 0, static_cast<size_t>(grid) * n * sizeof(float), s));
        P.acc = acc.p; P.touched = touched.p;
        sims_row_kernel<false><<<static_cast<unsigned>(grid), SIMS_WARPS * 32, list_bytes, s>>>(P);
    }
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace sphb
