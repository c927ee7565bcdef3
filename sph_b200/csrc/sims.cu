// Stage D, D1: per-level component similarities (replaces utils::createSimilarities,
// sph/utils/SparseMatrixAlgorithms.cpp:963-1378): D = -ln(sqrt(A) sqrt(A)ᵀ), pruned, per row the k
// smallest distances, column-sorted, L1-normalised (fp64 sum).
//
// Row-wise Gustavson SpGEMM that never materialises the product: S = sqrt(prune(A)) and its
// transpose (built by one radix sort on (column, row)) stay in HBM/L2; one persistent CTA per output
// row accumulates s_ic * s_jc into a dense accumulator row, then selects the bottom k of (-ln C, column).  Two accumulator homes:
//   CLASS (n up to ~52 k components: 16 x ceil(n / 16) floats beside the lists in 220 KB of SHARED memory): sims_class_kernel
//     below — rows j split into 16 classes (j mod 16), one warp per class, class-major transpose, no block barrier and no
//     atomics inside a row; every C[i][j] is summed in ascending c (the order of the reference's sparse dot), deterministic.
//     History at C2 level 1: L2 float atomics 381 ms (round 1) -> one column at a time with a block barrier per column 137 ->
//     register pipeline 91 -> class-partitioned 85 -> histogram candidate selection 56 -> 8-byte (row, value) records 52 ms.
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
    int64_t r0, r1;           // rows of the result this launch computes (a slab of output rows: SURVEY 8(e), D1 sharded by output row)
    const int64_t* srowptr;   // S rows
    const uint32_t* scol;
    const float* sval;
    const int64_t* colptr;    // Sᵀ: for column c the entries tkeys/tval[colptr[c] .. colptr[c+1]) (row = low 32 bits of the key)
    const uint64_t* tkeys;
    const float* tval;
    float prune_val;
    int hdi;                  // 1: createSimilaritiesHDI semantics, 0: createSimilaritiesEigen + bottom-k conversion
    int k, cap;               // cap: per-warp list capacity, power of two >= k + 32
    float* acc;               // SLAB: [grid][n] zero-initialised accumulator slabs
    uint32_t* touched;        // SLAB: [grid][n]
    uint32_t* out_cols;       // [n][k], padded 0xFFFFFFFF
    float* out_vals;          // [n][k]
    int32_t* row_nnz;
};

// testing hook (sphb_debug_force_sims_slab, not in the public header): run the L2-slab kernel where the shared-memory row would fit
static int g_force_slab = 0;
void sims_force_slab(int on) { g_force_slab = on; }

constexpr int SIMS_EB = 512;   // class kernel: row entries staged per batch (column, s_ic)


constexpr int SIMS_WARPS = 16;

// SLAB flavour (n too large for a shared-memory accumulator row): a private slab per CTA in HBM/L2, float atomics with return,
// first-touch list.
// BITMAP (n bits fit beside the lists in shared memory, n up to ~1.6 M): the products are fire-and-forget reductions
// (red.global.add.f32, no return value to wait for) and the rows touched are recorded in a shared-memory bitmap that is turned
// into the first-touch list afterwards.  Without it (larger n) every product is an atomic WITH return whose old value tells the
// first touch — the returning atomics bound the round-1 kernel at 1.8e10 products/s whatever the number of CTAs.
template <bool BITMAP>
__global__ void __launch_bounds__(SIMS_WARPS * 32) sims_row_kernel(const SimsParams P) {
    extern __shared__ uint64_t sm_keys[];  // SIMS_WARPS * cap keys | BITMAP: ceil(n / 32) words
    __shared__ unsigned int s_touched, s_valid, s_nsel;
    __shared__ double s_sum;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* acc = P.acc + static_cast<size_t>(blockIdx.x) * P.n;
    uint32_t* touched = P.touched + static_cast<size_t>(blockIdx.x) * P.n;
    uint64_t* mylist = sm_keys + static_cast<size_t>(warp) * P.cap;
    uint32_t* bm = reinterpret_cast<uint32_t*>(sm_keys + static_cast<size_t>(SIMS_WARPS) * P.cap);
    const int nwords = static_cast<int>((P.n + 31) >> 5);
    const int k = P.k, cap = P.cap;
    if constexpr (BITMAP)
        for (int t = threadIdx.x; t < nwords; t += blockDim.x) bm[t] = 0u;
    for (int64_t i = P.r0 + blockIdx.x; i < P.r1; i += gridDim.x) {
        if (threadIdx.x == 0) { s_touched = 0; s_valid = 0; s_nsel = 0; s_sum = 0.0; }
        __syncthreads();
        const int64_t rb = P.srowptr[i], re = P.srowptr[i + 1];
        // ---- 1. expand: C[i][j] += s_ic * s_jc, one warp per entry c of row i, lanes over column c of S
        for (int64_t e = rb + warp; e < re; e += SIMS_WARPS) {
            const uint32_t c = P.scol[e];
            const float sic = P.sval[e];
            const int64_t cb = P.colptr[c], ce = P.colptr[c + 1];
            if constexpr (BITMAP) {
#pragma unroll 4
                for (int64_t f = cb + lane; f < ce; f += 32) {
                    const uint32_t j = static_cast<uint32_t>(P.tkeys[f]);
                    atomicAdd(acc + j, sic * P.tval[f]);  // result unused: RED
                    const uint32_t bit = 1u << (j & 31u);
                    if (!(bm[j >> 5] & bit)) atomicOr(&bm[j >> 5], bit);
                }
            } else {
                for (int64_t f = cb + lane; f < ce; f += 32) {
                    const uint32_t j = static_cast<uint32_t>(P.tkeys[f]);
                    const float old = atomicAdd(acc + j, sic * P.tval[f]);
                    if (old == 0.f) touched[atomicAdd(&s_touched, 1u)] = j;
                }
            }
        }
        if constexpr (BITMAP) {
            __threadfence();  // the reductions of this thread are performed before the accumulators are read back
            __syncthreads();
            for (int wi = threadIdx.x; wi < nwords; wi += blockDim.x) {
                uint32_t bits = bm[wi];
                if (bits) {
                    bm[wi] = 0u;
                    unsigned int at = atomicAdd(&s_touched, static_cast<unsigned int>(__popc(bits)));
                    while (bits) {
                        touched[at++] = (static_cast<uint32_t>(wi) << 5) | static_cast<uint32_t>(__ffs(bits) - 1);
                        bits &= bits - 1;
                    }
                }
            }
        }
        __syncthreads();
        const unsigned int T = s_touched;
        // ---- 2. (Eigen variant) number of entries that survive the product prune, diagonal included
        if (!P.hdi) {
            unsigned int v = 0;
            for (unsigned int t = threadIdx.x; t < T; t += blockDim.x) v += eigen_keep(__ldcg(acc + touched[t]), P.prune_val) ? 1u : 0u;
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
                const uint32_t j = touched[t];
                const float c = __ldcg(acc + j);
                acc[j] = 0.f;
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
            P.out_cols[(i - P.r0) * k + t] = oc;
            P.out_vals[(i - P.r0) * k + t] = ov;
        }
        if (threadIdx.x == 0) P.row_nnz[i - P.r0] = nsel;
        __syncthreads();
    }
}

// ---- DENSE, class-partitioned (round 2, second version) -------------------------------------------------------------------
// The column-at-a-time kernel above needs a block barrier after every column of Sᵀ (two columns may hit the same row j from
// different threads) and is latency-bound by them: 1 233 cycles per column at C2 level 1, issue slots 35 % busy, more warps
// help, fewer hurt (16 warps 90 ms, 8 warps 109-140 ms).  Here the rows j are split into 16 CLASSES (j mod 16), warp w owns
// class w: Sᵀ is stored class-major (one radix sort on (class, column, row)), so the entries of column c that belong to class
// w are one contiguous segment, and warp w walks the columns of row i in order adding only its own segment into its own
// accumulator sub-array acc[w][j >> 4].  No two warps ever touch the same accumulator, so there is NO block barrier inside a
// row's expansion; inside a warp a column's rows are distinct and columns are processed in order (a __syncwarp between them),
// so every C[i][j] is still summed in ascending c — bit-identical to the kernel above, deterministic.
constexpr int SC_WARPS = 16;
constexpr int SC_D = 4;   // columns per register group
constexpr int SC_E = 2;   // entries of a (column, class) segment per lane held in registers (segments average n_col / 16)

struct ClassGroup {
    uint2 r[SC_D][SC_E];  // (row j, value bits) records of the class-major transpose
    float s[SC_D];
    int len[SC_D];
    int64_t b[SC_D];
};

__global__ void sims_zip_records(const uint64_t* __restrict__ keys, const float* __restrict__ val, int64_t num, uint2* __restrict__ rec) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        rec[t] = make_uint2(static_cast<uint32_t>(keys[t]), __float_as_uint(val[t]));
}

__global__ void sims_class_ptr(const uint64_t* __restrict__ keys, int64_t num, int64_t n, int cbits, int64_t* __restrict__ cptr) {
    const int64_t total = static_cast<int64_t>(SC_WARPS) * (n + 1);
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const int64_t w = t / (n + 1), c = t - w * (n + 1);
        cptr[t] = lower_bound_u64(keys, num, (static_cast<uint64_t>(w) << (32 + cbits)) | (static_cast<uint64_t>(c) << 32));
    }
}
__global__ void sims_class_keys(const uint64_t* __restrict__ tkeys, int64_t num, int cbits, uint64_t* __restrict__ out) {
    // (column, row) -> (row mod 16, column, row)
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t < num; t += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        const uint64_t k = tkeys[t];
        out[t] = ((k & 15ull) << (32 + cbits)) | k;
    }
}

// Two deeper-prefetch variants of the expansion were tried and lost against the register pipeline below (52 ms at C2 level 1):
// one cp.async.bulk + mbarrier per segment into a per-warp byte ring (139 ms: 300-byte bulk copies are far too small for the
// TMA unit) and 16-byte cp.async copies into a three-stage per-warp buffer (81 ms: the bookkeeping per stage costs more than the
// extra bytes in flight gain).  Both are deleted; tools/r02_sims_ab.sh holds the A/B.
__global__ void __launch_bounds__(SC_WARPS * 32) sims_class_kernel(const SimsParams P, const int64_t* __restrict__ cptr, const uint2* __restrict__ trec, int CH) {
    extern __shared__ uint64_t sm_keys[];  // SC_WARPS * cap keys | entry batch | SC_WARPS * CH accumulators
    __shared__ unsigned int s_valid, s_nsel, s_ncand, s_bin, s_above, s_nreg;
    __shared__ double s_sum;
    constexpr int THREADS = SC_WARPS * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* mylist = sm_keys + static_cast<size_t>(warp) * P.cap;
    uint32_t* e_c = reinterpret_cast<uint32_t*>(sm_keys + static_cast<size_t>(SC_WARPS) * P.cap);
    float* e_sic = reinterpret_cast<float*>(e_c + SIMS_EB);
    float* sacc = e_sic + SIMS_EB;
    float* myacc = sacc + static_cast<size_t>(warp) * CH;
    const int64_t* myptr = cptr + static_cast<int64_t>(warp) * (P.n + 1);
    const int k = P.k, cap = P.cap;
    for (int t = threadIdx.x; t < SC_WARPS * CH; t += THREADS) sacc[t] = 0.f;
    if (threadIdx.x == 0) s_nreg = 0;
    for (int64_t i = P.r0 + blockIdx.x; i < P.r1; i += gridDim.x) {
        if (threadIdx.x == 0) { s_valid = 0; s_nsel = 0; s_sum = 0.0; }
        __syncthreads();
        const int64_t rb = P.srowptr[i], re = P.srowptr[i + 1];
        // ---- 1. expand: warp w adds the class-w segment of every column of row i, columns in order
        for (int64_t eb = rb; eb < re; eb += SIMS_EB) {
            const int ne = static_cast<int>(min(static_cast<int64_t>(SIMS_EB), re - eb));
            for (int t = threadIdx.x; t < ne; t += THREADS) {
                e_c[t] = P.scol[eb + t];
                e_sic[t] = P.sval[eb + t];
            }
            __syncthreads();
            for (int t0 = 0; t0 < ne; t0 += 32) {
                // lane l looks up the segment of column t0 + l once; the columns are then consumed in order via shuffles
                int64_t b_l = 0;
                int len_l = 0;
                float s_l = 0.f;
                if (t0 + lane < ne) {
                    const uint32_t c = e_c[t0 + lane];
                    b_l = __ldg(myptr + c);
                    len_l = static_cast<int>(__ldg(myptr + c + 1) - b_l);
                    s_l = e_sic[t0 + lane];
                }
                const int nc = min(32, ne - t0);
                ClassGroup ga, gb;
                auto load_group = [&](ClassGroup& G, int tt0) {
#pragma unroll
                    for (int d = 0; d < SC_D; ++d) {
                        const int tt = tt0 + d;  // warp-uniform
                        G.len[d] = 0;
                        if (tt < nc) {
                            G.b[d] = __shfl_sync(0xffffffffu, b_l, tt);
                            G.len[d] = __shfl_sync(0xffffffffu, len_l, tt);
                            G.s[d] = __shfl_sync(0xffffffffu, s_l, tt);
#pragma unroll
                            for (int e = 0; e < SC_E; ++e) {
                                const int off = e * 32 + lane;
                                if (off < G.len[d]) G.r[d][e] = __ldg(trec + G.b[d] + off);
                            }
                        }
                    }
                };
                auto consume_group = [&](const ClassGroup& G) {
#pragma unroll
                    for (int d = 0; d < SC_D; ++d) {
                        const int len = G.len[d];  // warp-uniform
                        if (len == 0) continue;
#pragma unroll
                        for (int e = 0; e < SC_E; ++e) {
                            const int off = e * 32 + lane;
                            if (off < len) {
                                const uint32_t q = G.r[d][e].x >> 4;
                                myacc[q] = __fadd_rn(myacc[q], __fmul_rn(G.s[d], __uint_as_float(G.r[d][e].y)));
                            }
                        }
                        for (int off = SC_E * 32 + lane; off < len; off += 32) {  // long segments: the rows of one column are distinct
                            const uint2 r = __ldg(trec + G.b[d] + off);
                            myacc[r.x >> 4] = __fadd_rn(myacc[r.x >> 4], __fmul_rn(G.s[d], __uint_as_float(r.y)));
                        }
                        __syncwarp();  // the next column may hit the same rows from other lanes
                    }
                };
                load_group(ga, 0);
                for (int tt0 = 0; tt0 < nc; tt0 += 2 * SC_D) {
                    load_group(gb, tt0 + SC_D);
                    consume_group(ga);
                    load_group(ga, tt0 + 2 * SC_D);
                    consume_group(gb);
                }
            }
            __syncthreads();  // e_* are rewritten by the next batch; after the last batch: all classes complete
        }
        // ---- 2. (Eigen variant) number of entries that survive the product prune, diagonal included
        if (!P.hdi) {
            unsigned int v = 0;
            for (int t = threadIdx.x; t < SC_WARPS * CH; t += THREADS) v += eigen_keep(sacc[t], P.prune_val) ? 1u : 0u;  // padding slots are 0
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && v) atomicAdd(&s_valid, v);
            __syncthreads();
        }
        const bool multi = s_valid > 1;
        // ---- 3. candidates for the bottom-k of (D, column).  D = -ln C falls with C, so the row's k entries are among the k largest
        // C plus the few "special" entries whose D does not follow C (the diagonal, C == 1).  Two histogram levels over the
        // float bits of C (10 bits each: exponent + 12 mantissa bits) locate a threshold with at least k regular entries above it;
        // everything above (1 - 1e-5) of the threshold is a candidate (entries whose D could tie with the k-th after the fp32
        // rounding of the logarithm differ from it by less than 2e-6 in C, D <= 21), typically k + a few.  Only candidates pay the
        // fp64 logarithm and the sort.  (The per-warp sorted lists this replaces were 32 % of the kernel's instructions.)
        unsigned int* hist = reinterpret_cast<unsigned int*>(sm_keys);        // 1024 bins
        uint64_t* cand = sm_keys + 512;                                       // up to SIMS_NCAND keys behind the histogram
        const int SIMS_NCAND = min(1024, SC_WARPS * cap - 512);  // keys that fit behind the histogram (cap >= 64)
        auto is_special = [&](uint32_t j, float c) { return static_cast<int64_t>(j) == i || c == 1.f; };
        auto top_bin = [&](unsigned int need) -> unsigned int {  // highest bin B with (entries in bins >= B) >= need; warp 0, result in s_bin / s_above
            unsigned int above = 0, found = 0xffffffffu;
            for (int base = 1024 - 32; base >= 0 && found == 0xffffffffu; base -= 32) {
                const unsigned int h = hist[base + 31 - lane];  // lane 0 holds the highest bin of the chunk
                unsigned int incl = h;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned int up = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += up;
                }
                const unsigned m = __ballot_sync(0xffffffffu, above + incl >= need);
                if (m) {
                    const int l = __ffs(m) - 1;
                    found = static_cast<unsigned int>(base + 31 - l);
                    above += __shfl_sync(0xffffffffu, incl, l) - __shfl_sync(0xffffffffu, h, l);  // entries strictly above the bin
                } else {
                    above += __shfl_sync(0xffffffffu, incl, 31);
                }
            }
            if (lane == 0) { s_bin = found; s_above = above; }
            return found;
        };
        for (int t = threadIdx.x; t < 1024; t += THREADS) hist[t] = 0;
        if (threadIdx.x == 0) { s_ncand = 0; s_bin = 0xffffffffu; s_above = 0; }
        __syncthreads();
        {
            unsigned int nreg = 0;
            for (int q = lane; q < CH; q += 32) {
                const float c = myacc[q];
                const uint32_t j = (static_cast<uint32_t>(q) << 4) | static_cast<uint32_t>(warp);
                if (eigen_keep(c, P.prune_val) && c > 0.f && !is_special(j, c)) {
                    atomicAdd(&hist[(__float_as_uint(c) >> 21) & 1023u], 1u);
                    ++nreg;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) nreg += __shfl_xor_sync(0xffffffffu, nreg, o);
            if (lane == 0 && nreg) atomicAdd(&s_nreg, nreg);
        }
        __syncthreads();
        const unsigned int nreg_all = s_nreg;
        uint32_t thr_bits = 0;  // 0: every regular entry is a candidate
        if (nreg_all > static_cast<unsigned int>(k)) {
            if (warp == 0) top_bin(static_cast<unsigned int>(k));
            __syncthreads();
            const unsigned int b1 = s_bin, above1 = s_above;
            for (int t = threadIdx.x; t < 1024; t += THREADS) hist[t] = 0;
            __syncthreads();
            for (int q = lane; q < CH; q += 32) {
                const float c = myacc[q];
                const uint32_t j = (static_cast<uint32_t>(q) << 4) | static_cast<uint32_t>(warp);
                const uint32_t bits = __float_as_uint(c);
                if (eigen_keep(c, P.prune_val) && c > 0.f && !is_special(j, c) && ((bits >> 21) & 1023u) == b1) atomicAdd(&hist[(bits >> 11) & 1023u], 1u);
            }
            __syncthreads();
            if (warp == 0) top_bin(static_cast<unsigned int>(k) - above1);
            __syncthreads();
            thr_bits = (b1 << 21) | (s_bin << 11);
        }
        const float thr = __uint_as_float(thr_bits) * (1.f - 1e-5f);
        // collect + reset
        for (int q = lane; q < CH; q += 32) {
            const float c = myacc[q];
            const uint32_t j = (static_cast<uint32_t>(q) << 4) | static_cast<uint32_t>(warp);
            if (eigen_keep(c, P.prune_val)) {
                const bool diag = static_cast<int64_t>(j) == i;
                if (diag || c == 1.f || !(c > 0.f) || c >= thr) {
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
                        const unsigned int slot = atomicAdd(&s_ncand, 1u);
                        if (slot < static_cast<unsigned int>(SIMS_NCAND)) cand[slot] = make_key(dv, j);
                    }
                }
            }
        }
        __syncthreads();
        const unsigned int ncand = s_ncand;
        if (threadIdx.x == 0) s_nreg = 0;
        if (ncand <= static_cast<unsigned int>(SIMS_NCAND)) {
            for (int q = lane; q < CH; q += 32) myacc[q] = 0.f;
            int capc = 32;
            while (capc < static_cast<int>(ncand)) capc <<= 1;
            for (int t = static_cast<int>(ncand) + threadIdx.x; t < capc; t += THREADS) cand[t] = ~0ull;
            __syncthreads();
            block_bitonic_sort(cand, capc);
            for (int t = threadIdx.x; t < capc && t < k; t += THREADS) sm_keys[t] = cand[t];  // dst index = src index - 512: a thread's second read follows its first write
            __syncthreads();
            for (int t = capc + threadIdx.x; t < k; t += THREADS) sm_keys[t] = ~0ull;
            __syncthreads();
        } else {
        // more candidates than the list holds (massive ties at the threshold): the per-warp sorted lists of the first version
        int cnt = 0;
        uint64_t kth = ~0ull;
        float dk_screen = INFINITY;
        for (int base = 0; base < CH; base += 32) {
            const int q = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (q < CH) {
                const uint32_t j = (static_cast<uint32_t>(q) << 4) | static_cast<uint32_t>(warp);
                const float c = myacc[q];
                myacc[q] = 0.f;
                const bool diag = static_cast<int64_t>(j) == i;
                if (eigen_keep(c, P.prune_val) && (diag || c == 1.f || !(-__logf(c) > dk_screen))) {
                    float dv;
                    bool use = true;
                    if (P.hdi) {
                        use = !diag;
                        dv = c == 1.f ? c : -1.f * static_cast<float>(log(static_cast<double>(c)));
                    } else {
                        dv = diag ? c : -1.f * static_cast<float>(log(static_cast<double>(c)));
                        if (multi && (diag || dv == 0.f)) use = false;
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
        // ---- 4. merge the warps' lists: block-wide sort of all SC_WARPS * cap keys, the first k are the row
        block_bitonic_sort(sm_keys, SC_WARPS * cap);
        }
        int m = 0;
        for (int t = threadIdx.x; t < k; t += blockDim.x) m += sm_keys[t] != ~0ull ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
        if (lane == 0 && m) atomicAdd(&s_nsel, static_cast<unsigned int>(m));
        __syncthreads();
        const int nsel = static_cast<int>(s_nsel);
        if (warp == 0) {
            double sum = 0.0;
            for (int t = lane; t < nsel; t += 32) sum += static_cast<double>(key_val(sm_keys[t]));
            sum = warp_sum(sum);
            if (lane == 0) s_sum = sum;
        }
        __syncthreads();
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
            P.out_cols[(i - P.r0) * k + t] = oc;
            P.out_vals[(i - P.r0) * k + t] = ov;
        }
        if (threadIdx.x == 0) P.row_nnz[i - P.r0] = nsel;
        __syncthreads();
    }
}

// Rows [r0, r1) of the result (every output row needs the whole matrix and nothing of the other output rows, so slabs of rows
// computed anywhere concatenate to the one-call result: SURVEY 8(e), D1 sharded by output row); outputs hold (r1 - r0) rows.
int create_similarities_rows_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                                 int variant, const int64_t* d_comp_sizes, int64_t r0, int64_t r1, uint32_t* d_out_cols, float* d_out_vals,
                                 int32_t* d_row_nnz, cudaStream_t s) {
    if (r0 < 0 || r1 > n || r0 > r1) return fail(SPHB_ERR_INVALID, "create_similarities: row range outside [0, n]");
    if (r0 == r1) return 0;
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
    SimsParams P{};
    P.n = n; P.r0 = r0; P.r1 = r1; P.srowptr = srowptr.p; P.scol = scol.p; P.sval = sval.p; P.colptr = colptr.p; P.tkeys = tk1.p; P.tval = tval.p;
    P.prune_val = prune_val; P.hdi = hdi ? 1 : 0; P.k = static_cast<int>(k); P.cap = cap;
    P.out_cols = d_out_cols; P.out_vals = d_out_vals; P.row_nnz = d_row_nnz;
    DevBuf<float> acc;
    DevBuf<uint32_t> touched;
    const int cbits = bits_for(static_cast<uint64_t>(n));
    const int CH = static_cast<int>((n + SC_WARPS - 1) / SC_WARPS);
    const size_t class_smem = static_cast<size_t>(SC_WARPS) * cap * sizeof(uint64_t) + static_cast<size_t>(SIMS_EB) * 8 +
                              static_cast<size_t>(SC_WARPS) * CH * sizeof(float);
    if (class_smem <= 220 * 1024 && cbits <= 28 && !g_force_slab) {
        // class-major transpose: (row mod 16, column, row)
        DevBuf<uint64_t> ck0, ck1;
        DevBuf<float> cval;
        DevBuf<int64_t> cptr;
        SPHB_CUDA(ck0.alloc(std::max<int64_t>(snnz, 1), s));
        SPHB_CUDA(ck1.alloc(std::max<int64_t>(snnz, 1), s));
        SPHB_CUDA(cval.alloc(std::max<int64_t>(snnz, 1), s));
        SPHB_CUDA(cptr.alloc(static_cast<size_t>(SC_WARPS) * (n + 1), s));
        DevBuf<uint2> trec;
        SPHB_CUDA(trec.alloc(std::max<int64_t>(snnz, 1), s));
        sims_class_keys<<<g, 256, 0, s>>>(tk1.p, snnz, cbits, ck0.p);
        count_launch();
        // tk1 / tval are (column, row)-sorted already, so a stable sort on the 4 class bits alone finishes the job
        if (int rc = sort_pairs_bits<float>(ck0.p, ck1.p, tval.p, cval.p, snnz, 32 + cbits, 36 + cbits, s)) return rc;
        sims_class_ptr<<<g, 256, 0, s>>>(ck1.p, snnz, n, cbits, cptr.p);
        sims_zip_records<<<g, 256, 0, s>>>(ck1.p, cval.p, snnz, trec.p);
        count_launch(2);
        SPHB_CUDA(cudaFuncSetAttribute(sims_class_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(class_smem)));
        const int64_t grid = std::min<int64_t>(r1 - r0, static_cast<int64_t>(sms) * (class_smem <= 100 * 1024 ? 2 : 1));
        sims_class_kernel<<<static_cast<unsigned>(grid), SC_WARPS * 32, class_smem, s>>>(P, cptr.p, trec.p, CH);
        count_launch();
        SPHB_CUDA(cudaGetLastError());
        SPHB_CUDA(cudaStreamSynchronize(s));  // the class-major buffers are released on return
        return 0;
    }
    {
        const size_t list_bytes = static_cast<size_t>(SIMS_WARPS) * cap * sizeof(uint64_t);
        const size_t bitmap_bytes = static_cast<size_t>((n + 31) / 32) * 4;
        bool bitmap = list_bytes + bitmap_bytes <= 200 * 1024;
#ifdef SPHB_KNN_TUNING
        if (const char* e = getenv("SPHB_SIMS_BITMAP")) bitmap = bitmap && atoi(e) != 0;
#endif
        const size_t slab_smem = list_bytes + (bitmap ? bitmap_bytes : 0);
        if (bitmap) SPHB_CUDA(cudaFuncSetAttribute(sims_row_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(slab_smem)));
        else SPHB_CUDA(cudaFuncSetAttribute(sims_row_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(slab_smem)));
        int64_t grid = std::min<int64_t>(r1 - r0, static_cast<int64_t>(sms) * 2);
        // bound the accumulator slabs to 8 GiB.  (Keeping the slabs L2-resident by running fewer CTAs does not help: 768 x 768 x 128
        // level 1, 296 / 136 / 68 CTAs: 9.2 / 10.4 / 11.6 s — the returning atomics, not L2 misses, were the bound.)
        while (grid > 1 && static_cast<double>(grid) * n * 8.0 > 8.0 * (1ull << 30)) grid /= 2;
#ifdef SPHB_KNN_TUNING
        if (const char* e = getenv("SPHB_SIMS_SLAB_CTAS")) grid = std::max<int64_t>(1, std::min<int64_t>(r1 - r0, atoi(e)));
#endif
        SPHB_CUDA(acc.alloc(static_cast<size_t>(grid) * n, s));
        SPHB_CUDA(touched.alloc(static_cast<size_t>(grid) * n, s));
        SPHB_CUDA(cudaMemsetAsync(acc.p, 0, static_cast<size_t>(grid) * n * sizeof(float), s));
        P.acc = acc.p; P.touched = touched.p;
        if (bitmap) sims_row_kernel<true><<<static_cast<unsigned>(grid), SIMS_WARPS * 32, slab_smem, s>>>(P);
        else sims_row_kernel<false><<<static_cast<unsigned>(grid), SIMS_WARPS * 32, slab_smem, s>>>(P);
    }
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int create_similarities_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, int64_t nnz, int64_t k, float prune_val,
                            int variant, const int64_t* d_comp_sizes, uint32_t* d_out_cols, float* d_out_vals, int32_t* d_row_nnz, cudaStream_t s) {
    return create_similarities_rows_dev(n, d_rowptr, d_col, d_val, nnz, k, prune_val, variant, d_comp_sizes, 0, n, d_out_cols, d_out_vals, d_row_nnz, s);
}

}  // namespace sphb
