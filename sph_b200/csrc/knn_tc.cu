// Exact kNN graph on the 5th-gen tensor cores (Stage A; replaces the faiss call
// sites at sph/utils/Knn.cpp:56-136).
//
// Pipeline
//   1. knn_prep_*: column means (fp64), centre, scale by a power of two, split
//      every value into fp16 hi + lo, and lay out two K-major fp16 operand
//      matrices so that ONE plain GEMM yields  G_ij = |v_j|² - 2 u_i·v_j
//      (u = s(x-mean)):
//         A'[i] = [ hi | hi | lo | 1 1 1 | 0.. ]          (Kp columns)
//         B'[j] = [-2hi |-2lo |-2hi | n0 n1 n2 | 0.. ]     (n0+n1+n2 = |v_j|², fp16 split)
//      i.e. the three cross products hi·hi + hi·lo + lo·hi of a 22-bit mantissa
//      emulation, plus the database norm, inside the contraction.
//   2. knn_sweep_kernel: persistent, warp-specialised CTAs (TMA producer, MMA
//      issuer, 4 epilogue warps).  Each CTA owns a 128-row query tile and sweeps
//      the whole database in 256-row tiles: TMA -> 128B-swizzled smem ->
//      tcgen05.mma (M=128, N=256, fp32 accumulators double-buffered in TMEM) ->
//      tcgen05.ld epilogue.  The epilogue thread that owns a row compares each
//      G against the row's running threshold (one FMNMX per value on the fast
//      path); survivors are appended to a per-row candidate list.  When a list
//      fills up the warp re-ranks it EXACTLY (fp64 FMA chain on the original
//      fp32 data, bit-identical to the oracle), sorts (distance, index) keys,
//      keeps k and tightens the threshold.  Because the filter threshold carries
//      a rigorous bound on the GEMM error, the result is the exact kNN set.
//   3. the final per-row write-out fuses the reference's post-passes (eps clamp,
//      sqrt, self-first).
#include "knn_common.cuh"
#include "ptx_sm100.cuh"

#include <cuda_fp16.h>
#include <cmath>
#include <vector>

namespace sphb {

namespace tc {
constexpr int BM = 128;       // query rows per CTA tile (= TMEM lanes)
constexpr int BN = 256;       // database rows per accumulator stage (= TMEM columns)
constexpr int BK = 64;        // fp16 elements per smem row = 128 B swizzle span
constexpr int UMMA_K = 16;
constexpr int ACC_STAGES = 2;
constexpr int TMEM_COLS = ACC_STAGES * BN;  // 512
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int NUM_THREADS = 192;  // warp0 TMA, warp1 MMA, warps 2..5 epilogue
constexpr double VMAX2 = 16384.0;  // scaled squared-norm ceiling (fits fp16 with headroom)
constexpr double VMAX = 128.0;
}  // namespace tc

struct SweepParams {
    const float* X;          // original fp32 data [n, d]
    const double* un2;       // scaled centred squared norms [n]
    const double* scal;      // [0] = s² (scale squared)
    int64_t n;               // database size
    int d, k, cap, metric, l2_squared;
    int64_t q0, nq;          // query slab
    int n_qtiles, n_dbtiles, kchunks, k16_last, stages;
    double eps_c1, eps_c2;   // filter slack: eps_i = c1*(2|u_i|VMAX + VMAX²) + c2*(|u_i| + VMAX)
    uint64_t* cand;          // [grid, BM, cap] candidate scratch
    int64_t* ids;            // [nq, k]
    float* dist;             // [nq, k]
    unsigned long long* stats;  // [0] compactions, [1] accepted candidates
};

// ---------------------------------------------------------------------------
// prep kernels
// ---------------------------------------------------------------------------
__global__ void knn_prep_colsum(const float* __restrict__ X, int64_t n, int d, double* __restrict__ colsum) {
    // thread <-> column (coalesced across the row), block strides over rows
    const int c = blockIdx.y * blockDim.x + threadIdx.x;
    if (c >= d) return;
    double acc = 0.0;
    for (int64_t i = blockIdx.x; i < n; i += gridDim.x) acc += static_cast<double>(X[i * d + c]);
    atomicAdd(colsum + c, acc);
}

// per row: centred squared norm in fp64 (unscaled); global max via ordered-bits atomicMax
__global__ void knn_prep_norms(const float* __restrict__ X, int64_t n, int d, const double* __restrict__ colsum, int centre,
                               double* __restrict__ un2, unsigned long long* __restrict__ maxbits) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    double wmax = 0.0;
    for (int64_t i = warp; i < n; i += nwarps) {
        double acc = 0.0;
        for (int c = lane; c < d; c += 32) {
            const double mu = centre ? colsum[c] / static_cast<double>(n) : 0.0;
            const double u = static_cast<double>(X[i * d + c]) - mu;
            acc = fma(u, u, acc);
        }
        acc = warp_sum(acc);
        if (lane == 0) un2[i] = acc;
        wmax = fmax(wmax, acc);
    }
    if (lane == 0) atomicMax(maxbits, static_cast<unsigned long long>(__double_as_longlong(wmax)));
}

// scal[0] = s², scal[1] = s; s = 2^floor(log2(sqrt(VMAX2 / max_norm²)))
__global__ void knn_prep_scale(const unsigned long long* __restrict__ maxbits, double* __restrict__ scal) {
    const double mx = __longlong_as_double(static_cast<long long>(*maxbits));
    double s = 1.0;
    if (mx > 0.0 && isfinite(mx)) s = exp2(floor(0.5 * log2(tc::VMAX2 / mx)));
    scal[0] = s * s;
    scal[1] = s;
}

// one warp per row: write A'[i], B'[i]; rescale un2[i] by s².
__global__ void knn_prep_split(const float* __restrict__ X, int64_t n, int64_t np, int d, int kp, const double* __restrict__ colsum,
                               int centre, int with_norm, const double* __restrict__ scal, double* __restrict__ un2,
                               __half* __restrict__ A, __half* __restrict__ B) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const double s = scal[1], s2 = scal[0];
    for (int64_t i = warp; i < np; i += nwarps) {
        __half* a = A + i * kp;
        __half* b = B + i * kp;
        if (i >= n) {  // padding rows: zero queries, "infinitely far" database points
            if (lane < 3) b[3 * d + lane] = __float2half(65504.f);
            continue;
        }
        for (int c = lane; c < d; c += 32) {
            const double mu = centre ? colsum[c] / static_cast<double>(n) : 0.0;
            const double u = s * (static_cast<double>(X[i * d + c]) - mu);
            const __half hi = __double2half(u);
            const __half lo = __double2half(u - static_cast<double>(__half2float(hi)));
            const __half mhi = __float2half(-2.f * __half2float(hi));
            const __half mlo = __float2half(-2.f * __half2float(lo));
            a[c] = hi;         a[d + c] = hi;      a[2 * d + c] = lo;
            b[c] = mhi;        b[d + c] = mlo;     b[2 * d + c] = mhi;
        }
        if (lane == 0) {
            const double nn = un2[i] * s2;
            un2[i] = nn;
            const __half one = __float2half(1.f);
            a[3 * d] = one; a[3 * d + 1] = one; a[3 * d + 2] = one;
            if (with_norm) {
                const __half n0 = __double2half(nn);
                const double r1 = nn - static_cast<double>(__half2float(n0));
                const __half n1 = __double2half(r1);
                const double r2 = r1 - static_cast<double>(__half2float(n1));
                const __half n2 = __double2half(r2);
                b[3 * d] = n0; b[3 * d + 1] = n1; b[3 * d + 2] = n2;
            }
        }
    }
}

// ---------------------------------------------------------------------------
// the sweep
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(tc::NUM_THREADS, 1)
knn_sweep_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB, const SweepParams P) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages * STAGE_BYTES | sort buffers 4*cap*8 | barriers | tmem slot]
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int stages = P.stages;
    uint64_t* sortbuf = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(stages) * STAGE_BYTES);
    uint64_t* bars = sortbuf + 4 * static_cast<size_t>(P.cap);
    // bars: full[stages], empty[stages], tmem_full[2], tmem_empty[2]
    const uint32_t bar_full = ptx::smem_u32(bars);
    const uint32_t bar_empty = ptx::smem_u32(bars + stages);
    const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * stages);
    const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * stages + ACC_STAGES);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 2 * ACC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem_base = ptx::smem_u32(smem);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmapA);
        ptx::prefetch_tmap(&tmapB);
        for (int s = 0; s < stages; ++s) {
            ptx::mbar_init(bar_full + 8 * s, 1);
            ptx::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < ACC_STAGES; ++s) {
            ptx::mbar_init(bar_tfull + 8 * s, 1);
            ptx::mbar_init(bar_tempty + 8 * s, 4);  // one arrive per epilogue warp
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc(ptx::smem_u32(tmem_slot), TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
                const int32_t qrow = static_cast<int32_t>(P.q0 + static_cast<int64_t>(qt) * BM);
                for (int dt = 0; dt < P.n_dbtiles; ++dt) {
                    for (int kc = 0; kc < P.kchunks; ++kc) {
                        ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                        const uint32_t sa = smem_base + stage * STAGE_BYTES;
                        ptx::mbar_expect_tx(bar_full + 8 * stage, STAGE_BYTES);
                        ptx::tma_load_2d(sa, &tmapA, bar_full + 8 * stage, kc * BK, qrow);
                        ptx::tma_load_2d(sa + A_BYTES, &tmapB, bar_full + 8 * stage, kc * BK, dt * BN);
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_f16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t aphase = 0;
            for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
                for (int dt = 0; dt < P.n_dbtiles; ++dt) {
                    ptx::mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t tmem_d = tmem_base + as * BN;
                    for (int kc = 0; kc < P.kchunks; ++kc) {
                        ptx::mbar_wait(bar_full + 8 * stage, phase);
                        ptx::tc_fence_after();
                        const uint32_t sa = smem_base + stage * STAGE_BYTES;
                        const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
                        const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + A_BYTES);
                        const int nk = (kc == P.kchunks - 1) ? P.k16_last : (BK / UMMA_K);
                        for (int kk = 0; kk < nk; ++kk) {
                            // advance 16 fp16 = 32 B inside the 128 B swizzle span: +2 in the (addr >> 4) field
                            ptx::umma_f16(tmem_d, adesc + 2 * kk, bdesc + 2 * kk, idesc, (kc | kk) != 0);
                        }
                        ptx::umma_commit(bar_empty + 8 * stage);  // frees the smem slot when these MMAs retire
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                    ptx::umma_commit(bar_tfull + 8 * as);  // accumulator ready for the epilogue
                    if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
                }
            }
        }
    } else {
        // ===================== epilogue: filter + exact top-k =====================
        const int quarter = warp & 3;  // TMEM lane quarter this warp may access
        uint64_t* sm = sortbuf + static_cast<size_t>(warp - 2) * P.cap;
        const int cap = P.cap, k = P.k;
        const double s2 = P.scal[0];
        const double scaleG = (P.metric == METRIC_L2) ? s2 : 2.0 * s2;
        int as = 0;
        uint32_t aphase = 0;
        unsigned long long n_compact = 0, n_accept = 0;

        for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
            const int row_in_tile = quarter * 32 + lane;
            const int64_t r = static_cast<int64_t>(qt) * BM + row_in_tile;  // row within the slab
            const bool valid = r < P.nq;
            const int64_t gi = P.q0 + r;  // global point id of my row
            uint64_t* mybuf = P.cand + (static_cast<size_t>(blockIdx.x) * BM + row_in_tile) * cap;
            const double my_un2 = valid ? P.un2[gi] : 0.0;
            const double my_off = (P.metric == METRIC_L2) ? my_un2 : 0.0;
            const double my_eps = P.eps_c1 * (2.0 * sqrt(my_un2) * VMAX + VMAX2) + P.eps_c2 * (sqrt(my_un2) + VMAX);
            float thr = valid ? INFINITY : -INFINITY;
            int cnt = 0;  // entries in mybuf: [0,nk) exact sorted keys, [nk,cnt) new candidate indices
            int nk = 0;

            // warp-cooperative exact re-rank + sort + truncate of lane L's row
            auto compact_row = [&](int L) {
                const int cntL = __shfl_sync(0xffffffffu, cnt, L);
                const int nkL = __shfl_sync(0xffffffffu, nk, L);
                const int64_t giL = P.q0 + static_cast<int64_t>(qt) * BM + quarter * 32 + L;
                uint64_t* buf = P.cand + (static_cast<size_t>(blockIdx.x) * BM + quarter * 32 + L) * cap;
                for (int t = lane; t < cap; t += 32) {
                    uint64_t key = ~0ull;
                    if (t < nkL) key = buf[t];
                    else if (t < cntL) {
                        const uint32_t j = static_cast<uint32_t>(buf[t]);
                        key = make_key(exact_key(P.X, giL, static_cast<int64_t>(j), P.d, P.metric), j);
                    }
                    sm[t] = key;
                }
                __syncwarp();
                warp_bitonic_sort(sm, cap, lane);
                const int newnk = min(k, cntL);
                for (int t = lane; t < newnk; t += 32) buf[t] = sm[t];
                if (lane == L) {
                    cnt = newnk;
                    nk = newnk;
                    if (newnk == k) {
                        const double kv = static_cast<double>(key_val(sm[k - 1]));
                        const double g = scaleG * kv;
                        const double t = g + fabs(g) * 2.4e-7 - my_off + my_eps + (fabs(g) + my_off) * 2.4e-7;
                        thr = __double2float_ru(t);
                    }
                }
                __syncwarp();
            };

            for (int dt = 0; dt < P.n_dbtiles; ++dt) {
                ptx::mbar_wait(bar_tfull + 8 * as, aphase);
                ptx::tc_fence_after();
                const uint32_t jbase = static_cast<uint32_t>(dt) * BN;
#pragma unroll 1
                for (int c = 0; c < BN / 32; ++c) {
                    float v[32];
                    ptx::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN + c * 32, v);
                    float m = v[0];
#pragma unroll
                    for (int t = 1; t < 32; ++t) m = fminf(m, v[t]);
                    if (m <= thr) {
#pragma unroll
                        for (int t = 0; t < 32; ++t) {
                            if (v[t] <= thr) {
                                const uint32_t j = jbase + c * 32 + t;
                                if (j < P.n) {
                                    mybuf[cnt++] = j;
                                    ++n_accept;
                                }
                            }
                        }
                    }
                    unsigned full = __ballot_sync(0xffffffffu, cnt > cap - 32);
                    while (full) {
                        const int L = __ffs(full) - 1;
                        full &= full - 1;
                        compact_row(L);
                        ++n_compact;
                    }
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
                if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
            }

            // finalise the 32 rows of this warp
            for (int L = 0; L < 32; ++L) {
                const int64_t rL = static_cast<int64_t>(qt) * BM + quarter * 32 + L;
                if (rL >= P.nq) break;
                compact_row(L);
                const int nkL = __shfl_sync(0xffffffffu, nk, L);
                write_row(sm, nkL, k, P.q0 + rL, P.metric, P.l2_squared, P.ids + rL * k, P.dist + rL * k, lane);
                __syncwarp();
            }
        }
        if (lane == 0 && P.stats) {
            atomicAdd(P.stats + 0, n_compact);
        }
        n_accept = __reduce_add_sync(0xffffffffu, static_cast<unsigned>(n_accept & 0xffffffffu));
        if (lane == 0 && P.stats) atomicAdd(P.stats + 1, n_accept);
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn get_encode_fn() {
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeFn>(p);
    }
    return fn;
}

static int make_tmap(CUtensorMap* tm, const void* base, int64_t rows, int kp, int box_rows) {
    EncodeFn enc = get_encode_fn();
    if (!enc) return fail(3, "cuTensorMapEncodeTiled entry point not found");
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(kp), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(kp) * 2};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(tc::BK), static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(3, "cuTensorMapEncodeTiled failed: " + std::to_string(static_cast<int>(r)));
    return 0;
}

static int next_pow2_i(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

static int64_t g_knn_stats[6] = {0, 0, 0, 0, 0, 0};
void knn_set_stats(const int64_t* s) { for (int i = 0; i < 6; ++i) g_knn_stats[i] = s[i]; }
void knn_get_stats(int64_t* s) { for (int i = 0; i < 6; ++i) s[i] = g_knn_stats[i]; }

bool knn_tc_supported(int64_t n, int d, int k) {
    if (d < 8 || n < 1024) return false;
    if (n >= (1ll << 31)) return false;
    if (3 * d + 3 > 65536) return false;
    if (next_pow2_i(k + 96) > 1024) return false;
    return true;
}

int knn_tc_launch(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1, int64_t* d_ids,
                  float* d_dist, cudaStream_t stream) {
    using namespace tc;
    const int64_t nq = q1 - q0;
    if (nq <= 0) return 0;
    const int centre = metric == METRIC_L2 ? 1 : 0;
    const int kp = ((3 * d + 3) + 15) / 16 * 16;
    const int64_t np = (n + BN - 1) / BN * BN;
    const int cap = next_pow2_i(k + 96);
    const int kchunks = (kp + BK - 1) / BK;
    const int k16_last = (kp - (kchunks - 1) * BK) / UMMA_K;
    const int n_qtiles = static_cast<int>((nq + BM - 1) / BM);
    const int n_dbtiles = static_cast<int>(np / BN);
    const int grid = std::min(ctx().num_sms, n_qtiles);

    DevBuf<double> colsum, un2, scal;
    DevBuf<unsigned long long> maxbits, stats;
    DevBuf<__half> A, B;
    DevBuf<uint64_t> cand;
    SPHB_CUDA(colsum.alloc(d, stream));
    SPHB_CUDA(un2.alloc(n, stream));
    SPHB_CUDA(scal.alloc(2, stream));
    SPHB_CUDA(maxbits.alloc(1, stream));
    SPHB_CUDA(stats.alloc(2, stream));
    SPHB_CUDA(A.alloc(static_cast<size_t>(np) * kp, stream));
    SPHB_CUDA(B.alloc(static_cast<size_t>(np) * kp, stream));
    SPHB_CUDA(cand.alloc(static_cast<size_t>(grid) * BM * cap, stream));
    SPHB_CUDA(cudaMemsetAsync(colsum.p, 0, d * sizeof(double), stream));
    SPHB_CUDA(cudaMemsetAsync(maxbits.p, 0, sizeof(unsigned long long), stream));
    SPHB_CUDA(cudaMemsetAsync(stats.p, 0, 2 * sizeof(unsigned long long), stream));
    SPHB_CUDA(cudaMemsetAsync(A.p, 0, static_cast<size_t>(np) * kp * sizeof(__half), stream));
    SPHB_CUDA(cudaMemsetAsync(B.p, 0, static_cast<size_t>(np) * kp * sizeof(__half), stream));

    const int sms = ctx().num_sms;
    if (centre) {
        dim3 g(static_cast<unsigned>(std::min<int64_t>(n, sms * 8)), static_cast<unsigned>((d + 127) / 128));
        knn_prep_colsum<<<g, 128, 0, stream>>>(d_x, n, d, colsum.p);
        count_launch();
    }
    knn_prep_norms<<<sms * 8, 256, 0, stream>>>(d_x, n, d, colsum.p, centre, un2.p, maxbits.p);
    knn_prep_scale<<<1, 1, 0, stream>>>(maxbits.p, scal.p);
    knn_prep_split<<<sms * 8, 256, 0, stream>>>(d_x, n, np, d, kp, colsum.p, centre, metric == METRIC_L2 ? 1 : 0, scal.p, un2.p, A.p, B.p);
    count_launch(3);
    SPHB_CUDA(cudaGetLastError());

    CUtensorMap tmA, tmB;
    if (int rc = make_tmap(&tmA, A.p, np, kp, BM)) return rc;
    if (int rc = make_tmap(&tmB, B.p, np, kp, BN)) return rc;

    // shared memory budget: pipeline stages + sort buffers + barriers
    const size_t fixed = 4 * static_cast<size_t>(cap) * 8 + 256 + 1024;
    int stages = 4;
    while (stages > 2 && static_cast<size_t>(stages) * STAGE_BYTES + fixed > 227 * 1024) --stages;
    const size_t smem = static_cast<size_t>(stages) * STAGE_BYTES + fixed;
    SPHB_CUDA(cudaFuncSetAttribute(knn_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));

    SweepParams P{};
    P.X = d_x; P.un2 = un2.p; P.scal = scal.p; P.n = n; P.d = d; P.k = k; P.cap = cap; P.metric = metric; P.l2_squared = l2_squared;
    P.q0 = q0; P.nq = nq; P.n_qtiles = n_qtiles; P.n_dbtiles = n_dbtiles; P.kchunks = kchunks; P.k16_last = k16_last; P.stages = stages;
    // GEMM error bound: split truncation 6·2^-22 + accumulation (kp/16 steps)·2^-22,
    // times the magnitude (2|u||v| + |v|²); plus the fp16 subnormal floor; safety factor 2.
    const double nsteps = static_cast<double>(kp / UMMA_K);
    P.eps_c1 = 2.0 * (6.0 + nsteps) * std::ldexp(1.0, -22);
    P.eps_c2 = 2.0 * std::ldexp(1.0, -23) * std::sqrt(static_cast<double>(d));
    P.cand = cand.p; P.ids = d_ids; P.dist = d_dist; P.stats = stats.p;

    knn_sweep_kernel<<<grid, NUM_THREADS, smem, stream>>>(tmA, tmB, P);
    count_launch();
    SPHB_CUDA(cudaGetLastError());

    unsigned long long hstats[2] = {0, 0};
    SPHB_CUDA(cudaMemcpyAsync(hstats, stats.p, sizeof(hstats), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaStreamSynchronize(stream));
    const int64_t st[6] = {1, static_cast<int64_t>(hstats[0]), static_cast<int64_t>(hstats[1]), kp, n_dbtiles, n_qtiles};
    knn_set_stats(st);
    return 0;
}

}  // namespace sphb
