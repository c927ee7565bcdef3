// Exact kNN graph on the 5th-gen tensor cores (Stage A; replaces the faiss call
// sites at sph/utils/Knn.cpp:56-136).
//
// 1. knn_prep_*: column means (fp64), centre, scale by a power of two, split every
//    value into fp16 hi + lo (22-bit mantissa emulation) and build fp16 operand
//    matrices such that a GEMM yields  G_ij = |v_j|² - 2 u_i·v_j  (u = s(x-mean)):
//    three cross products hi·hi + hi·lo + lo·hi plus the database norm (itself a
//    3-term fp16 split multiplied by a column of ones) inside the contraction.
// 2. the sweep: persistent, warp-specialised CTAs (TMA producer warp, MMA issuer
//    warp, 4 or 8 epilogue warps).  Two flavours:
//      knn_sweep_resident<MH>  d <= 256: the query tile (MH x 128 rows, hi|lo) stays
//                              RESIDENT in shared memory for the whole database
//                              sweep; only database chunks stream through a TMA
//                              ring, so L2->SM traffic is ~(2d+64)*2 B per database
//                              row per CTA and the kernel is tensor-pipe bound;
//      knn_sweep_stream        any d: both operands stream in K chunks (K-concatenated
//                              layout [hi|hi|lo|1] x [-2hi|-2lo|-2hi|n]).
//    tcgen05.mma (M=128, fp32 accumulators double-buffered in TMEM) -> tcgen05.ld
//    epilogue -> threshold filter + per-row candidate lists (knn_epilogue.cuh).
// 3. knn_rerank_kernel: exact fp64 re-rank of each row's candidates (bit-identical
//    to the oracle), (distance, index) sort, fused post-passes, single write of
//    ids/dist.  Rows flagged as overflowing go through the exact SIMT kernel.
#include "knn_epilogue.cuh"

#include <cuda_fp16.h>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <vector>

namespace sphb {

namespace tc {
constexpr int BM = 128;       // query rows per MMA (= TMEM lanes)
constexpr int BK = 64;        // fp16 elements per smem row = 128 B swizzle span
constexpr int UMMA_K = 16;
constexpr int CHUNK_ROWS = 128;
constexpr int CHUNK_BYTES = CHUNK_ROWS * BK * 2;  // 16 KB: one 128-row x 64-col fp16 tile
constexpr double VMAX2 = 16384.0;  // scaled squared-norm ceiling (fits fp16 with headroom)
constexpr double VMAX = 128.0;
}  // namespace tc

struct SweepParams {
    const double* un2;     // scaled centred squared norms, indexed by GLOBAL row id [n]
    const float* dun;      // |u_i - fp16(u_i)| (scaled), by global row id [n]
    const float* dvmax;    // device scalar: max_i dun[i]
    const int64_t* qrows;  // optional explicit list of query rows [nq] (re-sweep of a subset); null: row r <-> q0 + r
    int64_t n;             // database size
    int k, cap, keep_max, metric;
    int64_t q0, nq;        // query slab (q0 = first global row when qrows is null)
    int64_t a_row0;        // row of the A operand matrix that holds query 0 of this launch
    int n_qtiles, n_dbtiles;
    int nseg, seg_tiles;   // resident flavour: database segments per query tile (0 / 1: none) and tiles per segment, see sweep_item
    int dp64;              // resident flavour: padded d / 64
    int passes;            // 1: single fp16 product, 3: split fp16 (hi·hi + lo·hi + hi·lo)
    int half_major;        // resident lean issue loop: all chunks for M half 0, then for half 1 (needs stages >= 2 chunks-per-tile)
    int norm_bytes;        // resident flavour: bytes of the norm chunk of one database tile (16 KB wide / 4 KB narrow)
    int stagger;           // database-tile start offset per CTA (desynchronises the CTAs' L2 reads)
    int lead;              // the sweep of a wave of query tiles starts `lead` database tiles before the wave's own rows
    int kchunks, k16_last, stages;  // stream flavour: K chunks per tile, k-steps in the last one; both: ring depth
    double eps_c1, eps_c2; // passes = 3 error model
    float acc_c, sub_c;    // passes = 1 error model (knn_epilogue.cuh BandConsts)
    uint64_t* cand;        // [grid, rows_per_cta, cap] scratch lists
    uint32_t* cand_out;    // [nq, keep_max]
    int* cnt_out;          // [nq]
    unsigned long long* stats;  // [0] compactions, [1] accepted candidates
    unsigned long long* progress;  // [1] zeroed: sum over CTAs of database tiles loaded (soft lock-step of the sweep, see sweep_pace)
    int pace_window;            // a CTA may run at most this many tiles ahead of the average (0: pacing off)
    int l2_prefetch;            // producer prefetches the database tile this many steps ahead into L2 (0: off)
    int compact_step;           // BandConsts::step
    int debug_skip_epilogue;    // debug: epilogue releases the accumulator without reading it (pure MMA rate)
    unsigned long long* prof;   // optional [grid][16] cycle counters (build -DSPHB_KNN_PROFILE, run SPHB_KNN_PROF=1)
};


__device__ __forceinline__ BandConsts make_band_consts(const SweepParams& P) {
    BandConsts bc;
    bc.passes = P.passes;
    bc.l2 = P.metric == METRIC_L2 ? 1 : 0;
    bc.vmax = static_cast<float>(tc::VMAX) * 1.001f;
    bc.dvmax = P.passes == 1 ? __ldg(P.dvmax) : 0.f;
    bc.acc_c = P.acc_c;
    bc.sub_c = P.sub_c;
    bc.step = P.compact_step;
    return bc;
}

// Soft lock-step of the CTAs' database sweeps.  Every CTA streams the whole database once per query tile; the CTAs of a
// wave start at (nearly) the same position, but without coupling they drift apart by more than the L2 holds and each
// then streams from DRAM on its own (measured: 18 TB of DRAM reads per C4 sweep launch, 92 % L2 misses, 3.1 TB/s of
// HBM traffic on a power-capped chip).  The TMA producer publishes its progress (tiles loaded since launch) in a global
// sum every 16 tiles and does not run more than pace_window tiles ahead of the average; the slowest CTA never waits,
// so there is no deadlock (all CTAs are resident: grid <= SM count, one CTA per SM).  A CTA that has finished
// publishes the progress of a complete run.
struct SweepPace {
    long long mine = 0, published = 0;
    bool off = false;  // this CTA gave up waiting (no global progress for ~5 ms: some CTA of the grid is not running)
};
__device__ __forceinline__ void sweep_pace(const SweepParams& P, SweepPace& sp, int lane);
// the same for a producer that is a single thread (streaming flavour)
__device__ __forceinline__ void sweep_pace_one(const SweepParams& P, SweepPace& sp) {
    if (P.pace_window > 0) {
        if (sp.mine > sp.published) {
            atomicAdd(P.progress, static_cast<unsigned long long>(sp.mine - sp.published));
            sp.published = sp.mine;
        }
        const long long g = gridDim.x;
        unsigned long long last = ~0ull;
        int idle = 0;
        while (!sp.off) {
            const unsigned long long tot = *reinterpret_cast<volatile unsigned long long*>(P.progress);
            if (sp.mine * g <= static_cast<long long>(tot) + static_cast<long long>(P.pace_window) * g) break;
            if (tot != last) { last = tot; idle = 0; }
            else if (++idle > 20000) sp.off = true;
            __nanosleep(256);
        }
    }
}
__device__ __forceinline__ void sweep_pace(const SweepParams& P, SweepPace& sp, int lane) {
    if (P.pace_window > 0 && lane == 0) {
        if (sp.mine > sp.published) {
            atomicAdd(P.progress, static_cast<unsigned long long>(sp.mine - sp.published));
            sp.published = sp.mine;
        }
        const long long g = gridDim.x;
        // The wait is bounded: the host only enables pacing when the whole grid fits the device (occupancy query), but a
        // CTA that is not scheduled all the same (MPS limit, another long kernel on the SMs) must not hang the others —
        // after ~5 ms without ANY global progress this CTA stops pacing itself for the rest of the launch.
        unsigned long long last = ~0ull;
        int idle = 0;
        while (!sp.off) {
            const unsigned long long tot = *reinterpret_cast<volatile unsigned long long*>(P.progress);
            if (sp.mine * g <= static_cast<long long>(tot) + static_cast<long long>(P.pace_window) * g) break;
            if (tot != last) { last = tot; idle = 0; }
            else if (++idle > 20000) sp.off = true;
            __nanosleep(256);
        }
    }
    __syncwarp();
}
__device__ __forceinline__ void sweep_pace_finish(const SweepParams& P, SweepPace& sp, int lane, int n_items) {
    if (P.pace_window > 0 && lane == 0) {
        const long long waves = (n_items + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
        const long long all = waves * P.n_dbtiles;
        if (all > sp.published) atomicAdd(P.progress, static_cast<unsigned long long>(all - sp.published));
    }
}

// Per-row state at the start of a query tile.
__device__ __forceinline__ void init_row(RowList& me, const SweepParams& P, int64_t r, bool valid) {
    const int64_t gi = valid ? (P.qrows ? P.qrows[r] : P.q0 + r) : 0;
    const double un2 = valid ? P.un2[gi] : 0.0;
    me.thr = valid ? INFINITY : -INFINITY;
    me.eps2 = row_eps2(un2, P.eps_c1, P.eps_c2, tc::VMAX, tc::VMAX2);
    me.un = __double2float_ru(sqrt(un2)) * 1.0001f;
    me.dun = (valid && P.passes == 1) ? P.dun[gi] : 0.f;
    me.cnt = 0;
    me.ovf = 0;
    me.trig = P.cap - 64;  // the first compaction happens on a full list (no threshold before it)
}

// First database tile of the sweep of query tile qt.  All CTAs of a wave (query tiles qt - blockIdx.x ...)
// start just before the wave's own rows — pixels with nearby indices are the likeliest neighbours, so the
// rows' thresholds tighten within the first few percent of the sweep and the filter's slow path becomes
// rare — each shifted by blockIdx.x * stagger tiles so that the CTAs read different L2 lines at any moment
// while the union of their positions stays inside an L2-sized window.
// r0: first query row of the wave, slot: position of the CTA (or CTA pair) inside the wave
__device__ __forceinline__ int sweep_start_at(const SweepParams& P, int64_t r0, int slot, int bn) {
    const int64_t anchor = P.qrows ? P.qrows[r0 < P.nq ? r0 : P.nq - 1] : P.q0 + r0;
    int64_t t = (anchor / bn - P.lead + static_cast<int64_t>(slot) * P.stagger) % P.n_dbtiles;
    if (t < 0) t += P.n_dbtiles;
    return static_cast<int>(t);
}
__device__ __forceinline__ int sweep_start(const SweepParams& P, int qt, int rows_per_cta, int bn) {
    return sweep_start_at(P, static_cast<int64_t>(qt - static_cast<int>(blockIdx.x)) * rows_per_cta, static_cast<int>(blockIdx.x), bn);
}

// ---------------------------------------------------------------------------
// prep kernels
// ---------------------------------------------------------------------------
__global__ void knn_prep_colsum(const float* __restrict__ X, int64_t n, int d, double* __restrict__ colsum) {
    const int c = blockIdx.y * blockDim.x + threadIdx.x;  // thread <-> column (coalesced), block strides over rows
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

__global__ void knn_prep_rescale(double* __restrict__ un2, int64_t n, const double* __restrict__ scal) {
    const double s2 = scal[0];
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * blockDim.x)
        un2[i] *= s2;
}

// Builds the fp16 operand rows, one warp per output row.  Output row r holds point i = rowlist ? rowlist[r] : r.
//   layout 0 (stream, K-concatenated), passes 3:  A = [hi|hi|lo|1 1 1|0..]   B = [-2hi|-2lo|-2hi|n0 n1 n2|0..]
//                                      passes 1:  A = [hi|1 1 1|0..]         B = [-2hi|n0 n1 n2|0..]
//   layout 1 (resident),               passes 3:  A = [hi(dp)|lo(dp)]        B = [-2hi(dp)|-2lo(dp)]    Nrm = [n0 n1 n2 0..]
//                                      passes 1:  A = [hi(dp)]               B = [-2hi(dp)]             Nrm = [n0 n1 n2 0..]
// (n0 n1 n2) = 3-term fp16 split of the scaled squared norm un2[i] (already rescaled).  A, B, Nrm may each be null.
// Rows r >= nvalid are padding: zero queries, "infinitely far" database points.  dun (optional) receives
// |u_i - hi_i| rounded up and dvmax its maximum (float bits, non-negative -> unsigned atomicMax).
__global__ void knn_prep_split(const float* __restrict__ X, int64_t n, int64_t nvalid, int64_t nrows, const int64_t* __restrict__ rowlist, int d,
                               int layout, int passes, int lda, int ldb, int ldn, int dp, const double* __restrict__ colsum, int centre,
                               int with_norm, const double* __restrict__ scal, const double* __restrict__ un2, __half* __restrict__ A,
                               __half* __restrict__ B, __half* __restrict__ Nrm, float* __restrict__ dun, unsigned int* __restrict__ dvmax) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    const double s = scal[1];
    const int ncol = layout == 0 ? (passes == 3 ? 3 * d : d) : 0;  // column of the norm / ones triple (layout 0)
    float wmax = 0.f;
    for (int64_t r = warp; r < nrows; r += nwarps) {
        __half* a = A ? A + r * lda : nullptr;
        __half* b = B ? B + r * ldb : nullptr;
        __half* nr = Nrm ? Nrm + r * ldn : (b && layout == 0 ? b + ncol : nullptr);
        if (r >= nvalid) {
            if (nr && lane < 3) nr[lane] = __float2half(65504.f);
            continue;
        }
        const int64_t i = rowlist ? rowlist[r] : r;
        double err2 = 0.0;
        for (int c = lane; c < d; c += 32) {
            const double mu = centre ? colsum[c] / static_cast<double>(n) : 0.0;
            const double u = s * (static_cast<double>(X[i * d + c]) - mu);
            const __half hi = __double2half(u);
            const double rem = u - static_cast<double>(__half2float(hi));
            const __half lo = __double2half(rem);
            err2 = fma(rem, rem, err2);
            const __half mhi = __float2half(-2.f * __half2float(hi));
            const __half mlo = __float2half(-2.f * __half2float(lo));
            if (layout == 0) {
                if (passes == 3) {
                    if (a) { a[c] = hi;  a[d + c] = hi;  a[2 * d + c] = lo; }
                    if (b) { b[c] = mhi; b[d + c] = mlo; b[2 * d + c] = mhi; }
                } else {
                    if (a) a[c] = hi;
                    if (b) b[c] = mhi;
                }
            } else {
                if (a) { a[c] = hi; if (passes == 3) a[dp + c] = lo; }
                if (b) { b[c] = mhi; if (passes == 3) b[dp + c] = mlo; }
            }
        }
        err2 = warp_sum(err2);
        if (lane == 0) {
            if (dun) {
                const float e = __double2float_ru(sqrt(err2)) * 1.0001f;
                dun[i] = e;
                wmax = fmaxf(wmax, e);
            }
            if (layout == 0 && a) {
                const __half one = __float2half(1.f);
                a[ncol] = one; a[ncol + 1] = one; a[ncol + 2] = one;
            }
            if (with_norm && nr) {
                const double nn = un2[i];
                const __half n0 = __double2half(nn);
                const double r1 = nn - static_cast<double>(__half2float(n0));
                const __half n1 = __double2half(r1);
                const double r2 = r1 - static_cast<double>(__half2float(n1));
                const __half n2 = __double2half(r2);
                nr[0] = n0; nr[1] = n1; nr[2] = n2;
            }
        }
    }
    if (dvmax && lane == 0 && wmax > 0.f) atomicMax(dvmax, __float_as_uint(wmax));
}

__global__ void knn_prep_ones(__half* __restrict__ ones) {  // [128, 64]: rows [1 1 1 0 ...]
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 128 * 64) ones[i] = __float2half((i & 63) < 3 ? 1.f : 0.f);
}

// ---------------------------------------------------------------------------
// sweep, resident-query flavour (1 pass: d <= 512, 3 passes: d <= 256)
// ---------------------------------------------------------------------------
// smem: [A: MH x nA chunks | ONES chunk | B ring: stages chunks | barriers | tmem slot]
// B chunk sequence per 128-row database tile: dp64 hi chunks, (3 passes: dp64 lo chunks,) 1 norm chunk.
// The norm chunk is a separate [np, 64] (wide, SWIZZLE_128B) or [np, 16] (narrow, SWIZZLE_32B) matrix;
// only its first 16 columns (one k-step against the ONES tile) are used.
// DP64T > 0: 1-pass specialisation with the chunk count known at compile time (the MMA issue loop is
// straight-line code with pre-built descriptors; every hi chunk runs all 4 k-steps, the zero padding of
// the operands makes that exact); DP64T = 0: generic (runtime chunk count, 1 or 3 passes, partial last chunk).
//
// Work items.  A launch processes n_qtiles * nseg items; item w = (query tile qt, database segment seg).
// nseg = 1 (every main sweep): an item is one query tile against the whole database, rotated to start near the
// tile's own rows.  nseg > 1 (the re-sweep of the few rows whose 1-pass band overflowed): the database is cut into
// nseg segments of seg_tiles tiles, so that a handful of query tiles still fills all SMs — each (tile, segment)
// keeps its own candidate lists (every true neighbour is among the k nearest of ITS segment, so the per-segment
// band argument holds unchanged) and the re-rank merges the nseg lists of a row.  Round 1 ran such a re-sweep as one
// database pass per query tile: 65 ms for 0.8 % of the rows at C4, however few rows a rank had.
struct SweepItem {
    int qt, seg, tile0, ntiles;
};
__device__ __forceinline__ SweepItem sweep_item(const SweepParams& P, int w) {
    SweepItem it;
    if (P.nseg <= 1) {
        it.qt = w; it.seg = 0; it.tile0 = 0; it.ntiles = P.n_dbtiles;
        return it;
    }
    it.seg = w / P.n_qtiles;  // segment-major: the CTAs of a wave read the same part of the database
    it.qt = w - it.seg * P.n_qtiles;
    it.tile0 = it.seg * P.seg_tiles;
    it.ntiles = min(P.seg_tiles, P.n_dbtiles - it.tile0);
    return it;
}
__device__ __forceinline__ int sweep_item_start(const SweepParams& P, const SweepItem& it, int rows_per_cta, int bn) {
    return P.nseg <= 1 ? sweep_start(P, it.qt, rows_per_cta, bn) : 0;
}
// database tile of step t of an item (dt0 from sweep_item_start)
__device__ __forceinline__ int sweep_tile(const SweepParams& P, const SweepItem& it, int dt0, int t) {
    if (P.nseg > 1) return it.tile0 + t;
    return dt0 + t < P.n_dbtiles ? dt0 + t : dt0 + t - P.n_dbtiles;
}

template <int MH, int R, int DP64T>
__global__ void __launch_bounds__(64 + 128 * MH, 1)
knn_sweep_resident(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                   const __grid_constant__ CUtensorMap tmapN, const __grid_constant__ CUtensorMap tmapO, const SweepParams P,
                   const int k16_tail) {
    using namespace tc;
    constexpr int BN = 128;
    constexpr int ACC_COLS = MH * BN;          // columns per accumulator stage
    constexpr int TMEM_COLS = 2 * ACC_COLS;    // double buffered
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int stages = P.stages, dp64 = P.dp64;
    const bool three = P.passes == 3;
    const int nA = three ? 2 * dp64 : dp64;  // A chunks per M half
    const int nB = nA + 1;                   // B chunks per database tile (the last one is the norm chunk)
    const int n_items = P.n_qtiles * (P.nseg > 1 ? P.nseg : 1);
    const uint32_t smem_base = ptx::smem_u32(smem);
    const uint32_t sA = smem_base;                                   // [m][c] -> sA + (m*nA + c)*CHUNK_BYTES
    const uint32_t sO = sA + MH * nA * CHUNK_BYTES;
    const uint32_t sB = sO + CHUNK_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(MH * nA + 1 + stages) * CHUNK_BYTES);
    const uint32_t bar_full = ptx::smem_u32(bars);
    const uint32_t bar_empty = ptx::smem_u32(bars + stages);
    // accumulator barriers are per (stage, M half): index as * MH + m; the 4 epilogue warps of a half release it
    const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * stages);
    const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * stages + 4);
    const uint32_t bar_afull = ptx::smem_u32(bars + 2 * stages + 8);
    const uint32_t bar_aempty = ptx::smem_u32(bars + 2 * stages + 9);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 10);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // warps: 0 TMA producer, 1 MMA issuer, 2.. epilogue.  (A second issuer warp, one per M half, was measured: no gain —
    // the MMA side is bound by the shared-memory operand bandwidth of M=128 x N=128 tiles, not by issue.)

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmapA);
        ptx::prefetch_tmap(&tmapB);
        ptx::prefetch_tmap(&tmapN);
        ptx::prefetch_tmap(&tmapO);
        for (int s = 0; s < stages; ++s) {
            ptx::mbar_init(bar_full + 8 * s, 1);
            ptx::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < 2 * MH; ++s) {
            ptx::mbar_init(bar_tfull + 8 * s, 1);
            ptx::mbar_init(bar_tempty + 8 * s, 4);
        }
        ptx::mbar_init(bar_afull, 1);
        ptx::mbar_init(bar_aempty, 1);
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
        // ===================== TMA producer (warp-uniform control flow, one elected lane issues) =====================
        int stage = 0;
        uint32_t phase = 0, aphase = 0;
        bool first = true;
        long long pw = 0, pt0 = SPHB_CLOCK();
        SweepPace pace;
        for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
            const SweepItem it = sweep_item(P, w);
            ptx::mbar_wait(bar_aempty, aphase ^ 1);  // previous query tile fully consumed
            if (ptx::elect_one_sync()) {
                ptx::mbar_expect_tx(bar_afull, (MH * nA + (first ? 1 : 0)) * CHUNK_BYTES);
                for (int m = 0; m < MH; ++m) {
                    const int32_t qrow = static_cast<int32_t>(P.a_row0 + (static_cast<int64_t>(it.qt) * MH + m) * BM);
                    for (int c = 0; c < nA; ++c) ptx::tma_load_2d(sA + (m * nA + c) * CHUNK_BYTES, &tmapA, bar_afull, c * BK, qrow);
                }
                if (first) ptx::tma_load_2d(sO, &tmapO, bar_afull, 0, 0);
            }
            __syncwarp();
            first = false;
            aphase ^= 1;
            const int dt0 = sweep_item_start(P, it, MH * BM, BN);
            for (int t = 0; t < it.ntiles; ++t) {
                const int dt = sweep_tile(P, it, dt0, t);
                if ((t & 15) == 0) sweep_pace(P, pace, lane);
                ++pace.mine;
                if (P.l2_prefetch > 0 && t + P.l2_prefetch < it.ntiles) {
                    // pull the tile this CTA needs l2_prefetch steps from now into L2
                    const int pd = sweep_tile(P, it, dt0, t + P.l2_prefetch);
                    if (ptx::elect_one_sync()) {
                        for (int c = 0; c < nB - 1; ++c) ptx::tma_prefetch_2d(&tmapB, c * BK, pd * BN);
                        ptx::tma_prefetch_2d(&tmapN, 0, pd * BN);
                    }
                    __syncwarp();
                }
                for (int c = 0; c < nB; ++c) {
                    const long long w0 = SPHB_CLOCK();
                    ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    pw += SPHB_CLOCK() - w0;
                    if (ptx::elect_one_sync()) {
                        if (c < nB - 1) {
                            ptx::mbar_expect_tx(bar_full + 8 * stage, CHUNK_BYTES);
                            ptx::tma_load_2d(sB + stage * CHUNK_BYTES, &tmapB, bar_full + 8 * stage, c * BK, dt * BN);
                        } else {
                            ptx::mbar_expect_tx(bar_full + 8 * stage, P.norm_bytes);
                            ptx::tma_load_2d(sB + stage * CHUNK_BYTES, &tmapN, bar_full + 8 * stage, 0, dt * BN);
                        }
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
        }
        sweep_pace_finish(P, pace, lane, P.n_qtiles);
        if (P.prof && lane == 0) {
            P.prof[blockIdx.x * 16 + 0] = pw;
            P.prof[blockIdx.x * 16 + 1] = SPHB_CLOCK() - pt0;
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (warp-uniform control flow, one elected lane issues) =====================
        if constexpr (DP64T > 0) {
            constexpr uint32_t idesc = ptx::make_idesc_f16(BM, BN);
            constexpr uint32_t STAGE_DESC = CHUNK_BYTES >> 4;  // descriptor address units (16 B) per ring slot
            uint64_t adesc[MH][DP64T];
#pragma unroll
            for (int m = 0; m < MH; ++m)
#pragma unroll
                for (int c = 0; c < DP64T; ++c) adesc[m][c] = ptx::make_smem_desc_sw128(sA + (m * DP64T + c) * CHUNK_BYTES);
            const uint64_t odesc = ptx::make_smem_desc_sw128(sO);
            const uint64_t bdesc0 = ptx::make_smem_desc_sw128(sB);
            const uint64_t ndesc0 = P.norm_bytes != CHUNK_BYTES ? ptx::make_smem_desc_sw32(sB) : bdesc0;
            int stage = 0, as = 0;
            uint32_t phase = 0, aphase = 0, accphase = 0;
            long long mw_te = 0, mw_f = 0, mt0 = SPHB_CLOCK();
            for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
                const SweepItem it = sweep_item(P, w);
                ptx::mbar_wait(bar_afull, aphase);
                aphase ^= 1;
                ptx::tc_fence_after();
                for (int t = 0; t < it.ntiles; ++t) {
                    const uint32_t tmem_d = tmem_base + as * ACC_COLS;
                    const bool last_tile = t == it.ntiles - 1;
                    if (P.half_major) {
                        // all chunks for M half 0, then for half 1: the epilogue warps of half 0 start while half 1 is
                        // still being multiplied, and a slow warp of one half only delays that half's next tile
#pragma unroll
                        for (int m = 0; m < MH; ++m) {
                            long long w0 = SPHB_CLOCK();
                            ptx::mbar_wait(bar_tempty + 8 * (as * MH + m), accphase ^ 1);
                            mw_te += SPHB_CLOCK() - w0;
                            ptx::tc_fence_after();
#pragma unroll
                            for (int c = 0; c <= DP64T; ++c) {
                                int st = stage + c;
                                uint32_t ph = phase;
                                if (st >= stages) { st -= stages; ph ^= 1; }
                                if (m == 0) {
                                    w0 = SPHB_CLOCK();
                                    ptx::mbar_wait(bar_full + 8 * st, ph);
                                    mw_f += SPHB_CLOCK() - w0;
                                    ptx::tc_fence_after();
                                }
                                if (ptx::elect_one_sync()) {
                                    if (c < DP64T) {
                                        const uint64_t bdesc = bdesc0 + static_cast<uint64_t>(st * STAGE_DESC);
#pragma unroll
                                        for (int kk = 0; kk < 4; ++kk)
                                            ptx::umma_f16(tmem_d + m * BN, adesc[m][c < DP64T ? c : 0] + 2 * kk, bdesc + 2 * kk, idesc, (c | kk) != 0 ? 1u : 0u);
                                    } else {
                                        ptx::umma_f16(tmem_d + m * BN, odesc, ndesc0 + static_cast<uint64_t>(st * STAGE_DESC), idesc, 1u);
                                    }
                                    if (m == MH - 1) ptx::umma_commit(bar_empty + 8 * st);
                                    if (c == DP64T) {
                                        ptx::umma_commit(bar_tfull + 8 * (as * MH + m));
                                        if (m == MH - 1 && last_tile) ptx::umma_commit(bar_aempty);  // query tile consumed
                                    }
                                }
                                __syncwarp();
                            }
                        }
                        stage += DP64T + 1;
                        if (stage >= stages) { stage -= stages; phase ^= 1; }
                    } else {
#pragma unroll
                        for (int m = 0; m < MH; ++m) ptx::mbar_wait(bar_tempty + 8 * (as * MH + m), accphase ^ 1);
                        ptx::tc_fence_after();
#pragma unroll
                        for (int c = 0; c < DP64T; ++c) {
                            ptx::mbar_wait(bar_full + 8 * stage, phase);
                            ptx::tc_fence_after();
                            if (ptx::elect_one_sync()) {
                                const uint64_t bdesc = bdesc0 + static_cast<uint64_t>(stage * STAGE_DESC);
#pragma unroll
                                for (int m = 0; m < MH; ++m)
#pragma unroll
                                    for (int kk = 0; kk < 4; ++kk)
                                        ptx::umma_f16(tmem_d + m * BN, adesc[m][c] + 2 * kk, bdesc + 2 * kk, idesc, (c | kk) != 0 ? 1u : 0u);
                                ptx::umma_commit(bar_empty + 8 * stage);
                            }
                            __syncwarp();
                            if (++stage == stages) { stage = 0; phase ^= 1; }
                        }
                        ptx::mbar_wait(bar_full + 8 * stage, phase);
                        ptx::tc_fence_after();
                        if (ptx::elect_one_sync()) {
                            const uint64_t ndesc = ndesc0 + static_cast<uint64_t>(stage * STAGE_DESC);
#pragma unroll
                            for (int m = 0; m < MH; ++m) ptx::umma_f16(tmem_d + m * BN, odesc, ndesc, idesc, 1u);  // ones x (n0 n1 n2)
                            ptx::umma_commit(bar_empty + 8 * stage);
#pragma unroll
                            for (int m = 0; m < MH; ++m) ptx::umma_commit(bar_tfull + 8 * (as * MH + m));
                            if (last_tile) ptx::umma_commit(bar_aempty);  // query tile consumed
                        }
                        __syncwarp();
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                    if (++as == 2) { as = 0; accphase ^= 1; }
                }
            }
            if (P.prof && lane == 0) {
                P.prof[blockIdx.x * 16 + 2] = mw_te;
                P.prof[blockIdx.x * 16 + 3] = mw_f;
                P.prof[blockIdx.x * 16 + 4] = SPHB_CLOCK() - mt0;
            }
        } else {
        constexpr uint32_t idesc = ptx::make_idesc_f16(BM, BN);
        int stage = 0;
        uint32_t phase = 0, aphase = 0;
        int as = 0;
        uint32_t accphase = 0;
        const uint64_t odesc = ptx::make_smem_desc_sw128(sO);
        const bool narrow = P.norm_bytes != CHUNK_BYTES;
        long long mw_te = 0, mw_f = 0, mt0 = SPHB_CLOCK();
        for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
            const SweepItem it = sweep_item(P, w);
            ptx::mbar_wait(bar_afull, aphase);
            aphase ^= 1;
            ptx::tc_fence_after();
            for (int t = 0; t < it.ntiles; ++t) {
                long long w0 = SPHB_CLOCK();
#pragma unroll
                for (int m = 0; m < MH; ++m) ptx::mbar_wait(bar_tempty + 8 * (as * MH + m), accphase ^ 1);
                mw_te += SPHB_CLOCK() - w0;
                ptx::tc_fence_after();
                for (int c = 0; c < nB; ++c) {
                    w0 = SPHB_CLOCK();
                    ptx::mbar_wait(bar_full + 8 * stage, phase);
                    mw_f += SPHB_CLOCK() - w0;
                    ptx::tc_fence_after();
                    const uint32_t sb = sB + stage * CHUNK_BYTES;
                    const uint64_t bdesc = ptx::make_smem_desc_sw128(sb);
                    if (ptx::elect_one_sync()) {
#pragma unroll
                        for (int m = 0; m < MH; ++m) {
                            const uint32_t tmem_d = tmem_base + as * ACC_COLS + m * BN;
                            if (c < dp64) {
                                // database hi chunk c  x  query hi chunk c (and, 3 passes, query lo chunk c)
                                const int nk = (c == dp64 - 1) ? k16_tail : 4;
                                const uint64_t ah = ptx::make_smem_desc_sw128(sA + (m * nA + c) * CHUNK_BYTES);
                                for (int kk = 0; kk < nk; ++kk) ptx::umma_f16(tmem_d, ah + 2 * kk, bdesc + 2 * kk, idesc, (c | kk) != 0);
                                if (three) {
                                    const uint64_t al = ptx::make_smem_desc_sw128(sA + (m * nA + dp64 + c) * CHUNK_BYTES);
                                    for (int kk = 0; kk < nk; ++kk) ptx::umma_f16(tmem_d, al + 2 * kk, bdesc + 2 * kk, idesc, 1);
                                }
                            } else if (c < nB - 1) {
                                // (3 passes) database lo chunk  x  query hi chunk
                                const int cc = c - dp64;
                                const int nk = (cc == dp64 - 1) ? k16_tail : 4;
                                const uint64_t ah = ptx::make_smem_desc_sw128(sA + (m * nA + cc) * CHUNK_BYTES);
                                for (int kk = 0; kk < nk; ++kk) ptx::umma_f16(tmem_d, ah + 2 * kk, bdesc + 2 * kk, idesc, 1);
                            } else {
                                // norm chunk: ones x (n0 n1 n2), one k-step
                                ptx::umma_f16(tmem_d, odesc, narrow ? ptx::make_smem_desc_sw32(sb) : bdesc, idesc, 1);
                            }
                        }
                        ptx::umma_commit(bar_empty + 8 * stage);
                        if (c == nB - 1) {
#pragma unroll
                            for (int m = 0; m < MH; ++m) ptx::umma_commit(bar_tfull + 8 * (as * MH + m));
                        }
                        if (c == nB - 1 && t == it.ntiles - 1) ptx::umma_commit(bar_aempty);  // query tile consumed
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                if (++as == 2) { as = 0; accphase ^= 1; }
            }
        }
        if (P.prof && lane == 0) {
            P.prof[blockIdx.x * 16 + 2] = mw_te;
            P.prof[blockIdx.x * 16 + 3] = mw_f;
            P.prof[blockIdx.x * 16 + 4] = SPHB_CLOCK() - mt0;
        }
        }
    } else {
        // ===================== epilogue: 4 warps per M half, one thread per query row =====================
        const int ew = warp - 2;       // 0 .. 4*MH-1
        const int quarter = warp & 3;  // TMEM lane quarter this warp may access
        const int mhalf = ew >> 2;
        constexpr int cap = 32 * R;
        const int k = P.k, keep_max = P.keep_max;
        const int out_mul = P.nseg > 1 ? P.nseg : 1;
        const BandConsts bc = make_band_consts(P);
        int as = 0;
        uint32_t accphase = 0;
        unsigned long long n_compact = 0, n_accept = 0;
        const uint32_t n32 = static_cast<uint32_t>(P.n);
        long long ew_tf = 0, e_ld = 0, e_filt = 0, e_exp = 0, e_cmp = 0, et0 = SPHB_CLOCK();
        for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
            const SweepItem it = sweep_item(P, w);
            const int row_in_tile = mhalf * BM + quarter * 32 + lane;
            const int64_t r = static_cast<int64_t>(it.qt) * (MH * BM) + row_in_tile;
            const bool valid = r < P.nq;
            uint64_t* warp_buf = P.cand + (static_cast<size_t>(blockIdx.x) * (MH * BM) + (row_in_tile - lane)) * cap;
            uint64_t* mybuf = warp_buf + static_cast<size_t>(lane) * cap;
            RowList me;
            init_row(me, P, r, valid);
            const int dt0 = sweep_item_start(P, it, MH * BM, BN);
            for (int t = 0; t < it.ntiles; ++t) {
                const int dt = sweep_tile(P, it, dt0, t);
                long long w0 = SPHB_CLOCK();
                ptx::mbar_wait(bar_tfull + 8 * (as * MH + mhalf), accphase);
                ew_tf += SPHB_CLOCK() - w0;
                ptx::tc_fence_after();
                const uint32_t jbase = static_cast<uint32_t>(dt) * BN;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * ACC_COLS + mhalf * BN;
#pragma unroll 1
                for (int c = 0; c < (P.debug_skip_epilogue ? 0 : BN / 32); c += 2) {
                    float v0[32], v1[32];
                    w0 = SPHB_CLOCK();
                    ptx::tmem_ld_32x32(taddr + c * 32, v0);       // two loads in flight before the wait
                    ptx::tmem_ld_32x32(taddr + c * 32 + 32, v1);
                    ptx::tmem_ld_wait(v0);
                    ptx::tmem_ld_wait(v1);
                    const long long w1 = SPHB_CLOCK();
                    e_ld += w1 - w0;
                    filter_pair(v0, v1, jbase + c * 32, n32, me, mybuf, n_accept);
                    compact_full_rows<R>(me, warp_buf, k, keep_max, lane, bc, n_compact, e_cmp);
                    e_filt += SPHB_CLOCK() - w1;
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * (as * MH + mhalf));
                if (++as == 2) { as = 0; accphase ^= 1; }
            }
            const long long x0 = SPHB_CLOCK();
            export_rows<R>(me, r - lane, P.nq, warp_buf, k, keep_max, P.cand_out, P.cnt_out, lane, bc, out_mul, it.seg);
            e_exp += SPHB_CLOCK() - x0;
        }
        if (P.prof && lane == 0) {
            atomicAdd(P.prof + blockIdx.x * 16 + 5, static_cast<unsigned long long>(ew_tf));
            atomicAdd(P.prof + blockIdx.x * 16 + 6, static_cast<unsigned long long>(e_ld));
            atomicAdd(P.prof + blockIdx.x * 16 + 7, static_cast<unsigned long long>(e_filt));
            atomicAdd(P.prof + blockIdx.x * 16 + 8, static_cast<unsigned long long>(e_exp));
            atomicAdd(P.prof + blockIdx.x * 16 + 9, static_cast<unsigned long long>(SPHB_CLOCK() - et0));
            atomicAdd(P.prof + blockIdx.x * 16 + 10, static_cast<unsigned long long>(e_cmp));
        }
        if (P.stats) {
            n_accept = warp_sum(static_cast<double>(n_accept));
            if (lane == 0) {
                atomicAdd(P.stats + 0, n_compact);
                atomicAdd(P.stats + 1, static_cast<unsigned long long>(n_accept));
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------
// sweep, streaming flavour (any d; used for d > 256)
// ---------------------------------------------------------------------------
namespace tcs {
constexpr int BN = 256;
constexpr int A_BYTES = tc::BM * tc::BK * 2;
constexpr int B_BYTES = BN * tc::BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TMEM_COLS = 2 * BN;
constexpr int NUM_THREADS = 192;
}  // namespace tcs

template <int R>
__global__ void __launch_bounds__(tcs::NUM_THREADS, 1)
knn_sweep_stream(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB, const SweepParams P) {
    using namespace tc;
    using namespace tcs;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    const int stages = P.stages;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(stages) * STAGE_BYTES);
    const uint32_t bar_full = ptx::smem_u32(bars);
    const uint32_t bar_empty = ptx::smem_u32(bars + stages);
    const uint32_t bar_tfull = ptx::smem_u32(bars + 2 * stages);
    const uint32_t bar_tempty = ptx::smem_u32(bars + 2 * stages + 2);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t smem_base = ptx::smem_u32(smem);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmapA);
        ptx::prefetch_tmap(&tmapB);
        for (int s = 0; s < stages; ++s) {
            ptx::mbar_init(bar_full + 8 * s, 1);
            ptx::mbar_init(bar_empty + 8 * s, 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(bar_tfull + 8 * s, 1);
            ptx::mbar_init(bar_tempty + 8 * s, 4);
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
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            // Soft lock-step as in the resident kernel (sweep_pace): a database tile of this flavour is 256 rows x ~2 KB, the L2
            // holds ~240 of them, and without coupling the CTAs drift apart until every CTA streams the database from DRAM on its
            // own (full C5: 582 TFLOP/s, HBM-bound; a short 8-wave slab, where they cannot drift far, reaches 984).
            SweepPace pace;
            for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
                const int32_t qrow = static_cast<int32_t>(P.a_row0 + static_cast<int64_t>(qt) * BM);
                const int dt0 = sweep_start(P, qt, BM, BN);
                for (int t = 0; t < P.n_dbtiles; ++t) {
                    const int dt = dt0 + t < P.n_dbtiles ? dt0 + t : dt0 + t - P.n_dbtiles;  // dt0, t < n_dbtiles
                    if ((t & 3) == 0) sweep_pace_one(P, pace);
                    ++pace.mine;
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
            if (P.pace_window > 0) {  // a CTA that has finished publishes the progress of a complete run
                const long long waves = (P.n_qtiles + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
                const long long all = waves * P.n_dbtiles;
                if (all > pace.published) atomicAdd(P.progress, static_cast<unsigned long long>(all - pace.published));
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = ptx::make_idesc_f16(BM, BN);
            int stage = 0;
            uint32_t phase = 0;
            int as = 0;
            uint32_t accphase = 0;
            for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
                for (int dt = 0; dt < P.n_dbtiles; ++dt) {
                    ptx::mbar_wait(bar_tempty + 8 * as, accphase ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t tmem_d = tmem_base + as * BN;
                    for (int kc = 0; kc < P.kchunks; ++kc) {
                        ptx::mbar_wait(bar_full + 8 * stage, phase);
                        ptx::tc_fence_after();
                        const uint32_t sa = smem_base + stage * STAGE_BYTES;
                        const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
                        const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + A_BYTES);
                        const int nk = (kc == P.kchunks - 1) ? P.k16_last : (BK / UMMA_K);
                        for (int kk = 0; kk < nk; ++kk) ptx::umma_f16(tmem_d, adesc + 2 * kk, bdesc + 2 * kk, idesc, (kc | kk) != 0);
                        ptx::umma_commit(bar_empty + 8 * stage);
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                    ptx::umma_commit(bar_tfull + 8 * as);
                    if (++as == 2) { as = 0; accphase ^= 1; }
                }
            }
        }
    } else {
        const int quarter = warp & 3;
        constexpr int cap = 32 * R;
        const int k = P.k, keep_max = P.keep_max;
        const BandConsts bc = make_band_consts(P);
        int as = 0;
        uint32_t accphase = 0;
        unsigned long long n_compact = 0, n_accept = 0;
        long long e_cmp = 0;
        const uint32_t n32 = static_cast<uint32_t>(P.n);
        for (int qt = blockIdx.x; qt < P.n_qtiles; qt += gridDim.x) {
            const int row_in_tile = quarter * 32 + lane;
            const int64_t r = static_cast<int64_t>(qt) * BM + row_in_tile;
            const bool valid = r < P.nq;
            uint64_t* warp_buf = P.cand + (static_cast<size_t>(blockIdx.x) * BM + quarter * 32) * cap;
            uint64_t* mybuf = warp_buf + static_cast<size_t>(lane) * cap;
            RowList me;
            init_row(me, P, r, valid);
            const int dt0 = sweep_start(P, qt, BM, BN);
            for (int t = 0; t < P.n_dbtiles; ++t) {
                const int dt = dt0 + t < P.n_dbtiles ? dt0 + t : dt0 + t - P.n_dbtiles;  // dt0, t < n_dbtiles
                ptx::mbar_wait(bar_tfull + 8 * as, accphase);
                ptx::tc_fence_after();
                const uint32_t jbase = static_cast<uint32_t>(dt) * BN;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + as * BN;
#pragma unroll 1
                for (int c = 0; c < BN / 32; c += 2) {
                    float v0[32], v1[32];
                    ptx::tmem_ld_32x32(taddr + c * 32, v0);       // two loads in flight before the wait
                    ptx::tmem_ld_32x32(taddr + c * 32 + 32, v1);
                    ptx::tmem_ld_wait(v0);
                    ptx::tmem_ld_wait(v1);
#ifdef SPHB_FILTER_SINGLE
                    filter_chunk(v0, jbase + c * 32, n32, me, mybuf, n_accept);
                    compact_full_rows<R>(me, warp_buf, k, keep_max, lane, bc, n_compact, e_cmp);
                    filter_chunk(v1, jbase + c * 32 + 32, n32, me, mybuf, n_accept);
                    compact_full_rows<R>(me, warp_buf, k, keep_max, lane, bc, n_compact, e_cmp);
#else
                    filter_pair(v0, v1, jbase + c * 32, n32, me, mybuf, n_accept);
                    compact_full_rows<R>(me, warp_buf, k, keep_max, lane, bc, n_compact, e_cmp);
#endif
                }
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(bar_tempty + 8 * as);
                if (++as == 2) { as = 0; accphase ^= 1; }
            }
            export_rows<R>(me, r - lane, P.nq, warp_buf, k, keep_max, P.cand_out, P.cnt_out, lane, bc);
        }
        if (P.stats) {
            n_accept = warp_sum(static_cast<double>(n_accept));
            if (lane == 0) {
                atomicAdd(P.stats + 0, n_compact);
                atomicAdd(P.stats + 1, static_cast<unsigned long long>(n_accept));
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------
// exact re-rank of the surviving candidates (one warp per query row)
// ---------------------------------------------------------------------------
// Row r of this launch is point gi = qrows ? qrows[r] : q0 + r; it is written to output row gi - out_q0.
// nseg > 1: the row's candidates are the union of its nseg per-segment lists (disjoint database ranges, sweep_item).
__global__ void __launch_bounds__(256) knn_rerank_kernel(const float* __restrict__ X, int d, int k, int cap, int keep_max, int nseg, int metric,
                                                         int l2_squared, int64_t q0, int64_t nq, const int64_t* __restrict__ qrows,
                                                         int64_t out_q0, const uint32_t* __restrict__ cand, const int* __restrict__ cnt,
                                                         int64_t* __restrict__ ids, float* __restrict__ dist,
                                                         int64_t* __restrict__ ovf_rows, int* __restrict__ ovf_count) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
    for (int64_t r = gw; r < nq; r += nw) {
        const int64_t gi = qrows ? qrows[r] : q0 + r;
        bool ovf = false;
        for (int sg = 0; sg < nseg; ++sg) ovf |= (cnt[r * nseg + sg] & OVERFLOW_FLAG) != 0;
        if (ovf) {
            if (lane == 0) ovf_rows[atomicAdd(ovf_count, 1)] = gi;
            continue;
        }
        int m = 0;
        for (int sg = 0; sg < nseg; ++sg) {
            const int c = cnt[r * nseg + sg];
            const uint32_t* src = cand + (r * nseg + sg) * keep_max;
            for (int t = lane; t < c; t += 32) {
                const uint32_t j = src[t];
                sm[m + t] = make_key(exact_key(X, gi, static_cast<int64_t>(j), d, metric), j);
            }
            m += c;
        }
        for (int t = m + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        const int64_t orow = gi - out_q0;
        write_row(sm, min(m, k), k, gi, metric, l2_squared, ids + orow * k, dist + orow * k, lane);
        __syncwarp();
    }
}

// The same re-rank with the candidate rows STAGED through shared memory (d % 4 == 0).  In knn_rerank_kernel every lane walks
// its own candidate row in global memory: per 4 channels one LDG.128 whose 32 lanes touch 32 different rows (32 L1 wavefronts).
// Here a group of 32 candidates is fetched chunk by chunk (32 channels = 128 B per row): one LDG.128 covers four rows with eight
// lanes each (4 wavefronts), the chunk lands in shared memory (row stride 36 floats: the quarter-warps of an LDS.128 hit disjoint
// banks), and lane l runs the exact fp64 chain of candidate l; the query chunk is converted to fp64 once and read as a broadcast.
// Same chain, same order (c = 0 .. d-1, one accumulator per candidate) as exact_key — bit-identical keys.  About 9 KB of shared
// memory per warp, so 24 warps per SM stay resident to cover the latency of the dependent DFMA chain.
constexpr int RR_DCH = 32;
constexpr int RR_LD = RR_DCH + 4;
constexpr int RR_WARPS = 8;
__host__ __device__ constexpr size_t rerank_staged_smem_per_warp(int cap) {
    return static_cast<size_t>(cap) * 8 + RR_DCH * 8 + 32 * RR_LD * 4;
}

__global__ void __launch_bounds__(RR_WARPS * 32, 3) knn_rerank_staged_kernel(const float* __restrict__ X, int d, int k, int cap, int keep_max, int nseg,
                                                                          int metric, int l2_squared, int64_t q0, int64_t nq,
                                                                          const int64_t* __restrict__ qrows, int64_t out_q0,
                                                                          const uint32_t* __restrict__ cand, const int* __restrict__ cnt,
                                                                          int64_t* __restrict__ ids, float* __restrict__ dist,
                                                                          int64_t* __restrict__ ovf_rows, int* __restrict__ ovf_count) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char* base = smem_raw + static_cast<size_t>(warp) * rerank_staged_smem_per_warp(cap);
    uint64_t* sm = reinterpret_cast<uint64_t*>(base);
    double* qd = reinterpret_cast<double*>(base + static_cast<size_t>(cap) * 8);
    float* stage = reinterpret_cast<float*>(base + static_cast<size_t>(cap) * 8 + RR_DCH * 8);
    const int sub = lane >> 3, col4 = (lane & 7) * 4;  // staging role: row 4c + sub, channels col4 .. col4 + 3 of the chunk
    const int64_t gw = static_cast<int64_t>(blockIdx.x) * RR_WARPS + warp;
    const int64_t nw = static_cast<int64_t>(gridDim.x) * RR_WARPS;
    for (int64_t r = gw; r < nq; r += nw) {
        const int64_t gi = qrows ? qrows[r] : q0 + r;
        bool ovf = false;
        for (int sg = 0; sg < nseg; ++sg) ovf |= (cnt[r * nseg + sg] & OVERFLOW_FLAG) != 0;
        if (ovf) {
            if (lane == 0) ovf_rows[atomicAdd(ovf_count, 1)] = gi;
            continue;
        }
        int m = 0;
        for (int sg = 0; sg < nseg; ++sg) {
            const int c = cnt[r * nseg + sg];
            const uint32_t* src = cand + (r * nseg + sg) * keep_max;
            for (int t = lane; t < c; t += 32) sm[m + t] = src[t];  // ids first; the keys replace them group by group
            m += c;
        }
        __syncwarp();
        const float* xq = X + gi * d;
        for (int g0 = 0; g0 < m; g0 += 32) {
            const int t = g0 + lane;
            const int nv = min(32, m - g0);
            const uint32_t jm = t < m ? static_cast<uint32_t>(sm[t]) : 0u;
            const float* xr[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) xr[c] = X + static_cast<int64_t>(__shfl_sync(0xffffffffu, jm, 4 * c + sub)) * d + col4;
            double acc = 0.0;
            for (int c0 = 0; c0 < d; c0 += RR_DCH) {
                const bool in = c0 + col4 < d;
                if (sub == 0 && in) {
                    const float4 q = __ldg(reinterpret_cast<const float4*>(xq + c0 + col4));
                    qd[col4] = q.x; qd[col4 + 1] = q.y; qd[col4 + 2] = q.z; qd[col4 + 3] = q.w;
                }
                float4 v[8];
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    if (in && 4 * c + sub < nv) v[c] = __ldg(reinterpret_cast<const float4*>(xr[c] + c0));
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    if (in && 4 * c + sub < nv) *reinterpret_cast<float4*>(stage + (4 * c + sub) * RR_LD + col4) = v[c];
                __syncwarp();
                if (lane < nv) {
                    const float* mine = stage + lane * RR_LD;
                    const int len = min(RR_DCH, d - c0);
                    if (metric == METRIC_L2) {
#pragma unroll
                        for (int e = 0; e < RR_DCH; e += 4) {
                            if (e < len) {
                                const float4 w = *reinterpret_cast<const float4*>(mine + e);
                                const double2 qa = *reinterpret_cast<const double2*>(qd + e), qb = *reinterpret_cast<const double2*>(qd + e + 2);
                                double tt;
                                tt = qa.x - static_cast<double>(w.x); acc = fma(tt, tt, acc);
                                tt = qa.y - static_cast<double>(w.y); acc = fma(tt, tt, acc);
                                tt = qb.x - static_cast<double>(w.z); acc = fma(tt, tt, acc);
                                tt = qb.y - static_cast<double>(w.w); acc = fma(tt, tt, acc);
                            }
                        }
                    } else {
#pragma unroll
                        for (int e = 0; e < RR_DCH; e += 4) {
                            if (e < len) {
                                const float4 w = *reinterpret_cast<const float4*>(mine + e);
                                const double2 qa = *reinterpret_cast<const double2*>(qd + e), qb = *reinterpret_cast<const double2*>(qd + e + 2);
                                acc = fma(qa.x, static_cast<double>(w.x), acc);
                                acc = fma(qa.y, static_cast<double>(w.y), acc);
                                acc = fma(qb.x, static_cast<double>(w.z), acc);
                                acc = fma(qb.y, static_cast<double>(w.w), acc);
                            }
                        }
                    }
                }
                __syncwarp();
            }
            const float v32 = static_cast<float>(acc);
            if (t < m) sm[t] = make_key(metric == METRIC_L2 ? v32 : -v32, jm);
        }
        for (int t = m + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        const int64_t orow = gi - out_q0;
        write_row(sm, min(m, k), k, gi, metric, l2_squared, ids + orow * k, dist + orow * k, lane);
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
int knn_simt_launch(const float*, int64_t, int, int, int, int, int64_t, int64_t, const int64_t*, const int*, int64_t, int64_t*, float*, cudaStream_t);

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

// 2-D fp16 row-major [rows, cols] tensor, box = box_cols columns x box_rows rows, zero OOB fill.
// box_cols = 64 -> SWIZZLE_128B (one 128 B swizzle span per row), box_cols = 16 -> SWIZZLE_32B.
static int make_tmap(CUtensorMap* tm, const void* base, int64_t rows, int64_t cols, int box_rows, int box_cols = tc::BK) {
    EncodeFn enc = get_encode_fn();
    if (!enc) return fail(3, "cuTensorMapEncodeTiled entry point not found");
    const cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * 2};
    const cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapSwizzle sw = box_cols == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(3, "cuTensorMapEncodeTiled failed: " + std::to_string(static_cast<int>(r)));
    return 0;
}

static int next_pow2_i(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// per host thread, like the context (api.cu): the workers of sphb_knn_graph_prow_multi run this file concurrently
static thread_local int64_t g_knn_stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
static thread_local float g_knn_ms[4] = {0.f, 0.f, 0.f, 0.f};  // prep, main sweep(s), re-rank(s), fallback (CUDA events on the launching stream)
void knn_get_timing(float* ms4) { for (int i = 0; i < 4; ++i) ms4[i] = g_knn_ms[i]; }
void knn_set_stats(const int64_t* s) { for (int i = 0; i < 8; ++i) g_knn_stats[i] = s[i]; }
void knn_get_stats(int64_t* s) { for (int i = 0; i < 8; ++i) s[i] = g_knn_stats[i]; }
// global ids of the rows of the last call that left the main sweep: [0] re-swept with the 3-product filter, [1] recomputed
// by the exact SIMT kernel (parity checks sample them, bench.py "parity")
static thread_local std::vector<int64_t> g_knn_ovf_rows[2];
void knn_clear_overflow_rows() { g_knn_ovf_rows[0].clear(); g_knn_ovf_rows[1].clear(); }
const std::vector<int64_t>& knn_overflow_rows(int which) { return g_knn_ovf_rows[which ? 1 : 0]; }
static int record_overflow_rows(int which, const int64_t* d_rows, int count, cudaStream_t stream) {
    if (count <= 0) return 0;
    std::vector<int64_t>& v = g_knn_ovf_rows[which];
    const size_t old = v.size();
    v.resize(old + static_cast<size_t>(count));
    SPHB_CUDA(cudaMemcpyAsync(v.data() + old, d_rows, static_cast<size_t>(count) * sizeof(int64_t), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaStreamSynchronize(stream));
    return 0;
}

bool knn_tc_supported(int64_t n, int d, int k) {
    if (d < 8 || n < 1024) return false;
    if (n >= (1ll << 31)) return false;
    if (3 * d + 3 > 65536) return false;
    if (next_pow2_i(k + 160) > 1024) return false;
    return true;
}

namespace {

struct EventTimer {
    cudaEvent_t a = nullptr, b = nullptr;
    EventTimer() { cudaEventCreate(&a); cudaEventCreate(&b); }
    ~EventTimer() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    void start(cudaStream_t s) { cudaEventRecord(a, s); }
    void stop(cudaStream_t s) { cudaEventRecord(b, s); }
    float ms() { float v = 0.f; cudaEventSynchronize(b); cudaEventElapsedTime(&v, a, b); return v; }
};

// fp16 operand matrices of one (passes, flavour) combination
struct Operands {
    int passes = 0, MH = 0;       // MH = 0: streaming flavour
    int lda = 0, ldb = 0, ldn = 0;
    int64_t a_rows = 0;           // rows of A (np for the full matrix, padded list length for a gathered one)
    DevBuf<__half> A, B, N;
    bool built_b = false;
};

// Everything one kNN call shares between its sweeps.
struct KnnJob {
    const float* x; int64_t n; int d, k, metric, l2_squared;
    int64_t np; int dp; int centre, sms;
    cudaStream_t stream;
    DevBuf<double> colsum, un2, scal;
    DevBuf<unsigned long long> maxbits, stats;
    DevBuf<float> dun;
    DevBuf<unsigned int> dvmax;
    DevBuf<__half> ones;
    bool narrow_norm;
    int64_t sum_compactions = 0, sum_accepted = 0;
};

int build_operands(KnnJob& J, Operands& op, int passes, int MH, const int64_t* d_rowlist, int64_t nlist, bool want_a, bool want_b) {
    const int d = J.d, dp = J.dp;
    const int layout = MH ? 1 : 0;
    op.passes = passes; op.MH = MH;
    const int kp = ((passes == 3 ? 3 * d + 3 : d + 3) + 15) / 16 * 16;
    op.lda = MH ? (passes == 3 ? 2 * dp : dp) : kp;
    op.ldb = op.lda;
    op.ldn = MH ? (J.narrow_norm ? 16 : 64) : 0;
    const int64_t rows_a = d_rowlist ? (nlist + 255) / 256 * 256 : J.np;
    cudaStream_t s = J.stream;
    if (want_a) {
        op.a_rows = rows_a;
        SPHB_CUDA(op.A.alloc(static_cast<size_t>(rows_a) * op.lda, s));
        SPHB_CUDA(cudaMemsetAsync(op.A.p, 0, static_cast<size_t>(rows_a) * op.lda * sizeof(__half), s));
    }
    if (want_b) {
        SPHB_CUDA(op.B.alloc(static_cast<size_t>(J.np) * op.ldb, s));
        SPHB_CUDA(cudaMemsetAsync(op.B.p, 0, static_cast<size_t>(J.np) * op.ldb * sizeof(__half), s));
        if (MH) {
            SPHB_CUDA(op.N.alloc(static_cast<size_t>(J.np) * op.ldn, s));
            SPHB_CUDA(cudaMemsetAsync(op.N.p, 0, static_cast<size_t>(J.np) * op.ldn * sizeof(__half), s));
        }
        op.built_b = true;
    }
    const int with_norm = J.metric == METRIC_L2 ? 1 : 0;
    if (d_rowlist) {
        // gathered A rows only (the database side is always the full matrix)
        knn_prep_split<<<J.sms * 8, 256, 0, s>>>(J.x, J.n, nlist, rows_a, d_rowlist, d, layout, passes, op.lda, op.ldb, op.ldn, dp, J.colsum.p,
                                                 J.centre, with_norm, J.scal.p, J.un2.p, op.A.p, nullptr, nullptr, nullptr, nullptr);
    } else {
        knn_prep_split<<<J.sms * 8, 256, 0, s>>>(J.x, J.n, J.n, J.np, nullptr, d, layout, passes, op.lda, op.ldb, op.ldn, dp, J.colsum.p, J.centre,
                                                 with_norm, J.scal.p, J.un2.p, want_a ? op.A.p : nullptr, want_b ? op.B.p : nullptr,
                                                 want_b && MH ? op.N.p : nullptr, J.dun.p, J.dvmax.p);
    }
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

// Tuning knobs of the sweep.  Release builds use the defaults; the environment is read only in builds with
// -DSPHB_KNN_TUNING (the profile build), so the library adds no public switches.
static int tune(const char* name, int def) {
#ifdef SPHB_KNN_TUNING
    if (const char* e = getenv(name)) return atoi(e);
#else
    (void)name;
#endif
    return def;
}

// One sweep + exact re-rank of query rows: either the slab [qa, qb) (A rows taken from the full operand
// matrix) or the explicit list d_qrows[0..nq) (A rows from the gathered matrix opA).  Rows whose band
// overflowed are appended to ovf_rows / ovf_count (device) and their number is returned in *n_ovf.
int sweep_and_rerank(KnnJob& J, const Operands& opA, const Operands& opB, int64_t qa, int64_t qb, const int64_t* d_qrows, int64_t out_q0,
                     int64_t* d_ids, float* d_dist, int64_t* d_ovf_rows, int* d_ovf_count, int* n_ovf, float* sweep_ms, float* rerank_ms) {
    using namespace tc;
    cudaStream_t stream = J.stream;
    const int passes = opB.passes, MH = opA.MH, d = J.d, k = J.k, dp = J.dp;  // opA.MH: rows per CTA (the B layout is the same for MH 1 and 2)
    const int64_t nq = qb - qa;
    if (nq <= 0) { *n_ovf = 0; return 0; }
    const int cap = passes == 1 ? std::max(512, next_pow2_i(k + 160)) : std::max(256, next_pow2_i(k + 96));
    const int keep_max = cap - 64;
    const int bn = MH ? 128 : tcs::BN;
    const int rows_per_cta = MH ? MH * BM : BM;
    const int n_qtiles = static_cast<int>((nq + rows_per_cta - 1) / rows_per_cta);
    const int n_dbtiles = static_cast<int>(J.np / bn);
    const bool lean = passes == 1 && MH == 2 && cap == 512 && !tune("SPHB_KNN_GENERIC_ISSUE", 0);
    // Database segments (resident flavour, re-sweep of a row list): enough (tile, segment) items for ~4 per SM, at most
    // what the re-rank's shared-memory sort holds (2048 keys per row) and segments of at least 32 tiles.
    int nseg = 1, seg_tiles = n_dbtiles;
    if (MH && d_qrows) {
        const int want = (4 * J.sms + n_qtiles - 1) / n_qtiles;
        nseg = std::max(1, std::min({want, 2048 / keep_max, n_dbtiles / 32, tune("SPHB_KNN_MAX_SEG", 64)}));
        seg_tiles = (n_dbtiles + nseg - 1) / nseg;
        nseg = (n_dbtiles + seg_tiles - 1) / seg_tiles;  // every segment holds at least one tile
    }
    const int n_items = n_qtiles * nseg;
    const int grid = std::min(J.sms, n_items);

    DevBuf<uint64_t> cand;
    DevBuf<uint32_t> cand_out;
    DevBuf<int> cnt_out;
    SPHB_CUDA(cand.alloc(static_cast<size_t>(grid) * rows_per_cta * cap, stream));
    SPHB_CUDA(cand_out.alloc(static_cast<size_t>(nq) * nseg * keep_max, stream));
    SPHB_CUDA(cnt_out.alloc(static_cast<size_t>(nq) * nseg, stream));
    SPHB_CUDA(cudaMemsetAsync(J.stats.p, 0, 3 * sizeof(unsigned long long), stream));

    SweepParams P{};
    P.un2 = J.un2.p; P.dun = J.dun.p; P.dvmax = reinterpret_cast<const float*>(J.dvmax.p); P.qrows = d_qrows;
    P.n = J.n; P.k = k; P.cap = cap; P.keep_max = keep_max; P.metric = J.metric;
    P.q0 = qa; P.nq = nq; P.a_row0 = d_qrows ? 0 : qa; P.n_qtiles = n_qtiles; P.n_dbtiles = n_dbtiles;
    P.nseg = nseg; P.seg_tiles = seg_tiles;
    P.passes = passes;
    P.cand = cand.p; P.cand_out = cand_out.p; P.cnt_out = cnt_out.p; P.stats = J.stats.p;
    P.debug_skip_epilogue = tune("SPHB_KNN_DEBUG_SKIP_EPI", 0);
    P.progress = J.stats.p + 2;
    P.l2_prefetch = tune("SPHB_KNN_L2PF", 0);
    P.compact_step = std::max(32, tune("SPHB_KNN_COMPACT_STEP", cap));
    // soft lock-step of the CTAs' sweeps (sweep_pace): only for whole-database items, and only when every CTA of the grid is
    // resident at once (one CTA per SM by shared memory; the occupancy query below guards partitioned / shared devices)
    P.pace_window = (MH && nseg == 1) ? tune("SPHB_KNN_PACE", 64) : (!MH ? tune("SPHB_KNN_STREAM_PACE", 16) : 0);
    DevBuf<unsigned long long> prof;
#ifdef SPHB_KNN_PROFILE
    const bool want_prof = getenv("SPHB_KNN_PROF") != nullptr;
#else
    const bool want_prof = false;
#endif
    if (want_prof) {
        SPHB_CUDA(prof.alloc(static_cast<size_t>(grid) * 16, stream));
        SPHB_CUDA(cudaMemsetAsync(prof.p, 0, static_cast<size_t>(grid) * 16 * sizeof(unsigned long long), stream));
        P.prof = prof.p;
    }
    // CTA b starts its database sweep b*stagger tiles in: the CTAs then read different lines at any
    // moment (no L2 hot lines) while the union of their positions stays inside an L2-sized window.
    P.stagger = MH ? tune("SPHB_KNN_STAGGER", 0) : tune("SPHB_KNN_STREAM_STAGGER", 0);  // streaming tiles are 256 rows x the whole K extent (516 KB at C5): any stagger spreads the CTAs over more than L2 holds
    P.lead = tune("SPHB_KNN_LEAD_ROWS", 16384) / bn;
    // Error models.  3 passes: split truncation 6·2^-22 + accumulation (one 2^-22 per
    // k-step), times the magnitude (2|u||v| + |v|²), plus the fp16 subnormal floor; safety factor 2.
    // 1 pass: operand rounding is bounded from the measured |u - fp16(u)| norms (knn_epilogue.cuh); the
    // accumulation coefficient is 8·2^-22 per k-step.
    const int kp = opB.lda;
    const int ksteps = MH ? ((passes == 3 ? 3 : 1) * dp / UMMA_K + 1) : kp / UMMA_K;
    P.eps_c1 = 2.0 * (6.0 + ksteps) * std::ldexp(1.0, -22);
    P.eps_c2 = 2.0 * std::ldexp(1.0, -23) * std::sqrt(static_cast<double>(d));
    P.acc_c = static_cast<float>(8.0 * (ksteps + 1) * std::ldexp(1.0, -22));
    P.sub_c = static_cast<float>(std::sqrt(static_cast<double>(d)) * std::ldexp(1.0, -25) * 1.01);

    CUtensorMap tmA, tmB, tmN, tmO;
    if (int rc = make_tmap(&tmA, opA.A.p, opA.a_rows, opA.lda, BM)) return rc;
    if (int rc = make_tmap(&tmB, opB.B.p, J.np, opB.ldb, bn)) return rc;
    if (int rc = make_tmap(&tmO, J.ones.p, 128, 64, 128)) return rc;

    const size_t smem_limit = 227 * 1024;
    EventTimer tsweep, trerank;
    tsweep.start(stream);
    if (MH) {
        if (int rc = make_tmap(&tmN, opB.N.p, J.np, opB.ldn, bn, opB.ldn == 16 ? 16 : 64)) return rc;
        const int dp64 = dp / 64;
        const int nA = (passes == 3 ? 2 : 1) * dp64;
        const size_t fixed = static_cast<size_t>(MH * nA + 1) * CHUNK_BYTES + 256 + 1024;
        int stages = 8;
        while (stages > 2 && fixed + static_cast<size_t>(stages) * CHUNK_BYTES > smem_limit) --stages;
        const size_t smem = fixed + static_cast<size_t>(stages) * CHUNK_BYTES;
        if (smem > smem_limit) return fail(5, "knn: shared memory budget exceeded for the resident kernel");
        P.dp64 = dp64; P.stages = stages;
        P.half_major = (stages >= 2 * (nA + 1) && !tune("SPHB_KNN_CHUNK_MAJOR", 0)) ? 1 : 0;
        P.norm_bytes = opB.ldn == 16 ? 128 * 32 : CHUNK_BYTES;
        const int tail = d - 64 * (dp64 - 1);
        const int k16_tail = (tail + 15) / 16;
#define SPHB_LAUNCH_RESIDENT(MH_, R_, DP_)                                                                                                      \
    do {                                                                                                                                        \
        auto kern = knn_sweep_resident<MH_, R_, DP_>;                                                                                           \
        SPHB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));                             \
        if (P.pace_window > 0) {                                                                                                                \
            /* the pacing spin needs every CTA of the grid resident: check what the device can host right now */                   \
            int per_sm = 0;                                                                                                                     \
            SPHB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 64 + 128 * MH_, smem));                                      \
            if (static_cast<int64_t>(per_sm) * J.sms < grid) P.pace_window = 0;                                                                 \
        }                                                                                                                                       \
        kern<<<grid, 64 + 128 * MH_, smem, stream>>>(tmA, tmB, tmN, tmO, P, k16_tail);                                                          \
    } while (0)
        if (lean && dp64 == 1) SPHB_LAUNCH_RESIDENT(2, 16, 1);
        else if (lean && dp64 == 2) SPHB_LAUNCH_RESIDENT(2, 16, 2);
        else if (lean && dp64 == 3) SPHB_LAUNCH_RESIDENT(2, 16, 3);
        else if (lean && dp64 == 4) SPHB_LAUNCH_RESIDENT(2, 16, 4);
        else if (MH == 2 && cap == 256) SPHB_LAUNCH_RESIDENT(2, 8, 0);
        else if (MH == 2 && cap == 512) SPHB_LAUNCH_RESIDENT(2, 16, 0);
        else if (MH == 2) SPHB_LAUNCH_RESIDENT(2, 32, 0);
        else if (cap == 256) SPHB_LAUNCH_RESIDENT(1, 8, 0);
        else if (cap == 512) SPHB_LAUNCH_RESIDENT(1, 16, 0);
        else SPHB_LAUNCH_RESIDENT(1, 32, 0);
#undef SPHB_LAUNCH_RESIDENT
    } else {
        const int kchunks = (kp + BK - 1) / BK;
        P.kchunks = kchunks;
        P.k16_last = (kp - (kchunks - 1) * BK) / UMMA_K;
        const size_t fixed = 256 + 1024;
        int stages = 4;
        while (stages > 2 && static_cast<size_t>(stages) * tcs::STAGE_BYTES + fixed > smem_limit) --stages;
        P.stages = stages;
        const size_t smem = static_cast<size_t>(stages) * tcs::STAGE_BYTES + fixed;
#define SPHB_LAUNCH_STREAM(R_)                                                                                                         \
    do {                                                                                                                               \
        SPHB_CUDA(cudaFuncSetAttribute(knn_sweep_stream<R_>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));     \
        if (P.pace_window > 0) { /* the pacing spin needs every CTA of the grid resident */                                             \
            int per_sm = 0;                                                                                                            \
            SPHB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, knn_sweep_stream<R_>, tcs::NUM_THREADS, smem));            \
            if (static_cast<int64_t>(per_sm) * J.sms < grid) P.pace_window = 0;                                                         \
        }                                                                                                                              \
        knn_sweep_stream<R_><<<grid, tcs::NUM_THREADS, smem, stream>>>(tmA, tmB, P);                                                   \
    } while (0)
        if (cap == 256) SPHB_LAUNCH_STREAM(8);
        else if (cap == 512) SPHB_LAUNCH_STREAM(16);
        else SPHB_LAUNCH_STREAM(32);
#undef SPHB_LAUNCH_STREAM
    }
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    tsweep.stop(stream);

    trerank.start(stream);
    {
        const int cap_sort = next_pow2_i(std::max(cap, nseg * keep_max));
        const size_t smem_staged = RR_WARPS * rerank_staged_smem_per_warp(cap_sort);
        if ((d & 3) == 0 && smem_staged <= 200 * 1024 && tune("SPHB_KNN_RERANK_STAGED", 1)) {
            SPHB_CUDA(cudaFuncSetAttribute(knn_rerank_staged_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_staged)));
            const int64_t blocks = std::min<int64_t>((nq + RR_WARPS - 1) / RR_WARPS, static_cast<int64_t>(J.sms) * 6);
            knn_rerank_staged_kernel<<<static_cast<unsigned>(blocks), RR_WARPS * 32, smem_staged, stream>>>(
                J.x, d, k, cap_sort, keep_max, nseg, J.metric, J.l2_squared, qa, nq, d_qrows, out_q0, cand_out.p, cnt_out.p, d_ids, d_dist,
                d_ovf_rows, d_ovf_count);
        } else {
            const size_t smem = 8 * static_cast<size_t>(cap_sort) * sizeof(uint64_t);
            SPHB_CUDA(cudaFuncSetAttribute(knn_rerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
            const int64_t blocks = std::min<int64_t>((nq + 7) / 8, static_cast<int64_t>(J.sms) * 8);
            knn_rerank_kernel<<<static_cast<unsigned>(blocks), 256, smem, stream>>>(J.x, d, k, cap_sort, keep_max, nseg, J.metric, J.l2_squared, qa,
                                                                                     nq, d_qrows, out_q0, cand_out.p, cnt_out.p, d_ids, d_dist,
                                                                                     d_ovf_rows, d_ovf_count);
        }
        count_launch();
        SPHB_CUDA(cudaGetLastError());
    }
    trerank.stop(stream);

    unsigned long long hstats[2] = {0, 0};
    SPHB_CUDA(cudaMemcpyAsync(hstats, J.stats.p, sizeof(hstats), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaMemcpyAsync(n_ovf, d_ovf_count, sizeof(int), cudaMemcpyDeviceToHost, stream));
    SPHB_CUDA(cudaStreamSynchronize(stream));
    J.sum_compactions += static_cast<int64_t>(hstats[0]);
    J.sum_accepted += static_cast<int64_t>(hstats[1]);
    *sweep_ms += tsweep.ms();
    *rerank_ms += trerank.ms();
    if (want_prof) {
        std::vector<unsigned long long> hp(static_cast<size_t>(grid) * 16);
        cudaMemcpy(hp.data(), prof.p, hp.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
        double acc[16] = {0};
        for (int b = 0; b < grid; ++b)
            for (int i = 0; i < 16; ++i) acc[i] += static_cast<double>(hp[b * 16 + i]) / grid;
        const int ew = MH ? 4 * MH : 4;
        fprintf(stderr,
                "[knn prof] passes %d nseg %d mean cycles per CTA: producer wait_empty %.3g of %.3g | mma wait_tempty %.3g wait_full %.3g of %.3g | "
                "epilogue(per warp) wait_tfull %.3g ld %.3g filter+compact %.3g (compact %.3g) export %.3g of %.3g | sweep %.2f ms\n",
                passes, nseg, acc[0], acc[1], acc[2], acc[3], acc[4], acc[5] / ew, acc[6] / ew, acc[7] / ew, acc[10] / ew, acc[8] / ew, acc[9] / ew,
                tsweep.ms());
    }
    return 0;
}

}  // namespace

// flavour: 0 auto, 1 streaming kernel 3 passes, 2 resident MH=1 3 passes, 3 resident 3 passes (MH auto),
//          4 one pass (resident when d allows, else streaming), 5 one pass streaming
int knn_tc_launch(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1, int64_t* d_ids,
                  float* d_dist, cudaStream_t stream, int flavour) {
    using namespace tc;
    const int64_t nq = q1 - q0;
    knn_clear_overflow_rows();
    if (nq <= 0) return 0;
    KnnJob J{};
    J.x = d_x; J.n = n; J.d = d; J.k = k; J.metric = metric; J.l2_squared = l2_squared; J.stream = stream;
    J.centre = metric == METRIC_L2 ? 1 : 0;
    J.sms = ctx().num_sms;
    J.dp = (d + 63) / 64 * 64;
    J.np = (n + 255) / 256 * 256;
    J.narrow_norm = tune("SPHB_KNN_NORM_WIDE", 0) == 0;
    const int dp = J.dp, sms = J.sms;

    // ---- which kernel: passes and flavour
    int passes_first = (flavour >= 1 && flavour <= 3) ? 3 : 1;
    if (flavour == 0 && tune("SPHB_KNN_PASSES", 1) == 3) passes_first = 3;
    auto pick_mh = [&](int passes) {
        if (flavour == 1 || flavour == 5) return 0;
        if (passes == 3) return dp <= 128 ? (flavour == 2 ? 1 : 2) : (dp <= 256 ? 1 : 0);
        return dp <= 256 ? 2 : (dp <= 512 ? 1 : 0);
    };

    EventTimer tprep;
    tprep.start(stream);
    SPHB_CUDA(J.colsum.alloc(d, stream));
    SPHB_CUDA(J.un2.alloc(n, stream));
    SPHB_CUDA(J.scal.alloc(2, stream));
    SPHB_CUDA(J.maxbits.alloc(1, stream));
    SPHB_CUDA(J.stats.alloc(3, stream));
    SPHB_CUDA(J.dun.alloc(n, stream));
    SPHB_CUDA(J.dvmax.alloc(1, stream));
    SPHB_CUDA(J.ones.alloc(128 * 64, stream));
    SPHB_CUDA(cudaMemsetAsync(J.colsum.p, 0, d * sizeof(double), stream));
    SPHB_CUDA(cudaMemsetAsync(J.maxbits.p, 0, sizeof(unsigned long long), stream));
    SPHB_CUDA(cudaMemsetAsync(J.dvmax.p, 0, sizeof(unsigned int), stream));
    if (J.centre) {
        dim3 g(static_cast<unsigned>(std::min<int64_t>(n, sms * 8)), static_cast<unsigned>((d + 127) / 128));
        knn_prep_colsum<<<g, 128, 0, stream>>>(d_x, n, d, J.colsum.p);
        count_launch();
    }
    knn_prep_norms<<<sms * 8, 256, 0, stream>>>(d_x, n, d, J.colsum.p, J.centre, J.un2.p, J.maxbits.p);
    knn_prep_scale<<<1, 1, 0, stream>>>(J.maxbits.p, J.scal.p);
    knn_prep_rescale<<<sms * 4, 256, 0, stream>>>(J.un2.p, n, J.scal.p);
    knn_prep_ones<<<32, 256, 0, stream>>>(J.ones.p);
    count_launch(4);
    SPHB_CUDA(cudaGetLastError());

    Operands op1, op3;  // full operand sets (A and B) for 1 pass / 3 passes, built on demand
    // need_a = false: only the database side (B, norms) is wanted — the re-sweep of a row list brings its own gathered A
    auto ensure_full = [&](Operands& op, int passes, bool need_a = true) -> int {
        if (op.built_b && (op.A.p || !need_a)) return 0;
        if (op.built_b) {  // B exists, A does not: build A alone
            Operands tmp;
            if (int rc = build_operands(J, tmp, passes, pick_mh(passes), nullptr, 0, true, false)) return rc;
            op.A.p = tmp.A.p; op.A.s = tmp.A.s; op.A.n = tmp.A.n; op.a_rows = tmp.a_rows;
            tmp.A.p = nullptr;
            return 0;
        }
        return build_operands(J, op, passes, pick_mh(passes), nullptr, 0, need_a, true);
    };
    if (int rc = ensure_full(passes_first == 1 ? op1 : op3, passes_first)) return rc;
    tprep.stop(stream);

    DevBuf<int> ovf_count;
    DevBuf<int64_t> ovf_rows;
    SPHB_CUDA(ovf_count.alloc(1, stream));
    SPHB_CUDA(ovf_rows.alloc(nq, stream));

    float sweep_ms = 0.f, rerank_ms = 0.f, fallback_ms = 0.f;
    int64_t resweep_rows = 0, simt_rows = 0, rows_1pass = 0;
    int n_qtiles_total = 0;

    // Sweeps [qa, qb) with `passes`.  Overflow rows of 1-pass sweeps accumulate in ovf_rows (all 1-pass sweeps of the call
    // share the list) and are re-swept ONCE with 3 passes at the end — a re-sweep costs one full pass over the database per
    // CTA however few rows it has; what overflows the 3-pass band (massive ties) goes to the exact SIMT kernel — never a
    // truncated answer.
    int pending1 = 0;  // overflow rows of the 1-pass sweeps so far
    SPHB_CUDA(cudaMemsetAsync(ovf_count.p, 0, sizeof(int), stream));
    auto simt_rows_exact = [&](const int64_t* rows_p, const int* count_p, int nrows) -> int {
        if (nrows <= 0) return 0;
        if (int rc = knn_simt_launch(d_x, n, d, k, metric, l2_squared, q0, q1, rows_p, count_p, nrows, d_ids, d_dist, stream)) return rc;
        SPHB_CUDA(cudaStreamSynchronize(stream));
        simt_rows += nrows;
        return 0;
    };
    auto process = [&](int64_t qa, int64_t qb, int passes, double* ovf_frac) -> int {
        Operands& op = passes == 1 ? op1 : op3;
        if (int rc = ensure_full(op, passes)) return rc;
        const int rpc = op.MH ? op.MH * BM : BM;
        n_qtiles_total += static_cast<int>((qb - qa + rpc - 1) / rpc);
        if (passes == 1) {
            int total = 0;
            if (int rc = sweep_and_rerank(J, op, op, qa, qb, nullptr, q0, d_ids, d_dist, ovf_rows.p, ovf_count.p, &total, &sweep_ms, &rerank_ms)) return rc;
            rows_1pass += qb - qa;
            if (ovf_frac) *ovf_frac = static_cast<double>(total - pending1) / static_cast<double>(qb - qa);
            pending1 = total;
            return 0;
        }
        DevBuf<int> ovf3_count;
        DevBuf<int64_t> ovf3_rows;
        SPHB_CUDA(ovf3_count.alloc(1, stream));
        SPHB_CUDA(ovf3_rows.alloc(qb - qa, stream));
        SPHB_CUDA(cudaMemsetAsync(ovf3_count.p, 0, sizeof(int), stream));
        int hovf = 0;
        if (int rc = sweep_and_rerank(J, op, op, qa, qb, nullptr, q0, d_ids, d_dist, ovf3_rows.p, ovf3_count.p, &hovf, &sweep_ms, &rerank_ms)) return rc;
        if (ovf_frac) *ovf_frac = static_cast<double>(hovf) / static_cast<double>(qb - qa);
        if (hovf == 0) return 0;
        if (int rc = record_overflow_rows(1, ovf3_rows.p, hovf, stream)) return rc;
        EventTimer tf;
        tf.start(stream);
        if (int rc = simt_rows_exact(ovf3_rows.p, ovf3_count.p, hovf)) return rc;
        tf.stop(stream);
        fallback_ms += tf.ms();
        return 0;
    };
    auto resolve_pending = [&]() -> int {
        if (pending1 == 0) return 0;
        EventTimer tf;
        tf.start(stream);
        const bool dbg = tune("SPHB_KNN_TIMING", 0) != 0;
        auto now = [&]() { if (dbg) cudaStreamSynchronize(stream); return std::chrono::steady_clock::now(); };
        auto ms_since = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
            return std::chrono::duration<double, std::milli>(b - a).count();
        };
        const auto t0 = now();
        bool recorded = false;  // the rows of this call's re-sweep are reported once (sphb_knn_last_overflow_rows), whatever stages they pass
        // First stage: the overflow rows once more through the 1-PRODUCT filter, database split into segments with a list
        // per (row, segment) — the lists of a row together hold nseg times the candidates of the main sweep at a third of the
        // 3-product sweep's tensor work.  What still overflows goes on to the 3-product stage below.
        if (tune("SPHB_KNN_RESWEEP_1PASS", 1) && op1.passes == 1 && op1.MH) {
            const int mh1 = (op1.MH == 2 && (pending1 + BM - 1) / BM <= J.sms) ? 1 : op1.MH;
            Operands sub1;
            if (int rc = build_operands(J, sub1, 1, mh1, ovf_rows.p, pending1, true, false)) return rc;
            DevBuf<int> ovfb_count;
            DevBuf<int64_t> ovfb_rows;
            SPHB_CUDA(ovfb_count.alloc(1, stream));
            SPHB_CUDA(ovfb_rows.alloc(pending1, stream));
            SPHB_CUDA(cudaMemsetAsync(ovfb_count.p, 0, sizeof(int), stream));
            float sm1 = 0.f, rm1 = 0.f;
            int hovfb = 0;
            if (int rc = sweep_and_rerank(J, sub1, op1, 0, pending1, ovf_rows.p, q0, d_ids, d_dist, ovfb_rows.p, ovfb_count.p, &hovfb, &sm1, &rm1)) return rc;
            if (dbg) fprintf(stderr, "[knn timing] 1-product re-sweep of %d rows: %d still overflow (kernel %.1f + %.1f ms)\n", pending1, hovfb, sm1, rm1);
            if (int rc = record_overflow_rows(0, ovf_rows.p, pending1, stream)) return rc;
            recorded = true;
            resweep_rows += pending1;
            if (hovfb == 0) {
                tf.stop(stream);
                fallback_ms += tf.ms();
                pending1 = 0;
                return 0;
            }
            SPHB_CUDA(cudaMemcpyAsync(ovf_rows.p, ovfb_rows.p, static_cast<size_t>(hovfb) * sizeof(int64_t), cudaMemcpyDeviceToDevice, stream));
            pending1 = hovfb;
            resweep_rows -= pending1;  // counted again below
        }
        if (int rc = ensure_full(op3, 3, false)) return rc;
        const auto t1 = now();
        // few rows: 128-row query tiles (MH = 1) put twice as many CTAs to work, each with half the MMA work per database tile
        const int mh_re = (op3.MH == 2 && (pending1 + BM - 1) / BM <= J.sms) ? 1 : op3.MH;
        Operands sub;
        if (int rc = build_operands(J, sub, 3, mh_re, ovf_rows.p, pending1, true, false)) return rc;
        DevBuf<int> ovf2_count;
        DevBuf<int64_t> ovf2_rows;
        SPHB_CUDA(ovf2_count.alloc(1, stream));
        SPHB_CUDA(ovf2_rows.alloc(pending1, stream));
        SPHB_CUDA(cudaMemsetAsync(ovf2_count.p, 0, sizeof(int), stream));
        float sm = 0.f, rm = 0.f;
        int hovf2 = 0;
        const auto t2 = now();
        if (int rc = sweep_and_rerank(J, sub, op3, 0, pending1, ovf_rows.p, q0, d_ids, d_dist, ovf2_rows.p, ovf2_count.p, &hovf2, &sm, &rm)) return rc;
        const auto t3 = now();
        resweep_rows += pending1;
        if (!recorded)
            if (int rc = record_overflow_rows(0, ovf_rows.p, pending1, stream)) return rc;
        if (int rc = record_overflow_rows(1, ovf2_rows.p, hovf2, stream)) return rc;
        if (int rc = simt_rows_exact(ovf2_rows.p, ovf2_count.p, hovf2)) return rc;
        if (dbg)
            fprintf(stderr, "[knn timing] re-sweep of %d rows: 3-pass operands %.1f ms, gathered A %.1f ms, sweep+rerank %.1f ms (kernel %.1f + %.1f), simt %.1f ms\n",
                    pending1, ms_since(t0, t1), ms_since(t1, t2), ms_since(t2, t3), sm, rm, ms_since(t3, now()));
        tf.stop(stream);
        fallback_ms += tf.ms();
        pending1 = 0;
        return 0;
    };

    if (passes_first == 1 && flavour == 0) {
        // pilot: one wave of query tiles with the 1-pass filter; if its band overflows for too many rows
        // (distance-concentrated data) the rest of the slab uses the split-precision filter directly.
        const int rpc = (pick_mh(1) ? pick_mh(1) * BM : BM);
        const int64_t pilot = static_cast<int64_t>(sms) * rpc;
        if (nq > 3 * pilot) {
            double frac = 0.0;
            if (int rc = process(q0, q0 + pilot, 1, &frac)) return rc;
            if (int rc = process(q0 + pilot, q1, frac > 0.25 ? 3 : 1, nullptr)) return rc;
        } else {
            if (int rc = process(q0, q1, 1, nullptr)) return rc;
        }
    } else {
        if (int rc = process(q0, q1, passes_first, nullptr)) return rc;
    }
    if (int rc = resolve_pending()) return rc;

    g_knn_ms[0] = tprep.ms();
    g_knn_ms[1] = sweep_ms;
    g_knn_ms[2] = rerank_ms;
    g_knn_ms[3] = fallback_ms;
    const Operands& opm = rows_1pass > 0 ? op1 : op3;
    const int kexec = opm.MH ? ((opm.passes == 3 ? 3 : 1) * dp + 16) : opm.lda;
    const int64_t st[8] = {1 + opm.MH, J.sum_compactions, J.sum_accepted, kexec, simt_rows, n_qtiles_total, rows_1pass > 0 ? 1 : 3, resweep_rows};
    knn_set_stats(st);
    return 0;
}

}  // namespace sphb
