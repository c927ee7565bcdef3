// Epilogue shared by the two tcgen05 sweep kernels: threshold filter on the fp32
// accumulator tile (TMEM -> registers) and the per-row approximate candidate list.
//
// Each epilogue thread owns one query row (= one TMEM lane).  A database point j
// survives the filter iff its GEMM value G_ij <= thr_i.  Survivors are appended to
// the row's list in L2-resident scratch as (G, j) keys.  When a list is nearly full
// the warp sorts it by (G, j), reads the k-th smallest value G_(k) and keeps every
// entry with G <= G_(k) + 2*eps_i; that limit becomes the new threshold.  Because
// |G - G_true| <= eps_i (error bound), every point whose TRUE
// distance is among the k smallest is still in the list at the end of the sweep;
// the exact fp64 re-rank (knn_rerank_kernel) then picks the final k.  If more than
// keep_max entries fall inside the band (massive ties) the row is flagged and is
// recomputed by the exact SIMT kernel instead — never silently truncated.
//
// Two error models:
//   passes = 3  split-fp16 operands (hi·hi + hi·lo + lo·hi): eps_i is a per-row constant.
//   passes = 1  single fp16 product: the operand rounding error is bounded with the ACTUAL
//               rounding-error norms |u_i - fp16(u_i)| computed by the prep kernel, and with an
//               adaptive radius R_i >= |v_j| for every database point that can matter, derived from
//               the row's current k-th value (band_1pass below).
#pragma once
#include "knn_common.cuh"
#include "ptx_sm100.cuh"

// Cycle-counter instrumentation of the sweep (build with -DSPHB_KNN_PROFILE, run with SPHB_KNN_PROF=1).
#ifdef SPHB_KNN_PROFILE
#define SPHB_CLOCK() clock64()
#else
#define SPHB_CLOCK() 0ll
#endif

namespace sphb {

constexpr int OVERFLOW_FLAG = 1 << 30;

struct RowList {
    float thr;   // accept iff G <= thr
    float eps2;  // passes = 3: 2 * eps_i, rounded up
    float un;    // passes = 1: |u_i| (scaled, centred), rounded up
    float dun;   // passes = 1: |u_i - fp16(u_i)|, rounded up
    int cnt;     // entries currently in the row's scratch list
    int ovf;     // overflow flag
    int trig;    // compact when cnt exceeds this: min(cap - 64, survivors of the last compaction + BandConsts::step)
};

// Constants of the band computation, identical for every row of a launch.
struct BandConsts {
    int passes;    // 1 or 3
    int l2;        // 1: L2 metric (the adaptive radius applies), 0: inner product
    float vmax;    // upper bound on every |v_j| (scaled)
    float dvmax;   // max_j |v_j - fp16(v_j)|, rounded up
    float acc_c;   // accumulation-error coefficient (relative to 2|u||v| + |v|²)
    float sub_c;   // fp16 subnormal floor of a row's rounding-error norm: sqrt(d) * 2^-25
    int step;      // appends between two compactions of a row (the threshold is stale in between: a short step means
                   // fewer accepted candidates and cheaper, half-width compactions, but more of them)
};

// passes = 1.  Error of one computed G_ij against the true value, for |v_j| <= R:
//   2 (|du_i| |fp16(v_j)| + |u_i| |dv_j|) + accumulation, with |dv_j| <= min(dvmax, 2^-11 R + sub_c).
__device__ __forceinline__ float pair_err_1pass(float R, float un, float dun, const BandConsts& bc) {
    const float Rh = R * 1.0005f;  // |fp16(v)| <= |v| (1 + 2^-11)
    const float dv = fminf(bc.dvmax, 0.00048877f * R + bc.sub_c);
    return 2.f * (dun * Rh + un * dv) + bc.acc_c * (2.f * un * Rh + Rh * Rh);
}
// Width of the band above the current k-th value Gk that must be kept so that no true neighbour is
// lost: E(j') + E(j*) for the two database points of the proof in DESIGN.md, both of which have
// |v| <= |u_i| + sqrt(Gk + Emax + |u_i|²) (L2 only; for the inner product R = vmax).
__device__ __forceinline__ float band_1pass(float Gk, float un, float dun, const BandConsts& bc) {
    float R = bc.vmax;
    if (bc.l2) {
        const float emax = pair_err_1pass(R, un, dun, bc);
        const float D = fmaxf(Gk + emax + un * un * 1.0001f, 0.f) * 1.0001f;
        R = fminf(R, (un + sqrtf(D)) * 1.0005f);
    }
    return 2.f * pair_err_1pass(R, un, dun, bc) * 1.001f;
}

// Ascending bitonic sort of 32*R u64 keys held in registers, element e = lane*R + r.
// Partner distances below R are register swaps, the others one 64-bit shuffle per key.
// The stage loops are real loops (only the per-register work is unrolled): a fully unrolled
// network is ~3000 instructions per call site and thrashes the instruction cache.
template <int R, int J2>
__device__ __forceinline__ void bitonic_inlane_stage(uint64_t (&x)[R], int lane, int k2) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if ((r & J2) == 0) {
            const bool up = (((lane * R + r) & k2) == 0);
            const uint64_t a = x[r], b = x[r | J2];
            const bool sw = (a > b) == up;
            x[r] = sw ? b : a;
            x[r | J2] = sw ? a : b;
        }
    }
}
template <int R, int J2>
__device__ __forceinline__ void bitonic_inlane_from(uint64_t (&x)[R], int lane, int k2) {
    // stages j2 = J2, J2/2, ..., 1, skipping those with j2 >= k2
    if constexpr (J2 >= 1) {
        if (J2 < k2) bitonic_inlane_stage<R, J2>(x, lane, k2);
        bitonic_inlane_from<R, J2 / 2>(x, lane, k2);
    }
}
template <int R>
__device__ __forceinline__ void warp_bitonic_sort_regs(uint64_t (&x)[R], int lane) {
#pragma unroll 1
    for (int k2 = 2; k2 <= 32 * R; k2 <<= 1) {
#pragma unroll 1
        for (int lj = (k2 >> 1) / R; lj > 0; lj >>= 1) {  // cross-lane stages, partner lane = lane ^ lj
            const bool lower = (lane & lj) == 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint64_t other = __shfl_xor_sync(0xffffffffu, x[r], lj);
                const bool up = (((lane * R + r) & k2) == 0);
                const bool keep_min = (lower == up);
                const uint64_t mn = x[r] < other ? x[r] : other;
                const uint64_t mx = x[r] < other ? other : x[r];
                x[r] = keep_min ? mn : mx;
            }
        }
        bitonic_inlane_from<R, R / 2>(x, lane, k2);
    }
}

// Sort + band-truncate the list of lane L's row (whole warp participates); the sorted keys stay in
// x[] (element e = lane*R + r) for the caller.  cap = 32*R.
template <int R>
__device__ __forceinline__ void compact_row_approx(RowList& me, int L, uint64_t* __restrict__ buf_L, uint64_t (&x)[R], int k, int keep_max,
                                                   int lane, const BandConsts& bc) {
    const int cntL = __shfl_sync(0xffffffffu, me.cnt, L);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int e = lane * R + r;
        x[r] = (e < cntL) ? buf_L[e] : ~0ull;
    }
    warp_bitonic_sort_regs<R>(x, lane);
    int m = cntL;
    float limit = INFINITY;
    int ovf = 0;
    if (cntL >= k) {
        // k-th smallest key lives in lane (k-1)/R, register (k-1)%R
        uint64_t kth = 0;
        const int kr = (k - 1) % R;
#pragma unroll
        for (int r = 0; r < R; ++r) kth = (r == kr) ? x[r] : kth;
        kth = __shfl_sync(0xffffffffu, kth, (k - 1) / R);
        // every lane evaluates the band with its own row constants; lane L's result is the one used
        const float band = bc.passes == 3 ? me.eps2 : band_1pass(key_val(kth), me.un, me.dun, bc);
        limit = __fadd_ru(key_val(kth), __shfl_sync(0xffffffffu, band, L));
        int c = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) c += (lane * R + r < cntL && key_val(x[r]) <= limit) ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        m = c;
        if (m > keep_max) {
            m = keep_max;
            ovf = 1;
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int e = lane * R + r;
        if (e < m) buf_L[e] = x[r];
    }
    if (lane == L) {
        me.cnt = m;
        me.thr = limit;
        me.ovf |= ovf;
        me.trig = min(32 * R - 64, m + bc.step);
    }
    __syncwarp();
}

// ---- in-sweep compaction: selection on the 32-bit values only --------------------------------------------
// What a compaction during the sweep has to deliver is the row's new threshold (k-th smallest value + band) and the
// list without the entries above it; the order of the survivors is irrelevant (export_rows sorts once at the end).
// So only the ordered 32-bit value halves are sorted — one shuffle and a min/max per element and stage instead of
// a 64-bit compare-exchange — and the u64 entries, kept in registers, are stream-compacted in place.
template <int R, int J2>
__device__ __forceinline__ void bitonic32_inlane_stage(uint32_t (&x)[R], int lane, int k2) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if ((r & J2) == 0) {
            const bool up = (((lane * R + r) & k2) == 0);
            const uint32_t mn = min(x[r], x[r | J2]), mx = max(x[r], x[r | J2]);
            x[r] = up ? mn : mx;
            x[r | J2] = up ? mx : mn;
        }
    }
}
template <int R, int J2>
__device__ __forceinline__ void bitonic32_inlane_from(uint32_t (&x)[R], int lane, int k2) {
    if constexpr (J2 >= 1) {
        if (J2 < k2) bitonic32_inlane_stage<R, J2>(x, lane, k2);
        bitonic32_inlane_from<R, J2 / 2>(x, lane, k2);
    }
}
template <int R>
__device__ __forceinline__ void warp_bitonic_sort_regs32(uint32_t (&x)[R], int lane) {
#pragma unroll 1
    for (int k2 = 2; k2 <= 32 * R; k2 <<= 1) {
#pragma unroll 1
        for (int lj = (k2 >> 1) / R; lj > 0; lj >>= 1) {
            const bool lower = (lane & lj) == 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const uint32_t other = __shfl_xor_sync(0xffffffffu, x[r], lj);
                const bool up = (((lane * R + r) & k2) == 0);
                x[r] = (lower == up) ? min(x[r], other) : max(x[r], other);
            }
        }
        bitonic32_inlane_from<R, R / 2>(x, lane, k2);
    }
}

template <int R>
__device__ __forceinline__ void compact_row_select_n(RowList& me, int L, uint64_t* __restrict__ buf_L, int k, int keep_max, int cap, int lane,
                                                     const BandConsts& bc) {
    const int cntL = __shfl_sync(0xffffffffu, me.cnt, L);
    uint64_t x[R];
    uint32_t v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int e = r * 32 + lane;  // coalesced
        x[r] = (e < cntL) ? buf_L[e] : ~0ull;
        v[r] = static_cast<uint32_t>(x[r] >> 32);
    }
    int m = cntL;
    float limit = INFINITY;
    int ovf = 0;
    if (cntL >= k) {
        warp_bitonic_sort_regs32<R>(v, lane);  // as a multiset; sorted position p = lane*R + r
        uint32_t kth = 0;
        const int kr = (k - 1) % R;
#pragma unroll
        for (int r = 0; r < R; ++r) kth = (r == kr) ? v[r] : kth;
        kth = __shfl_sync(0xffffffffu, kth, (k - 1) / R);
        const float gk = ord2f(kth);
        const float band = bc.passes == 3 ? me.eps2 : band_1pass(gk, me.un, me.dun, bc);
        limit = __fadd_ru(gk, __shfl_sync(0xffffffffu, band, L));
        const uint32_t lim = f2ord(limit);
        const unsigned lt = (1u << lane) - 1u;
        int base = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool keep = (r * 32 + lane < cntL) && static_cast<uint32_t>(x[r] >> 32) <= lim;
            const unsigned mk = __ballot_sync(0xffffffffu, keep);
            const int pos = base + __popc(mk & lt);
            if (keep && pos < keep_max) buf_L[pos] = x[r];  // pos <= r*32 + lane: never ahead of an unread slot (all reads are done)
            base += __popc(mk);
        }
        m = base;
        if (m > keep_max) {
            m = keep_max;
            ovf = 1;
        }
    }
    if (lane == L) {
        me.cnt = m;
        me.thr = limit;
        me.ovf |= ovf;
        me.trig = min(cap - 64, m + bc.step);
    }
    __syncwarp();
}
// ---- in-sweep compaction, round 2: selection by bisection on the ordered 32-bit values ---------------------
// The sorting network above costs ~2.5 k dependent warp instructions (10.6 k cycles measured) per compaction; with one
// compaction per ~20 database tiles somewhere in the CTA and only two accumulator stages per M half to absorb it, those
// cycles were the MMA issuer's whole wait for a free accumulator (26 % of the sweep, profiles/r02_sweep_prof.txt).
// What a compaction needs is ANY upper bound T of the row's k-th smallest value (thr = T + band(T) then still keeps every
// true neighbour; a looser T only keeps a few more candidates): bisect the ordered-u32 value range, counting the
// list's values <= pivot with one compare per register and a REDUX per probe, and stop as soon as the count lies in
// [k, k + 8].  ~15 probes of ~40 independent instructions instead of a 45-stage network.
template <int R>
__device__ __forceinline__ void compact_row_bisect_n(RowList& me, int L, uint64_t* __restrict__ buf_L, int k, int keep_max, int cap, int lane,
                                                     const BandConsts& bc) {
    const int cntL = __shfl_sync(0xffffffffu, me.cnt, L);
    uint64_t x[R];
    uint32_t v[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int e = r * 32 + lane;  // coalesced
        x[r] = (e < cntL) ? buf_L[e] : ~0ull;
        v[r] = static_cast<uint32_t>(x[r] >> 32);
    }
    int m = cntL;
    float limit = INFINITY;
    int ovf = 0;
    if (cntL >= k) {
        uint32_t lo = 0xffffffffu, hi = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            lo = min(lo, v[r]);
            hi = max(hi, (r * 32 + lane < cntL) ? v[r] : 0u);
        }
        lo = __reduce_min_sync(0xffffffffu, lo);
        hi = __reduce_max_sync(0xffffffffu, hi);  // count(v <= hi) = cntL >= k: the invariant of the search
        while (lo < hi) {
            const uint32_t mid = lo + ((hi - lo) >> 1);
            int c = 0;
#pragma unroll
            for (int r = 0; r < R; ++r) c += (v[r] <= mid) ? 1 : 0;  // padding is 0xffffffff > mid (mid < hi)
            c = __reduce_add_sync(0xffffffffu, c);
            if (c >= k) {
                hi = mid;
                if (c <= k + 8) break;
            } else {
                lo = mid + 1;
            }
        }
        const float gk = ord2f(hi);  // >= the k-th smallest value of the list
        const float band = bc.passes == 3 ? me.eps2 : band_1pass(gk, me.un, me.dun, bc);
        limit = __fadd_ru(gk, __shfl_sync(0xffffffffu, band, L));
        const uint32_t lim = f2ord(limit);
        const unsigned lt = (1u << lane) - 1u;
        int base = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool keep = (r * 32 + lane < cntL) && v[r] <= lim;
            const unsigned mk = __ballot_sync(0xffffffffu, keep);
            const int pos = base + __popc(mk & lt);
            if (keep && pos < keep_max) buf_L[pos] = x[r];  // pos <= r*32 + lane: never ahead of an unread slot (all reads are done)
            base += __popc(mk);
        }
        m = base;
        if (m > keep_max) {
            m = keep_max;
            ovf = 1;
        }
    }
    if (lane == L) {
        me.cnt = m;
        me.thr = limit;
        me.ovf |= ovf;
        me.trig = min(cap - 64, m + bc.step);
    }
    __syncwarp();
}

// lists of at most half the capacity take the half-width network
template <int R>
__device__ __forceinline__ void compact_row_select(RowList& me, int L, uint64_t* __restrict__ buf_L, int k, int keep_max, int lane,
                                                   const BandConsts& bc) {
    const int cntL = __shfl_sync(0xffffffffu, me.cnt, L);
#ifdef SPHB_COMPACT_SORT_SELECT
    if constexpr (R >= 8) {
        if (cntL <= 16 * R) {
            compact_row_select_n<R / 2>(me, L, buf_L, k, keep_max, 32 * R, lane, bc);
            return;
        }
    }
    compact_row_select_n<R>(me, L, buf_L, k, keep_max, 32 * R, lane, bc);
#else
    if constexpr (R >= 8) {
        if (cntL <= 16 * R) {
            compact_row_bisect_n<R / 2>(me, L, buf_L, k, keep_max, 32 * R, lane, bc);
            return;
        }
    }
    compact_row_bisect_n<R>(me, L, buf_L, k, keep_max, 32 * R, lane, bc);
#endif
}

// Filter one 32-column chunk of the accumulator tile held in v[] (columns jbase..jbase+31).
// Fast path: 8 group minima (4 columns each) + their minimum, one compare.  Slow path (some
// column passes): only the groups whose minimum passes are scanned.
__device__ __forceinline__ void filter_chunk(const float (&v)[32], uint32_t jbase, uint32_t n, RowList& me, uint64_t* __restrict__ mybuf,
                                             unsigned long long& n_accept) {
    float g[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) g[q] = fminf(fminf(v[4 * q], v[4 * q + 1]), fminf(v[4 * q + 2], v[4 * q + 3]));
    const float m = fminf(fminf(fminf(g[0], g[1]), fminf(g[2], g[3])), fminf(fminf(g[4], g[5]), fminf(g[6], g[7])));
    if (m <= me.thr) {
        const float thr = me.thr;
        int cnt = me.cnt;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            if (g[q] <= thr) {
#pragma unroll
                for (int t = 4 * q; t < 4 * q + 4; ++t) {
                    const uint32_t j = jbase + t;
                    if (v[t] <= thr && j < n) mybuf[cnt++] = make_key(v[t], j);
                }
            }
        }
        n_accept += static_cast<unsigned long long>(cnt - me.cnt);
        me.cnt = cnt;
    }
}

// Two 32-column chunks at once: the two min trees are independent (twice the ILP of two filter_chunk
// calls) and there is one branch per 64 columns.  Appends at most 64 keys.
__device__ __forceinline__ void filter_pair(const float (&v0)[32], const float (&v1)[32], uint32_t jbase, uint32_t n, RowList& me,
                                            uint64_t* __restrict__ mybuf, unsigned long long& n_accept) {
    float g0[8], g1[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        g0[q] = fminf(fminf(v0[4 * q], v0[4 * q + 1]), fminf(v0[4 * q + 2], v0[4 * q + 3]));
        g1[q] = fminf(fminf(v1[4 * q], v1[4 * q + 1]), fminf(v1[4 * q + 2], v1[4 * q + 3]));
    }
    const float m0 = fminf(fminf(fminf(g0[0], g0[1]), fminf(g0[2], g0[3])), fminf(fminf(g0[4], g0[5]), fminf(g0[6], g0[7])));
    const float m1 = fminf(fminf(fminf(g1[0], g1[1]), fminf(g1[2], g1[3])), fminf(fminf(g1[4], g1[5]), fminf(g1[6], g1[7])));
    if (fminf(m0, m1) <= me.thr) {
        const float thr = me.thr;
        int cnt = me.cnt;
        // usually one lane passes with one value: only the 32-column half (and 4-column group) that holds it is scanned
        if (m0 <= thr) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (g0[q] <= thr) {
#pragma unroll
                    for (int t = 4 * q; t < 4 * q + 4; ++t) {
                        const uint32_t j = jbase + t;
                        if (v0[t] <= thr && j < n) mybuf[cnt++] = make_key(v0[t], j);
                    }
                }
            }
        }
        if (m1 <= thr) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                if (g1[q] <= thr) {
#pragma unroll
                    for (int t = 4 * q; t < 4 * q + 4; ++t) {
                        const uint32_t j = jbase + 32 + t;
                        if (v1[t] <= thr && j < n) mybuf[cnt++] = make_key(v1[t], j);
                    }
                }
            }
        }
        n_accept += static_cast<unsigned long long>(cnt - me.cnt);
        me.cnt = cnt;
    }
}

// After filter_chunk: compact every row of this warp whose list is close to full.
template <int R>
__device__ __forceinline__ void compact_full_rows(RowList& me, uint64_t* __restrict__ warp_buf, int k, int keep_max, int lane,
                                                  const BandConsts& bc, unsigned long long& n_compact, long long& t_compact) {
    constexpr int cap = 32 * R;
    unsigned full = __ballot_sync(0xffffffffu, me.cnt > me.trig);  // trig <= cap - 64: room for one more filter_pair (<= 64 appends)
    if (full == 0) return;
    const long long c0 = SPHB_CLOCK();
    __syncwarp();  // the owner lanes' list appends must be visible to the whole warp
    while (full) {
        const int L = __ffs(full) - 1;
        full &= full - 1;
#ifdef SPHB_COMPACT_FULL_SORT
        uint64_t x[R];
        compact_row_approx<R>(me, L, warp_buf + static_cast<size_t>(L) * cap, x, k, keep_max, lane, bc);
#else
        compact_row_select<R>(me, L, warp_buf + static_cast<size_t>(L) * cap, k, keep_max, lane, bc);
#endif
        ++n_compact;
    }
    t_compact += SPHB_CLOCK() - c0;
}

// End of sweep: final compaction of the warp's 32 rows and export of the candidate ids.
//   r0: slab-row of lane 0; list of row r goes to slot r * out_mul + out_add of cand_out [slots, keep_max] uint32 and
//   cnt_out [slots] (bit 30 = overflow) — out_mul = database segments per row, out_add = this item's segment.
template <int R>
__device__ __forceinline__ void export_rows(RowList& me, int64_t r0, int64_t nq, uint64_t* __restrict__ warp_buf, int k, int keep_max,
                                            uint32_t* __restrict__ cand_out, int* __restrict__ cnt_out, int lane, const BandConsts& bc,
                                            int out_mul = 1, int out_add = 0) {
    constexpr int cap = 32 * R;
    __syncwarp();
    for (int L = 0; L < 32; ++L) {
        const int64_t rL = r0 + L;
        if (rL >= nq) break;
        uint64_t x[R];
        compact_row_approx<R>(me, L, warp_buf + static_cast<size_t>(L) * cap, x, k, keep_max, lane, bc);
        const int m = __shfl_sync(0xffffffffu, me.cnt, L);
        const int ovf = __shfl_sync(0xffffffffu, me.ovf, L);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int e = lane * R + r;
            if (e < m) cand_out[(rL * out_mul + out_add) * keep_max + e] = key_idx(x[r]);
        }
        if (lane == 0) cnt_out[rL * out_mul + out_add] = m | (ovf ? OVERFLOW_FLAG : 0);
        __syncwarp();
    }
}

// eps_i in GEMM units: c1*(2|u_i|VMAX + VMAX²) + c2*(|u_i| + VMAX)
__device__ __forceinline__ float row_eps2(double un2, double c1, double c2, double vmax, double vmax2) {
    const double un = sqrt(un2);
    const double e = c1 * (2.0 * un * vmax + vmax2) + c2 * (un + vmax);
    return __double2float_ru(2.0 * e);
}

}  // namespace sphb
