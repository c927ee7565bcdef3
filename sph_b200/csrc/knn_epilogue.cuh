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
#pragma once
#include "knn_common.cuh"
#include "ptx_sm100.cuh"

namespace sphb {

constexpr int OVERFLOW_FLAG = 1 << 30;

struct RowList {
    float thr;   // accept iff G <= thr
    float eps2;  // 2 * eps_i, rounded up
    int cnt;     // entries currently in the row's scratch list
    int ovf;     // overflow flag
};

// Sort + band-truncate the list of lane L's row (whole warp participates).
__device__ __forceinline__ void compact_row_approx(RowList& me, int L, uint64_t* __restrict__ buf_L, uint64_t* sm, int cap, int k, int keep_max,
                                                   int lane) {
    const int cntL = __shfl_sync(0xffffffffu, me.cnt, L);
    const float eps2L = __shfl_sync(0xffffffffu, me.eps2, L);
    for (int t = lane; t < cap; t += 32) sm[t] = (t < cntL) ? buf_L[t] : ~0ull;
    __syncwarp();
    warp_bitonic_sort(sm, cap, lane);
    int m = cntL;
    float limit = INFINITY;
    int ovf = 0;
    if (cntL >= k) {
        limit = __fadd_ru(key_val(sm[k - 1]), eps2L);
        int c = 0;
        for (int t = lane; t < cap; t += 32) c += (t < cntL && key_val(sm[t]) <= limit) ? 1 : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        m = c;
        if (m > keep_max) {
            m = keep_max;
            ovf = 1;
        }
    }
    for (int t = lane; t < m; t += 32) buf_L[t] = sm[t];
    if (lane == L) {
        me.cnt = m;
        me.thr = limit;
        me.ovf |= ovf;
    }
    __syncwarp();
}

// Filter one 32-column chunk of the accumulator tile held in v[] (columns jbase..jbase+31).
__device__ __forceinline__ void filter_chunk(const float (&v)[32], uint32_t jbase, int64_t n, RowList& me, uint64_t* __restrict__ mybuf,
                                             unsigned long long& n_accept) {
    float m = v[0];
#pragma unroll
    for (int t = 1; t < 32; ++t) m = fminf(m, v[t]);
    if (m <= me.thr) {
#pragma unroll
        for (int t = 0; t < 32; ++t) {
            if (v[t] <= me.thr) {
                const uint32_t j = jbase + t;
                if (j < n) {
                    mybuf[me.cnt++] = make_key(v[t], j);
                    ++n_accept;
                }
            }
        }
    }
}

// After filter_chunk: compact every row of this warp whose list is close to full.
__device__ __forceinline__ void compact_full_rows(RowList& me, uint64_t* __restrict__ warp_buf, uint64_t* sm, int cap, int k, int keep_max,
                                                  int lane, unsigned long long& n_compact) {
    unsigned full = __ballot_sync(0xffffffffu, me.cnt > cap - 32);
    while (full) {
        const int L = __ffs(full) - 1;
        full &= full - 1;
        compact_row_approx(me, L, warp_buf + static_cast<size_t>(L) * cap, sm, cap, k, keep_max, lane);
        ++n_compact;
    }
}

// End of sweep: final compaction of the warp's 32 rows and export of the candidate ids.
//   r0: slab-row of lane 0; cand_out [nq, keep_max] uint32, cnt_out [nq] (bit 30 = overflow).
__device__ __forceinline__ void export_rows(RowList& me, int64_t r0, int64_t nq, uint64_t* __restrict__ warp_buf, uint64_t* sm, int cap, int k,
                                            int keep_max, uint32_t* __restrict__ cand_out, int* __restrict__ cnt_out, int lane) {
    for (int L = 0; L < 32; ++L) {
        const int64_t rL = r0 + L;
        if (rL >= nq) break;
        compact_row_approx(me, L, warp_buf + static_cast<size_t>(L) * cap, sm, cap, k, keep_max, lane);
        const int m = __shfl_sync(0xffffffffu, me.cnt, L);
        const int ovf = __shfl_sync(0xffffffffu, me.ovf, L);
        for (int t = lane; t < m; t += 32) cand_out[rL * keep_max + t] = key_idx(sm[t]);
        if (lane == 0) cnt_out[rL] = m | (ovf ? OVERFLOW_FLAG : 0);
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
