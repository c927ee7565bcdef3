// Stage C, RW1: Monte-Carlo random-walk transition sums (replaces utils::doRandomWalks,
// sph/utils/SparseMatrixAlgorithms.cpp:34-290).
//
// One warp per start point.  Lane l runs walks l, l+32, ...; every step draws a counter-based
// uniform (Philox4x32-10 keyed by the seed, counter = (start, walk, step)), scans the current row in
// ascending column order accumulating fp64 probability until u < cum (stays put otherwise, :129-135)
// and records the event (column, walk*L + step) in the warp's shared-memory event list.  The list is
// then sorted by (column, event order) and reduced run by run IN EVENT ORDER with the step weights
// recomputed from the step index — the same fp32 addition order as the reference's
// `hits[next] += stepWeight` for the same paths, so sums are deterministic and layout independent.
// Row post-processing follows the reference: FIRST_VISIT transform (:176-218), prune (:221-233),
// removeDiagonalElements(keepSingleEntry) (:704-718), L1 normalisation (:253-267).
// The HBM traffic is the random row gathers (8 B x half a row per step); rows are written once.
#include "../../include/sphb.h"
#include "graph_common.cuh"

namespace sphb {

struct WalkParams {
    int64_t n;
    const int64_t* rowptr;
    const uint32_t* col;
    const float* val;
    const double* cum;  // SEARCH kernel: per-entry running fp64 sum of the row (sequential order), see walk_build_cum
    uint32_t W, L;
    int weighting;
    float prune_value;
    uint64_t seed;
    int remove_diagonal, normalize;
    int cap_ev;         // events per start: scan kernel power of two >= W*L; search kernel W*L rounded up to 32
    int col_bits, radix_bits;  // search kernel: bits of a column id, digit width of the radix passes
    int64_t start0, nstarts;
    int tmp_stride;     // entries per start in the temp buffers
    uint32_t* tmp_col;
    float* tmp_val;
    int* row_nnz;       // [nstarts]
};

__device__ __forceinline__ float step_weight(int weighting, uint32_t step, uint32_t L) {
    switch (weighting) {
        case SPHB_W_LINEAR: return static_cast<float>(1.0 - (static_cast<double>(step) / L));
        case SPHB_W_NORMAL: {
            const float index = step * 3.f / L;
            return expf(-0.5f * (index * index));
        }
        case SPHB_W_ONLYLAST: return (step < L - 1) ? 0.f : 1.f;
        case SPHB_W_FIRST_VISIT: return static_cast<float>(step + 1);
        default: return 1.f;
    }
}

// testing hook (sphb_debug_force_walk_scan, not in the public header): always run the serial-scan kernel
static int g_force_scan = 0;
void walks_force_scan(int on) { g_force_scan = on; }

constexpr int FV_MAX_LEN = 64;  // FIRST_VISIT keeps the visited set of a walk in registers/local memory

// Running sums of every row, fp64, accumulated in storage (ascending column) order exactly as the reference's scan does
// (`inc += val[e]`, SparseMatrixAlgorithms.cpp:129-135) — one thread per row, sequential.  With them a step is a binary
// search for the first entry with u < cum instead of a serial scan of half the row: the same entry for every u as long as
// the sums are non-decreasing, i.e. no value is negative (flagged otherwise; the scan kernel then runs).
__global__ void walk_build_cum(int64_t n, const int64_t* __restrict__ rowptr, const float* __restrict__ val, double* __restrict__ cum,
                               int* __restrict__ negative) {
    for (int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; r < n; r += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        double inc = 0.0;
        bool neg = false;
        for (int64_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
            const float v = val[e];
            neg |= v < 0.f;
            inc += static_cast<double>(v);
            cum[e] = inc;
        }
        if (neg) atomicExch(negative, 1);
    }
}

constexpr int WALK_CHAINS = 3;  // SEARCH kernel: independent walks per lane in flight (96 walks per warp pass; W = 90 fits one pass)

// The serial-scan kernel: lane = walk, every step scans the current row until u < inc.  Runs for matrices with negative
// values and for FIRST_VISIT weighting (which keeps a visited set per walk); everything else takes random_walk_search_kernel
// below.  Both draw the same Philox uniform per (start, walk, step) and pick the same entry, so their results are identical.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) random_walk_kernel(const WalkParams P) {
    extern __shared__ uint64_t smem_ev[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* ev = smem_ev + static_cast<size_t>(warp) * P.cap_ev;
    const uint32_t W = P.W, L = P.L;
    const uint32_t total = W * L;
    const uint2 key = make_uint2(static_cast<uint32_t>(P.seed), static_cast<uint32_t>(P.seed >> 32));
    for (int64_t s = static_cast<int64_t>(blockIdx.x) * WARPS + warp; s < P.nstarts; s += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t start = P.start0 + s;
        // ---- 1. the walks
        for (uint32_t w = lane; w < W; w += 32) {
            int64_t cur = start;
            uint32_t visited[FV_MAX_LEN + 1];
            int nvis = 0;
            if (P.weighting == SPHB_W_FIRST_VISIT) visited[nvis++] = static_cast<uint32_t>(cur);
            for (uint32_t step = 0; step < L; ++step) {
                const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(start), static_cast<uint32_t>(start >> 32), w, step), key);
                const double u = u01_from_bits(r.x, r.y);
                int64_t next = cur;
                double inc = 0.0;
                const int64_t e1 = P.rowptr[cur + 1];
                for (int64_t e = P.rowptr[cur]; e < e1; ++e) {
                    inc += static_cast<double>(__ldg(P.val + e));
                    if (u < inc) {
                        next = __ldg(P.col + e);
                        break;
                    }
                }
                bool record = true;
                if (P.weighting == SPHB_W_ONLYLAST) record = step == L - 1;
                if (P.weighting == SPHB_W_FIRST_VISIT) {
                    bool seen = false;
                    for (int t = 0; t < nvis; ++t) seen |= visited[t] == static_cast<uint32_t>(next);
                    record = !seen;
                    if (!seen && nvis <= FV_MAX_LEN) visited[nvis++] = static_cast<uint32_t>(next);
                }
                const uint32_t order = w * L + step;
                ev[order] = record ? ((static_cast<uint64_t>(next) << 32) | order) : ~0ull;
                cur = next;
            }
        }
        for (int t = total + lane; t < P.cap_ev; t += 32) ev[t] = ~0ull;
        __syncwarp();
        // ---- 2. sort by (column, event order)
        warp_bitonic_sort(ev, P.cap_ev, lane);
        // ---- 3. reduce every run of equal columns in event order; compact (col, value) in place
        int m = 0;  // entries written so far
        uint32_t prev_col = 0xffffffffu;
        bool have_prev = false;
        for (int base = 0; base < static_cast<int>(total); base += 32) {
            const int t = base + lane;
            uint64_t k = t < static_cast<int>(total) ? ev[t] : ~0ull;
            const bool valid = k != ~0ull;
            const uint32_t c = static_cast<uint32_t>(k >> 32);
            uint32_t pc = __shfl_up_sync(0xffffffffu, c, 1);
            bool pv = __shfl_up_sync(0xffffffffu, valid ? 1 : 0, 1) != 0;
            if (lane == 0) { pc = prev_col; pv = have_prev; }
            const bool head = valid && !(pv && pc == c);
            float sum = 0.f;
            int cnt = 0;
            if (head) {
                for (int u = t; u < static_cast<int>(total); ++u) {
                    const uint64_t ku = ev[u];
                    if (ku == ~0ull || static_cast<uint32_t>(ku >> 32) != c) break;
                    sum += step_weight(P.weighting, static_cast<uint32_t>(ku) % L, L);
                    ++cnt;
                }
                if (P.weighting == SPHB_W_FIRST_VISIT) {
                    // hits /= hitCount, then the linear map onto [0, 1] (SparseMatrixAlgorithms.cpp:176-218)
                    sum /= static_cast<float>(cnt);
                    const float mm = -1.f / (L - 1.f), cc = L / (L - 1.f);
                    sum = fmaxf(0.f, mm * sum + cc);
                }
            }
            bool keep = head;
            if (keep && P.prune_value > 0.f && (sum <= P.prune_value || sum == 0.f)) keep = false;
            prev_col = __shfl_sync(0xffffffffu, c, 31);
            have_prev = __shfl_sync(0xffffffffu, valid ? 1 : 0, 31) != 0;
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            __syncwarp();  // all reads of this chunk (and the forward run scans) are done before the in-place writes
            if (keep) ev[m + __popc(mask & ((1u << lane) - 1u))] = (static_cast<uint64_t>(c) << 32) | __float_as_uint(sum);
            m += __popc(mask);
            __syncwarp();
            if (!__any_sync(0xffffffffu, valid)) break;  // sorted: only sentinels follow
        }
        // ---- 4. removeDiagonalElements(keepSingleEntry = true): with more than one entry, drop the diagonal and every exact zero
        if (P.remove_diagonal && m > 1) {
            int m2 = 0;
            for (int base = 0; base < m; base += 32) {
                const int t = base + lane;
                const uint64_t k = t < m ? ev[t] : 0ull;
                const bool keep = t < m && static_cast<int64_t>(k >> 32) != start && __uint_as_float(static_cast<uint32_t>(k)) != 0.f;
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                __syncwarp();
                if (keep) ev[m2 + __popc(mask & ((1u << lane) - 1u))] = k;
                m2 += __popc(mask);
                __syncwarp();
            }
            m = m2;
        }
        // ---- 5. L1 normalisation: fp32 sum in column order (Eigen sum()), then the single write of the row
        float nv = 1.f;
        if (P.normalize) {
            float acc = 0.f;
            if (lane == 0)
                for (int t = 0; t < m; ++t) acc += __uint_as_float(static_cast<uint32_t>(ev[t]));
            nv = __shfl_sync(0xffffffffu, acc, 0);
        }
        uint32_t* oc = P.tmp_col + s * P.tmp_stride;
        float* ov = P.tmp_val + s * P.tmp_stride;
        for (int t = lane; t < m; t += 32) {
            const uint64_t k = ev[t];
            oc[t] = static_cast<uint32_t>(k >> 32);
            const float v = __uint_as_float(static_cast<uint32_t>(k));
            ov[t] = P.normalize ? v / nv : v;
        }
        if (lane == 0) P.row_nnz[s] = m;
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------
// The production kernel.  lane = WALK_CHAINS interleaved walks; every step is a binary search in the row's running sums
// (7 dependent 8-byte probes for a 90-entry row instead of a serial scan of ~45 entries; the chains of a lane hide each
// other's latency).  The events of a start are kept as (column u32, step u16) in event order — the array index IS the
// event order — and grouped by column with a STABLE LSD radix sort in shared memory (digits of <= 9 bits; per 32-event chunk
// a match.any gives the rank among equal digits), so every run of equal columns is still in event order and its fp32 sum is
// the reference's `hits[next] += stepWeight` sequence.  Step weights come from a table built once per block.
// (ncu on the first version — u64 keys, bitonic sort of 2048, expf per event: 90 k warp instructions per start, half of them
// the sort, a third the weights; profiles/r02_walk_*.)
// Shared memory per warp: colA, colB u32[cap] | stepA, stepB u16[cap] | cnt u32[512]; per block: wt float[L].
// ---------------------------------------------------------------------------------------------
constexpr int WALK_RADIX_BITS = 9;

__device__ __forceinline__ void warp_radix_pass(const uint32_t* __restrict__ ksrc, const uint16_t* __restrict__ psrc, uint32_t* __restrict__ kdst,
                                                uint16_t* __restrict__ pdst, uint32_t* __restrict__ cnt, int nev, int shift, int bits, int lane) {
    const int bins = 1 << bits;
    const uint32_t dmask = static_cast<uint32_t>(bins - 1);
    for (int b = lane; b < bins; b += 32) cnt[b] = 0;
    __syncwarp();
    for (int t = lane; t < nev; t += 32) atomicAdd(&cnt[(ksrc[t] >> shift) & dmask], 1u);
    __syncwarp();
    // exclusive scan of the bins: lane owns bins [lane * per, (lane + 1) * per)
    const int per = bins >> 5;  // bins >= 32
    uint32_t local = 0;
    for (int b = 0; b < per; ++b) local += cnt[lane * per + b];
    uint32_t incl = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
    }
    uint32_t run = incl - local;
    for (int b = 0; b < per; ++b) {
        const uint32_t c = cnt[lane * per + b];
        cnt[lane * per + b] = run;
        run += c;
    }
    __syncwarp();
    // stable scatter, chunk by chunk in event order
    for (int base = 0; base < nev; base += 32) {
        const int t = base + lane;
        const bool valid = t < nev;
        const uint32_t key = valid ? ksrc[t] : 0u;
        const uint16_t pay = valid ? psrc[t] : static_cast<uint16_t>(0);
        const uint32_t dg = valid ? ((key >> shift) & dmask) : static_cast<uint32_t>(bins + lane);  // invalid lanes group with nobody
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        const uint32_t off = valid ? cnt[dg] : 0u;
        __syncwarp();
        if (valid && (__ffs(peers) - 1) == lane) cnt[dg] = off + __popc(peers);
        __syncwarp();
        if (valid) {
            const uint32_t pos = off + __popc(peers & ((1u << lane) - 1u));
            kdst[pos] = key;
            pdst[pos] = pay;
        }
    }
    __syncwarp();
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) random_walk_search_kernel(const WalkParams P) {
    extern __shared__ uint64_t smem_ev[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t W = P.W, L = P.L;
    const bool all_steps = P.weighting != SPHB_W_ONLYLAST;       // ONLYLAST records one event per walk
    const int nev = static_cast<int>(all_steps ? W * L : W);
    const int cap = P.cap_ev;                                     // >= nev, multiple of 32
    uint8_t* base8 = reinterpret_cast<uint8_t*>(smem_ev) + static_cast<size_t>(warp) * (static_cast<size_t>(cap) * 12 + 2048);
    uint32_t* colA = reinterpret_cast<uint32_t*>(base8);
    uint32_t* colB = colA + cap;
    uint16_t* stepA = reinterpret_cast<uint16_t*>(colB + cap);
    uint16_t* stepB = stepA + cap;
    uint32_t* cnt = reinterpret_cast<uint32_t*>(stepB + cap);
    float* wt = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(smem_ev) + static_cast<size_t>(WARPS) * (static_cast<size_t>(cap) * 12 + 2048));
    for (uint32_t t = threadIdx.x; t < L; t += WARPS * 32) wt[t] = step_weight(P.weighting, t, L);
    __syncthreads();
    const uint2 key = make_uint2(static_cast<uint32_t>(P.seed), static_cast<uint32_t>(P.seed >> 32));
    for (int64_t s = static_cast<int64_t>(blockIdx.x) * WARPS + warp; s < P.nstarts; s += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t start = P.start0 + s;
        // ---- 1. the walks
        for (uint32_t w0 = 0; w0 < W; w0 += 32 * WALK_CHAINS) {
            int64_t cur[WALK_CHAINS];
            bool live[WALK_CHAINS];
#pragma unroll
            for (int c = 0; c < WALK_CHAINS; ++c) {
                cur[c] = start;
                live[c] = w0 + c * 32 + lane < W;
            }
            for (uint32_t step = 0; step < L; ++step) {
                double u[WALK_CHAINS];
                int64_t lo[WALK_CHAINS], hi[WALK_CHAINS], e1[WALK_CHAINS];
#pragma unroll
                for (int c = 0; c < WALK_CHAINS; ++c) {
                    const uint32_t w = w0 + c * 32 + lane;
                    const uint4 r = philox4x32_10(make_uint4(static_cast<uint32_t>(start), static_cast<uint32_t>(start >> 32), w, step), key);
                    u[c] = u01_from_bits(r.x, r.y);
                    lo[c] = live[c] ? __ldg(P.rowptr + cur[c]) : 0;
                    e1[c] = live[c] ? __ldg(P.rowptr + cur[c] + 1) : 0;
                    hi[c] = e1[c];
                }
                // first entry with u < cum (lower bound on a non-decreasing array); the chains advance together
                while (true) {
                    bool any = false;
                    double cm[WALK_CHAINS];
                    int64_t mid[WALK_CHAINS];
#pragma unroll
                    for (int c = 0; c < WALK_CHAINS; ++c) {
                        mid[c] = (lo[c] + hi[c]) >> 1;
                        if (lo[c] < hi[c]) {
                            cm[c] = __ldg(P.cum + mid[c]);
                            any = true;
                        }
                    }
                    if (!any) break;
#pragma unroll
                    for (int c = 0; c < WALK_CHAINS; ++c) {
                        if (lo[c] < hi[c]) {
                            if (u[c] < cm[c]) hi[c] = mid[c];
                            else lo[c] = mid[c] + 1;
                        }
                    }
                }
#pragma unroll
                for (int c = 0; c < WALK_CHAINS; ++c) {
                    if (!live[c]) continue;
                    const uint32_t w = w0 + c * 32 + lane;
                    const int64_t next = lo[c] < e1[c] ? static_cast<int64_t>(__ldg(P.col + lo[c])) : cur[c];  // never found: stays put
                    if (all_steps) {
                        colA[w * L + step] = static_cast<uint32_t>(next);
                        stepA[w * L + step] = static_cast<uint16_t>(step);
                    } else if (step == L - 1) {
                        colA[w] = static_cast<uint32_t>(next);
                        stepA[w] = static_cast<uint16_t>(step);
                    }
                    cur[c] = next;
                }
            }
        }
        __syncwarp();
        // ---- 2. group by column, stable (event order survives inside every run)
        uint32_t* ck = colA;
        uint16_t* cp = stepA;
        uint32_t* ok = colB;
        uint16_t* op = stepB;
        for (int shift = 0; shift < P.col_bits; shift += P.radix_bits) {
            int bits = min(P.radix_bits, P.col_bits - shift);
            if (bits < 5) bits = 5;  // at least 32 bins (one per lane in the scan); the extra digit bits are zero
            warp_radix_pass(ck, cp, ok, op, cnt, nev, shift, bits, lane);
            uint32_t* tk = ck; ck = ok; ok = tk;
            uint16_t* tp = cp; cp = op; op = tp;
        }
        // ---- 3. reduce every run of equal columns in event order; (column, value) -> (ok, float view of ck) compacted in place
        float* vout = reinterpret_cast<float*>(ck);
        int m = 0;
        uint32_t prev_col = 0xffffffffu;
        bool have_prev = false;
        for (int base = 0; base < nev; base += 32) {
            const int t = base + lane;
            const bool valid = t < nev;
            const uint32_t c = valid ? ck[t] : 0xffffffffu;
            uint32_t pc = __shfl_up_sync(0xffffffffu, c, 1);
            bool pv = __shfl_up_sync(0xffffffffu, valid ? 1 : 0, 1) != 0;
            if (lane == 0) { pc = prev_col; pv = have_prev; }
            const bool head = valid && !(pv && pc == c);
            float sum = 0.f;
            if (head) {
                for (int u = t; u < nev && ck[u] == c; ++u) sum += all_steps ? wt[cp[u]] : 1.f;
            }
            bool keep = head;
            if (keep && P.prune_value > 0.f && (sum <= P.prune_value || sum == 0.f)) keep = false;
            prev_col = __shfl_sync(0xffffffffu, c, 31);
            have_prev = __shfl_sync(0xffffffffu, valid ? 1 : 0, 31) != 0;
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            __syncwarp();  // all reads of this chunk (and the forward run scans) are done before the in-place writes
            if (keep) {
                const int o = m + __popc(mask & ((1u << lane) - 1u));
                ok[o] = c;
                vout[o] = sum;
            }
            m += __popc(mask);
            __syncwarp();
        }
        // ---- 4. removeDiagonalElements(keepSingleEntry = true): with more than one entry, drop the diagonal and every exact zero
        if (P.remove_diagonal && m > 1) {
            int m2 = 0;
            for (int base = 0; base < m; base += 32) {
                const int t = base + lane;
                const uint32_t c = t < m ? ok[t] : 0u;
                const float v = t < m ? vout[t] : 0.f;
                const bool keep = t < m && static_cast<int64_t>(c) != start && v != 0.f;
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                __syncwarp();
                if (keep) {
                    const int o = m2 + __popc(mask & ((1u << lane) - 1u));
                    ok[o] = c;
                    vout[o] = v;
                }
                m2 += __popc(mask);
                __syncwarp();
            }
            m = m2;
        }
        // ---- 5. L1 normalisation: fp32 sum in column order (Eigen sum()), then the single write of the row
        float nv = 1.f;
        if (P.normalize) {
            float acc = 0.f;
            if (lane == 0)
                for (int t = 0; t < m; ++t) acc += vout[t];
            nv = __shfl_sync(0xffffffffu, acc, 0);
        }
        uint32_t* oc = P.tmp_col + s * P.tmp_stride;
        float* ov = P.tmp_val + s * P.tmp_stride;
        for (int t = lane; t < m; t += 32) {
            oc[t] = ok[t];
            ov[t] = P.normalize ? vout[t] / nv : vout[t];
        }
        if (lane == 0) P.row_nnz[s] = m;
        __syncwarp();
    }
}

// fixed-stride temp rows -> CSR at the scanned offsets
__global__ void walks_compact(int64_t nstarts, int tmp_stride, const uint32_t* __restrict__ tmp_col, const float* __restrict__ tmp_val,
                              const int64_t* __restrict__ offs, int64_t base, int64_t cap, int64_t* __restrict__ rowptr_out,
                              uint32_t* __restrict__ out_col, float* __restrict__ out_val) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
    for (int64_t s = warp; s < nstarts; s += nwarps) {
        const int64_t b = base + offs[s], e = base + offs[s + 1];
        if (lane == 0) rowptr_out[s + 1] = e;
        if (e > cap) continue;
        for (int64_t t = lane; t < e - b; t += 32) {
            out_col[b + t] = tmp_col[s * tmp_stride + t];
            out_val[b + t] = tmp_val[s * tmp_stride + t];
        }
    }
}
__global__ void widen_counts(const int* __restrict__ in, int64_t num, int64_t* __restrict__ out) {
    for (int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; t <= num; t += static_cast<int64_t>(gridDim.x) * blockDim.x)
        out[t] = t < num ? in[t] : 0;
}

// Start points [q0, q1) only (the multi-GPU shard: every walk may read any row of the replicated matrix, the start points
// are independent, and the counter-based RNG is keyed by the GLOBAL start id, so the slabs of a sharded run concatenate to
// exactly the single-GPU result).  d_out_rowptr: q1 - q0 + 1 entries, relative to the slab.
int random_walks_slab_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks, uint64_t walk_len,
                          int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize, int64_t q0,
                          int64_t q1, int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    *out_nnz = 0;
    if (q0 < 0 || q1 > n || q1 < q0) return fail(SPHB_ERR_INVALID, "random_walks: bad start-point range");
    SPHB_CUDA(cudaMemsetAsync(d_out_rowptr, 0, (q1 - q0 + 1) * sizeof(int64_t), s));
    if (n == 0 || num_walks == 0 || q1 == q0) return 0;
    const uint64_t total = num_walks * walk_len;
    if (total > 8192) return fail(SPHB_ERR_UNSUPPORTED, "random_walks: numRandomWalks * singleWalkLength > 8192 unsupported");
    if (weighting == SPHB_W_FIRST_VISIT && walk_len > FV_MAX_LEN) return fail(SPHB_ERR_UNSUPPORTED, "random_walks: FIRST_VISIT with walk length > 64 unsupported");
    // pruneValue from pruneSteps (SparseMatrixAlgorithms.cpp:48-66)
    float pv = prune_value;
    if (prune_steps > 0) {
        const uint64_t ps = std::min<uint64_t>(walk_len - 1, prune_steps);
        const uint64_t inv = walk_len - ps;
        if (weighting == SPHB_W_LINEAR) pv = static_cast<float>(1.0 - (static_cast<double>(inv) / walk_len));
        else if (weighting == SPHB_W_NORMAL) { const float idx = inv * 3.f / walk_len; pv = std::exp(-0.5f * std::pow(idx, 2.f)); }
        else if (weighting == SPHB_W_CONSTANT) pv = static_cast<float>(inv);
        else if (weighting == SPHB_W_FIRST_VISIT) pv = static_cast<float>(ps) / walk_len;
        if (pv > 0.5f) pv = 0.5f;
    }
    int cap_ev = 32;
    while (cap_ev < static_cast<int>(total)) cap_ev <<= 1;
    const int sms = ctx().num_sms;
    constexpr int WARPS = 4;
    int warps = WARPS;
    size_t smem = static_cast<size_t>(WARPS) * cap_ev * sizeof(uint64_t);
    if (smem > 200 * 1024) { warps = 1; smem = static_cast<size_t>(cap_ev) * sizeof(uint64_t); }
    const int tmp_stride = static_cast<int>(std::min<uint64_t>(total, static_cast<uint64_t>(n)));
    // chunk of start points per launch: bounded temp storage (<= 4 GiB), rows compacted chunk by chunk
    int64_t chunk = std::max<int64_t>(1024, (4ll << 30) / (static_cast<int64_t>(tmp_stride) * 8));
    chunk = std::min<int64_t>(chunk, q1 - q0);
    DevBuf<uint32_t> tcol;
    DevBuf<float> tval;
    DevBuf<int> cnt;
    DevBuf<int64_t> cnt64, offs;
    SPHB_CUDA(tcol.alloc(static_cast<size_t>(chunk) * tmp_stride, s));
    SPHB_CUDA(tval.alloc(static_cast<size_t>(chunk) * tmp_stride, s));
    SPHB_CUDA(cnt.alloc(chunk, s));
    SPHB_CUDA(cnt64.alloc(chunk + 1, s));
    SPHB_CUDA(offs.alloc(chunk + 1, s));
    // running sums for the binary-search kernel (8 B per entry); FIRST_VISIT and matrices with negative entries use the scan kernel
    DevBuf<double> cum;
    bool search = weighting != SPHB_W_FIRST_VISIT && !g_force_scan;
    if (search) {
        int64_t nnz_in = 0;
        SPHB_CUDA(cudaMemcpyAsync(&nnz_in, d_rowptr + n, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaStreamSynchronize(s));
        DevBuf<int> negative;
        SPHB_CUDA(negative.alloc(1, s));
        SPHB_CUDA(cudaMemsetAsync(negative.p, 0, sizeof(int), s));
        SPHB_CUDA(cum.alloc(static_cast<size_t>(std::max<int64_t>(nnz_in, 1)), s));
        walk_build_cum<<<static_cast<unsigned>(std::max<int64_t>(1, std::min<int64_t>((n + 127) / 128, static_cast<int64_t>(sms) * 16))), 128, 0, s>>>(
            n, d_rowptr, d_val, cum.p, negative.p);
        count_launch();
        int hneg = 0;
        SPHB_CUDA(cudaMemcpyAsync(&hneg, negative.p, sizeof(int), cudaMemcpyDeviceToHost, s));
        SPHB_CUDA(cudaStreamSynchronize(s));
        if (hneg) search = false;
    }
    // shared memory: scan kernel u64 events (power of two for the bitonic sort); search kernel (col u32, step u16) x 2 buffers + 512 radix
    // counters per warp + the step-weight table
    int cap_s = 0, col_bits = 0, radix_bits = 0;
    if (search) {
        const int nev = static_cast<int>(weighting == SPHB_W_ONLYLAST ? num_walks : total);
        cap_s = (nev + 31) / 32 * 32;
        smem = static_cast<size_t>(WARPS) * (static_cast<size_t>(cap_s) * 12 + 2048) + walk_len * sizeof(float);
        warps = WARPS;
        if (smem > 200 * 1024) { warps = 1; smem = static_cast<size_t>(cap_s) * 12 + 2048 + walk_len * sizeof(float); }
        col_bits = std::max(5, bits_for(static_cast<uint64_t>(n)));
        const int passes = (col_bits + WALK_RADIX_BITS - 1) / WALK_RADIX_BITS;
        radix_bits = (col_bits + passes - 1) / passes;
    }
    auto set_attr = [&](auto kern) { return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)); };
    if (warps == WARPS) SPHB_CUDA(search ? set_attr(random_walk_search_kernel<WARPS>) : set_attr(random_walk_kernel<WARPS>));
    else SPHB_CUDA(search ? set_attr(random_walk_search_kernel<1>) : set_attr(random_walk_kernel<1>));
    int64_t base = 0;
    for (int64_t s0 = q0; s0 < q1; s0 += chunk) {
        const int64_t ns = std::min(chunk, q1 - s0);
        WalkParams P{};
        P.n = n; P.rowptr = d_rowptr; P.col = d_col; P.val = d_val;
        P.W = static_cast<uint32_t>(num_walks); P.L = static_cast<uint32_t>(walk_len);
        P.weighting = weighting; P.prune_value = pv; P.seed = seed; P.remove_diagonal = remove_diagonal; P.normalize = normalize;
        P.cap_ev = cap_ev; P.start0 = s0; P.nstarts = ns; P.tmp_stride = tmp_stride;
        P.tmp_col = tcol.p; P.tmp_val = tval.p; P.row_nnz = cnt.p; P.cum = cum.p;
        if (search) { P.cap_ev = cap_s; P.col_bits = col_bits; P.radix_bits = radix_bits; }
        const int64_t blocks = std::min<int64_t>((ns + warps - 1) / warps, static_cast<int64_t>(sms) * 16);
        if (warps == WARPS && search) random_walk_search_kernel<WARPS><<<static_cast<unsigned>(blocks), WARPS * 32, smem, s>>>(P);
        else if (warps == WARPS) random_walk_kernel<WARPS><<<static_cast<unsigned>(blocks), WARPS * 32, smem, s>>>(P);
        else if (search) random_walk_search_kernel<1><<<static_cast<unsigned>(blocks), 32, smem, s>>>(P);
        else random_walk_kernel<1><<<static_cast<unsigned>(blocks), 32, smem, s>>>(P);
        widen_counts<<<static_cast<unsigned>(std::min<int64_t>((ns + 256) / 256, sms * 8)), 256, 0, s>>>(cnt.p, ns, cnt64.p);
        count_launch(2);
        if (int rc = exclusive_sum<int64_t>(cnt64.p, offs.p, ns + 1, s)) return rc;
        int64_t chunk_nnz = 0;
        SPHB_CUDA(cudaMemcpyAsync(&chunk_nnz, offs.p + ns, sizeof(int64_t), cudaMemcpyDeviceToHost, s));
        walks_compact<<<static_cast<unsigned>(std::min<int64_t>((ns * 32 + 255) / 256, sms * 8)), 256, 0, s>>>(ns, tmp_stride, tcol.p, tval.p, offs.p, base, cap,
                                                                                                              d_out_rowptr + (s0 - q0), d_out_col, d_out_val);
        count_launch();
        SPHB_CUDA(cudaGetLastError());
        SPHB_CUDA(cudaStreamSynchronize(s));
        base += chunk_nnz;
        if (base > cap) return fail(SPHB_ERR_CAPACITY, "random_walks: output capacity too small");
    }
    *out_nnz = base;
    return 0;
}

int random_walks_dev(int64_t n, const int64_t* d_rowptr, const uint32_t* d_col, const float* d_val, uint64_t num_walks, uint64_t walk_len,
                     int weighting, float prune_value, uint64_t prune_steps, uint64_t seed, int remove_diagonal, int normalize,
                     int64_t* d_out_rowptr, uint32_t* d_out_col, float* d_out_val, int64_t cap, int64_t* out_nnz, cudaStream_t s) {
    return random_walks_slab_dev(n, d_rowptr, d_col, d_val, num_walks, walk_len, weighting, prune_value, prune_steps, seed, remove_diagonal, normalize, 0, n,
                                 d_out_rowptr, d_out_col, d_out_val, cap, out_nnz, s);
}

}  // namespace sphb
