// Exact kNN, SIMT flavour: one warp per query row, fp64 pair keys evaluated
// directly (no GEMM), running top-k kept as sorted u64 keys in shared memory.
// Used where the tensor-core sweep does not pay: tiny n, very small d (the
// 3-channel RGB configuration is FP64/ALU bound) — and as the
// on-device cross-check of the tcgen05 path in the GPU tests.
#include "../../include/sphb.h"
#include "knn_common.cuh"

namespace sphb {

// fp32 screen in front of the exact fp64 key (L2, d <= SIMT_SCREEN_D): the fp32 sum of squared differences has a relative error
// of at most (d + 4) 2^-24 (all terms >= 0), so a pair whose fp32 value exceeds (1 + band) x the current k-th exact value cannot
// enter the list and skips the fp64 chain (two conversions + DADD + DFMA per channel).  At d = 3 (C3, RGB) the fp64 chain of
// every pair bounded the kernel at 3.2e11 pairs/s.  The result is unchanged: every pair that could pass is still keyed exactly.
constexpr int SIMT_SCREEN_D = 16;
__device__ __forceinline__ float screen_limit(uint64_t kth, int d) {
    if (kth == ~0ull) return INFINITY;
    const float band = static_cast<float>(d + 4) * 5.9604645e-8f;
    return key_val(kth) * (1.f + band) * (1.f + 2.4e-7f) + 1e-37f;
}
template <int DMAX>  // query values in registers: the loops are unrolled over DMAX >= d
__device__ __forceinline__ float l2_fp32(const float (&q)[DMAX > 0 ? DMAX : 1], const float* __restrict__ xj, int d) {
    float f = 0.f;
#pragma unroll
    for (int c = 0; c < DMAX; ++c) {
        if (c < d) {
            const float t = q[c] - __ldg(xj + c);
            f = fmaf(t, t, f);
        }
    }
    return f;
}

template <int WARPS, int DMAX>  // DMAX > 0: fp32 screen for L2 with d <= DMAX
__global__ void __launch_bounds__(WARPS * 32) knn_simt_kernel(const float* __restrict__ X, int64_t n, int d, int k, int cap, int metric,
                                                              int l2_squared, int64_t q0, int64_t q1, const int64_t* __restrict__ rowlist,
                                                              const int* __restrict__ rowcount, int64_t* __restrict__ ids,
                                                              float* __restrict__ dist) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    // rows to process: the slab [q0,q1), or (rowlist != null) an explicit list of global row ids
    const int64_t nq = rowlist ? static_cast<int64_t>(*rowcount) : q1 - q0;
    for (int64_t rr = static_cast<int64_t>(blockIdx.x) * WARPS + warp; rr < nq; rr += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t i = rowlist ? rowlist[rr] : q0 + rr;
        const int64_t r = i - q0;  // output row within the slab
        int cnt = 0;
        uint64_t kth = ~0ull;
        constexpr bool screen = DMAX > 0;
        float q[DMAX > 0 ? DMAX : 1];
        if constexpr (screen) {
#pragma unroll
            for (int c = 0; c < DMAX; ++c) q[c] = c < d ? __ldg(X + i * d + c) : 0.f;
        }
        float lim = INFINITY;
        int64_t base0 = 0;
        if constexpr (DMAX == 4) {
            // d == 3 (RGB): a lane takes four consecutive points = 48 bytes = three aligned float4 loads.  With one point per lane
            // every 4-byte load instruction of a warp spans 384 bytes (4 L1 wavefronts for 128 useful bytes) and the L1 bounds the
            // scan at ~2.7 pairs per SM and cycle (measured 2.15); the vector form moves the same bytes in a quarter of the wavefronts.
            if (d == 3 && (reinterpret_cast<uintptr_t>(X) & 15) == 0) {
                const int64_t nvec = n & ~static_cast<int64_t>(127);
                for (; base0 < nvec; base0 += 128) {
                    const int64_t j0 = base0 + 4 * lane;
                    const float4* p4 = reinterpret_cast<const float4*>(X + j0 * 3);
                    const float4 a = __ldg(p4), b = __ldg(p4 + 1), c4 = __ldg(p4 + 2);
                    const float px[4] = {a.x, a.w, b.z, c4.y}, py[4] = {a.y, b.x, b.w, c4.z}, pz[4] = {a.z, b.y, c4.x, c4.w};
                    float f4[4];
#pragma unroll
                    for (int sub = 0; sub < 4; ++sub) {
                        const float tx = q[0] - px[sub], ty = q[1] - py[sub], tz = q[2] - pz[sub];
                        f4[sub] = fmaf(tz, tz, fmaf(ty, ty, tx * tx));
                    }
                    if (!__any_sync(0xffffffffu, fminf(fminf(f4[0], f4[1]), fminf(f4[2], f4[3])) <= lim)) continue;  // the common case: nothing near
#pragma unroll
                    for (int sub = 0; sub < 4; ++sub) {
                        const float f = f4[sub];
                        bool pass = false;
                        uint64_t key = ~0ull;
                        if (f <= lim) {
                            key = make_key(exact_key(X, i, j0 + sub, d, metric), static_cast<uint32_t>(j0 + sub));
                            pass = key <= kth;
                        }
                        const unsigned m = __ballot_sync(0xffffffffu, pass);
                        if (m) {
                            if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                            cnt += __popc(m);
                            if (cnt + 32 > cap) {
                                for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
                                __syncwarp();
                                warp_bitonic_sort(sm, cap, lane);
                                cnt = min(cnt, k);
                                kth = (cnt == k) ? sm[k - 1] : ~0ull;
                                lim = screen_limit(kth, d);
                            }
                        }
                    }
                }
            }
        }
        for (int64_t base = base0; base < n; base += 32) {
            const int64_t j = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (j < n && (!screen || l2_fp32<DMAX>(q, X + j * d, d) <= lim)) {
                key = make_key(exact_key(X, i, j, d, metric), static_cast<uint32_t>(j));
                pass = key <= kth;
            }
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(sm, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? sm[k - 1] : ~0ull;
                    lim = screen_limit(kth, d);
                }
            }
        }
        for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        cnt = min(cnt, k);
        write_row(sm, cnt, k, i, metric, l2_squared, ids + r * k, dist + r * k, lane);
        __syncwarp();
    }
}

// Few rows (the overflow fallback of the tensor-core path): one warp per (row, database segment) keeps the best k of
// its segment, a second kernel merges the S sorted segment lists of a row.  Same keys, same (value, index) order.
template <int WARPS, int DMAX>
__global__ void __launch_bounds__(WARPS * 32) knn_simt_seg_kernel(const float* __restrict__ X, int64_t n, int d, int k, int cap, int metric,
                                                                  const int64_t* __restrict__ rowlist, const int* __restrict__ rowcount, int S,
                                                                  int64_t seglen, uint64_t* __restrict__ tmp) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t nwork = static_cast<int64_t>(*rowcount) * S;
    for (int64_t w = static_cast<int64_t>(blockIdx.x) * WARPS + warp; w < nwork; w += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t i = rowlist[w / S];
        const int64_t j0 = (w % S) * seglen, j1 = min(n, j0 + seglen);
        int cnt = 0;
        uint64_t kth = ~0ull;
        constexpr bool screen = DMAX > 0;
        float q[DMAX > 0 ? DMAX : 1];
        if constexpr (screen) {
#pragma unroll
            for (int c = 0; c < DMAX; ++c) q[c] = c < d ? __ldg(X + i * d + c) : 0.f;
        }
        float lim = INFINITY;
        for (int64_t base = j0; base < j1; base += 32) {
            const int64_t j = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (j < j1 && (!screen || l2_fp32<DMAX>(q, X + j * d, d) <= lim)) {
                key = make_key(exact_key(X, i, j, d, metric), static_cast<uint32_t>(j));
                pass = key <= kth;
            }
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(sm, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? sm[k - 1] : ~0ull;
                    lim = screen_limit(kth, d);
                }
            }
        }
        for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        for (int t = lane; t < k; t += 32) tmp[w * k + t] = sm[t];  // padded with ~0
        __syncwarp();
    }
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) knn_simt_merge_kernel(const uint64_t* __restrict__ tmp, int S, int k, int cap, int metric, int l2_squared,
                                                                    int64_t q0, const int64_t* __restrict__ rowlist, const int* __restrict__ rowcount,
                                                                    int64_t* __restrict__ ids, float* __restrict__ dist) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t nq = *rowcount;
    for (int64_t rr = static_cast<int64_t>(blockIdx.x) * WARPS + warp; rr < nq; rr += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t i = rowlist[rr];
        const uint64_t* src = tmp + rr * S * k;
        const int total = S * k;
        int cnt = 0;
        uint64_t kth = ~0ull;
        for (int base = 0; base < total; base += 32) {
            const int t = base + lane;
            const uint64_t key = t < total ? src[t] : ~0ull;
            const bool pass = key != ~0ull && key <= kth;
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int u = cnt + lane; u < cap; u += 32) sm[u] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(sm, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? sm[k - 1] : ~0ull;
                }
            }
        }
        for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        cnt = min(cnt, k);
        const int64_t r = i - q0;
        write_row(sm, cnt, k, i, metric, l2_squared, ids + r * k, dist + r * k, lane);
        __syncwarp();
    }
}

static int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

template <int DMAX>
static int knn_simt_launch_t(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1, const int64_t* d_rowlist,
                             const int* d_rowcount, int64_t max_rows, int64_t* d_ids, float* d_dist, cudaStream_t stream) {
    constexpr int WARPS = 4;
    const int cap = next_pow2(k + 64);
    const size_t smem = static_cast<size_t>(WARPS) * cap * sizeof(uint64_t);
    if (smem > 200 * 1024) return fail(5, "knn_simt: k too large");
    SPHB_CUDA(cudaFuncSetAttribute(knn_simt_kernel<WARPS, DMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t nq = d_rowlist ? max_rows : q1 - q0;
    if (nq <= 0) return 0;
    const int64_t target_warps = static_cast<int64_t>(ctx().num_sms) * 32;
    if (d_rowlist && nq * 2 <= target_warps && n >= 4096) {
        int S = static_cast<int>(std::min<int64_t>((target_warps + nq - 1) / nq, std::min<int64_t>(256, n / 1024)));
        if (S < 2) S = 2;
        const int64_t seglen = ((n + S - 1) / S + 31) / 32 * 32;
        DevBuf<uint64_t> tmp;
        SPHB_CUDA(tmp.alloc(static_cast<size_t>(nq) * S * k, stream));
        SPHB_CUDA(cudaFuncSetAttribute(knn_simt_seg_kernel<WARPS, DMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        SPHB_CUDA(cudaFuncSetAttribute(knn_simt_merge_kernel<WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        const int64_t b1 = std::min<int64_t>((nq * S + WARPS - 1) / WARPS, static_cast<int64_t>(ctx().num_sms) * 16);
        knn_simt_seg_kernel<WARPS, DMAX><<<static_cast<unsigned>(b1), WARPS * 32, smem, stream>>>(d_x, n, d, k, cap, metric, d_rowlist, d_rowcount, S, seglen,
                                                                                                  tmp.p);
        count_launch();
        SPHB_CUDA(cudaGetLastError());
        const int64_t b2 = (nq + WARPS - 1) / WARPS;
        knn_simt_merge_kernel<WARPS><<<static_cast<unsigned>(b2), WARPS * 32, smem, stream>>>(tmp.p, S, k, cap, metric, l2_squared, q0, d_rowlist, d_rowcount,
                                                                                              d_ids, d_dist);
        count_launch();
        SPHB_CUDA(cudaGetLastError());
        return 0;
    }
    int64_t blocks = (nq + WARPS - 1) / WARPS;
    const int64_t maxb = static_cast<int64_t>(ctx().num_sms) * 16;
    if (blocks > maxb) blocks = maxb;
    knn_simt_kernel<WARPS, DMAX><<<static_cast<unsigned>(blocks), WARPS * 32, smem, stream>>>(d_x, n, d, k, cap, metric, l2_squared, q0, q1, d_rowlist,
                                                                                                  d_rowcount, d_ids, d_dist);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

int knn_simt_launch(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1, const int64_t* d_rowlist,
                    const int* d_rowcount, int64_t max_rows, int64_t* d_ids, float* d_dist, cudaStream_t stream) {
    // the fp32 screen pays where the exact fp64 chain is most of a pair's cost: L2, few channels
    if (metric == METRIC_L2 && d <= 4) return knn_simt_launch_t<4>(d_x, n, d, k, metric, l2_squared, q0, q1, d_rowlist, d_rowcount, max_rows, d_ids, d_dist, stream);
    if (metric == METRIC_L2 && d <= SIMT_SCREEN_D)
        return knn_simt_launch_t<SIMT_SCREEN_D>(d_x, n, d, k, metric, l2_squared, q0, q1, d_rowlist, d_rowcount, max_rows, d_ids, d_dist, stream);
    return knn_simt_launch_t<0>(d_x, n, d, k, metric, l2_squared, q0, q1, d_rowlist, d_rowcount, max_rows, d_ids, d_dist, stream);
}

// ---------------------------------------------------------------------------------------------
// No-allocation variant of the segmented exact search (sphb_knn_rows_noalloc_dev, include/sphb.h).
// No counterpart in the reference; added for callers that cannot
// allocate device memory (CUDA graph capture, tight serving loops).  The scratch of knn_simt_seg_kernel —
// the per-(row, segment) sorted lists — lives in a file-scope device array of fixed size instead of a
// stream-ordered allocation.  Every GPU that loads the library holds the array until exit; all calls share
// it (calls that overlap on the GPU race), and it keeps the last call's contents (every slot a call reads is
// written by the same call first; the row count slot is set per call).
// ---------------------------------------------------------------------------------------------
constexpr int NOALLOC_MAX_ROWS = 256;   // query rows per call
constexpr int NOALLOC_MAX_K = 128;      // neighbours per row
constexpr int NOALLOC_SEGMENTS = 32;    // database segments per row
constexpr size_t NOALLOC_LIST_WORDS = static_cast<size_t>(NOALLOC_MAX_ROWS) * NOALLOC_SEGMENTS * NOALLOC_MAX_K;  // 1 Mi u64 = 8 MiB
// [0, 32): header (word 0 holds the row count as an int), then the segment lists
static __device__ __align__(256) unsigned long long g_noalloc_scratch[32 + NOALLOC_LIST_WORDS];

__global__ void knn_noalloc_set_count(int* count, int v) { *count = v; }

// merge of the segment lists; output row = position in the row list (the existing merge kernel writes to row id - q0)
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32) knn_simt_merge_list_kernel(const uint64_t* __restrict__ tmp, int S, int k, int cap, int metric,
                                                                         int l2_squared, const int64_t* __restrict__ rowlist,
                                                                         const int* __restrict__ rowcount, int64_t* __restrict__ ids,
                                                                         float* __restrict__ dist) {
    extern __shared__ uint64_t smem_keys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t* sm = smem_keys + static_cast<size_t>(warp) * cap;
    const int64_t nq = *rowcount;
    for (int64_t rr = static_cast<int64_t>(blockIdx.x) * WARPS + warp; rr < nq; rr += static_cast<int64_t>(gridDim.x) * WARPS) {
        const int64_t i = rowlist[rr];
        const uint64_t* src = tmp + rr * S * k;
        const int total = S * k;
        int cnt = 0;
        uint64_t kth = ~0ull;
        for (int base = 0; base < total; base += 32) {
            const int t = base + lane;
            const uint64_t key = t < total ? src[t] : ~0ull;
            const bool pass = key != ~0ull && key <= kth;
            const unsigned m = __ballot_sync(0xffffffffu, pass);
            if (m) {
                if (pass) sm[cnt + __popc(m & ((1u << lane) - 1u))] = key;
                cnt += __popc(m);
                if (cnt + 32 > cap) {
                    for (int u = cnt + lane; u < cap; u += 32) sm[u] = ~0ull;
                    __syncwarp();
                    warp_bitonic_sort(sm, cap, lane);
                    cnt = min(cnt, k);
                    kth = (cnt == k) ? sm[k - 1] : ~0ull;
                }
            }
        }
        for (int t = cnt + lane; t < cap; t += 32) sm[t] = ~0ull;
        __syncwarp();
        warp_bitonic_sort(sm, cap, lane);
        cnt = min(cnt, k);
        write_row(sm, cnt, k, i, metric, l2_squared, ids + rr * k, dist + rr * k, lane);
        __syncwarp();
    }
}

namespace {
// device address of the scratch array per device ordinal (the array's name is a host-side stand-in in host code)
constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned long long* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

int noalloc_scratch_ptr(unsigned long long** out) {
    int dev = -1;
    SPHB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(SPHB_ERR_UNSUPPORTED, "knn_rows_noalloc: device ordinal out of range");
    if (!g_noalloc_ptr[dev]) {
        void* p = nullptr;
        SPHB_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));  // once per device, with that device current
        g_noalloc_ptr[dev] = static_cast<unsigned long long*>(p);
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}
}  // namespace

}  // namespace sphb

using namespace sphb;

extern "C" SPHB_API int sphb_knn_rows_noalloc_dev(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, const int64_t* d_rows,
                                                  int64_t nrows, int64_t* d_ids, float* d_dist, void* stream) {
    if (int rc = ensure_init()) return rc;
    if (!d_x || !d_rows || !d_ids || !d_dist || n <= 0 || d <= 0 || k <= 0 || nrows < 0) return fail(SPHB_ERR_INVALID, "knn_rows_noalloc: bad arguments");
    if (metric != SPHB_METRIC_L2 && metric != SPHB_METRIC_INNER_PRODUCT) return fail(SPHB_ERR_INVALID, "knn_rows_noalloc: unknown metric");
    if (nrows > NOALLOC_MAX_ROWS) return fail(SPHB_ERR_CAPACITY, "knn_rows_noalloc: more than 256 query rows per call (fixed scratch); split the call");
    if (k > NOALLOC_MAX_K) return fail(SPHB_ERR_CAPACITY, "knn_rows_noalloc: k > 128 exceeds the fixed scratch; use sphb_knn_exact_dev");
    if (n >= (1ll << 32) - 1) return fail(SPHB_ERR_UNSUPPORTED, "knn_rows_noalloc: n must fit in uint32");
    if (nrows == 0) return SPHB_OK;
    unsigned long long* scratch = nullptr;
    if (int rc = noalloc_scratch_ptr(&scratch)) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    constexpr int WARPS = 4;
    int cap = 1;
    while (cap < k + 64) cap <<= 1;  // <= 256: 8 KB of dynamic shared memory, below the 48 KB default limit
    const size_t smem = static_cast<size_t>(WARPS) * cap * sizeof(uint64_t);
    int S = static_cast<int>(std::min<int64_t>(NOALLOC_SEGMENTS, std::max<int64_t>(1, n / 1024)));
    const int64_t seglen = ((n + S - 1) / S + 31) / 32 * 32;
    int* d_count = reinterpret_cast<int*>(scratch);
    uint64_t* lists = reinterpret_cast<uint64_t*>(scratch + 32);
    knn_noalloc_set_count<<<1, 1, 0, s>>>(d_count, static_cast<int>(nrows));
    const int64_t b1 = std::min<int64_t>((nrows * S + WARPS - 1) / WARPS, static_cast<int64_t>(ctx().num_sms) * 16);
    knn_simt_seg_kernel<WARPS, 0><<<static_cast<unsigned>(b1), WARPS * 32, smem, s>>>(d_x, n, d, k, cap, metric, d_rows, d_count, S, seglen, lists);
    const int64_t b2 = (nrows + WARPS - 1) / WARPS;
    knn_simt_merge_list_kernel<WARPS><<<static_cast<unsigned>(b2), WARPS * 32, smem, s>>>(lists, S, k, cap, metric, l2_squared, d_rows, d_count, d_ids, d_dist);
    count_launch(3);
    SPHB_CUDA(cudaGetLastError());
    return SPHB_OK;
}
