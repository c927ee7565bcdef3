// Exact kNN, SIMT flavour: one warp per query row, fp64 pair keys evaluated
// directly (no GEMM), running top-k kept as sorted u64 keys in shared memory.
// Used where the tensor-core sweep does not pay: tiny n, very small d (the
// 3-channel RGB configuration is FP64/ALU bound) — and as the
// on-device cross-check of the tcgen05 path in the GPU tests.
#include "knn_common.cuh"

namespace sphb {

template <int WARPS>
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
        for (int64_t base = 0; base < n; base += 32) {
            const int64_t j = base + lane;
            bool pass = false;
            uint64_t key = ~0ull;
            if (j < n) {
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

static int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

int knn_simt_launch(const float* d_x, int64_t n, int d, int k, int metric, int l2_squared, int64_t q0, int64_t q1, const int64_t* d_rowlist,
                    const int* d_rowcount, int64_t max_rows, int64_t* d_ids, float* d_dist, cudaStream_t stream) {
    constexpr int WARPS = 4;
    const int cap = next_pow2(k + 64);
    const size_t smem = static_cast<size_t>(WARPS) * cap * sizeof(uint64_t);
    if (smem > 200 * 1024) return fail(5, "knn_simt: k too large");
    SPHB_CUDA(cudaFuncSetAttribute(knn_simt_kernel<WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    const int64_t nq = d_rowlist ? max_rows : q1 - q0;
    if (nq <= 0) return 0;
    int64_t blocks = (nq + WARPS - 1) / WARPS;
    const int64_t maxb = static_cast<int64_t>(ctx().num_sms) * 16;
    if (blocks > maxb) blocks = maxb;
    knn_simt_kernel<WARPS><<<static_cast<unsigned>(blocks), WARPS * 32, smem, stream>>>(d_x, n, d, k, cap, metric, l2_squared, q0, q1, d_rowlist, d_rowcount, d_ids,
                                                                                            d_dist);
    count_launch();
    SPHB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace sphb
