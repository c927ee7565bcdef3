// Shared host/device helpers for the sm_100a back-end.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace sphb {

// ---- host-side context -----------------------------------------------------
struct Context {
    int device = -1;
    cudaStream_t stream = nullptr;
    int num_sms = 0;
    bool ready = false;
};
Context& ctx();
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
void count_launch(int64_t n = 1);
int ensure_init();

#define SPHB_CUDA(call)                                                                            \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            return ::sphb::fail(3, std::string(#call) + ": " + cudaGetErrorString(_e));            \
        }                                                                                          \
    } while (0)

// Workspace cache (api.cu): device blocks released by the library are kept per stream and handed out again for
// requests of a similar size on the same stream (stream order makes the reuse safe).  A call of the kNN path makes
// ~20 stream-ordered allocations of up to several GB; going to cudaMallocAsync for each of them showed sporadic
// stalls of 0.1-2 s (pool growth / remapping) between kernels.  Trimmed by sphb_shutdown and when more than 64 GB idle.
cudaError_t ws_alloc(void** p, size_t bytes, cudaStream_t stream);
void ws_free(void* p, cudaStream_t stream);
void ws_trim();

// RAII stream-ordered device buffer.
template <class T>
struct DevBuf {
    T* p = nullptr;
    cudaStream_t s = nullptr;
    size_t n = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    cudaError_t alloc(size_t count, cudaStream_t stream) {
        release();
        s = stream;
        n = count;
        if (count == 0) return cudaSuccess;
        return ws_alloc(reinterpret_cast<void**>(&p), count * sizeof(T), stream);
    }
    void release() {
        if (p) ws_free(p, s);
        p = nullptr;
    }
    ~DevBuf() { release(); }
};

// ---- device helpers --------------------------------------------------------
#ifdef __CUDACC__

// Order-preserving map float -> uint32 (works for negatives; NaN sorts last-ish).
__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    const uint32_t b = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
    return __uint_as_float(b);
}
__device__ __forceinline__ uint64_t make_key(float v, uint32_t idx) {
    return (static_cast<uint64_t>(f2ord(v)) << 32) | idx;
}
__device__ __forceinline__ float key_val(uint64_t k) { return ord2f(static_cast<uint32_t>(k >> 32)); }
__device__ __forceinline__ uint32_t key_idx(uint64_t k) { return static_cast<uint32_t>(k); }

// Ascending bitonic sort of `cap` (power of two) u64 keys held in shared memory,
// executed by one warp.
__device__ __forceinline__ void warp_bitonic_sort(uint64_t* sm, int cap, int lane) {
    for (int k2 = 2; k2 <= cap; k2 <<= 1) {
        for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
            for (int t = lane; t < (cap >> 1); t += 32) {
                const int i = ((t & ~(j2 - 1)) << 1) | (t & (j2 - 1));
                const int l = i | j2;
                const uint64_t a = sm[i], b = sm[l];
                const bool up = (i & k2) == 0;
                if ((a > b) == up) {
                    sm[i] = b;
                    sm[l] = a;
                }
            }
            __syncwarp();
        }
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

#endif  // __CUDACC__

}  // namespace sphb
