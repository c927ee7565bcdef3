// Helpers shared by the CSR graph passes (Stage B6 / C / D): CUB-backed sort / scan wrappers that
// allocate their temporary storage stream-ordered, row lookup by binary search, Philox4x32-10.
#pragma once
#include "common.cuh"

#include <cub/cub.cuh>

namespace sphb {

inline int bits_for(uint64_t v) {  // number of bits needed to represent values < v
    int b = 0;
    while (b < 64 && (v >> b) != 0) ++b;
    return b;
}

// Stable LSD radix sort of (key, value) pairs on bits [0, end_bit).
template <class V>
int sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const V* vals_in, V* vals_out, int64_t num, int end_bit, cudaStream_t s) {
    if (num <= 0) return 0;
    if (num > 0x7fffffffll) return fail(5, "sort_pairs: more than 2^31 - 1 records");
    size_t bytes = 0;
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(num), 0, end_bit, s));
    DevBuf<uint8_t> tmp;
    SPHB_CUDA(tmp.alloc(bytes, s));
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(num), 0, end_bit, s));
    count_launch(3);
    return 0;
}

// The same on bits [begin_bit, end_bit) only (input already ordered on the lower bits: LSD sorts are stable).
template <class V>
int sort_pairs_bits(const uint64_t* keys_in, uint64_t* keys_out, const V* vals_in, V* vals_out, int64_t num, int begin_bit, int end_bit, cudaStream_t s) {
    if (num <= 0) return 0;
    if (num > 0x7fffffffll) return fail(5, "sort_pairs: more than 2^31 - 1 records");
    size_t bytes = 0;
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(num), begin_bit, end_bit, s));
    DevBuf<uint8_t> tmp;
    SPHB_CUDA(tmp.alloc(bytes, s));
    SPHB_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, bytes, keys_in, keys_out, vals_in, vals_out, static_cast<int>(num), begin_bit, end_bit, s));
    count_launch(3);
    return 0;
}

template <class T>
int exclusive_sum(const T* in, T* out, int64_t num, cudaStream_t s) {
    if (num <= 0) return 0;
    if (num > 0x7fffffffll) return fail(5, "exclusive_sum: more than 2^31 - 1 items");
    size_t bytes = 0;
    SPHB_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, static_cast<int>(num), s));
    DevBuf<uint8_t> tmp;
    SPHB_CUDA(tmp.alloc(bytes, s));
    SPHB_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, bytes, in, out, static_cast<int>(num), s));
    count_launch(2);
    return 0;
}

#ifdef __CUDACC__
// first index t in [0, num) with keys[t] >= key (keys ascending)
__device__ __forceinline__ int64_t lower_bound_u64(const uint64_t* __restrict__ keys, int64_t num, uint64_t key) {
    int64_t lo = 0, hi = num;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}
// position of column c in the column-sorted row [b, e), or -1
__device__ __forceinline__ int64_t find_col(const uint32_t* __restrict__ col, int64_t b, int64_t e, uint32_t c) {
    int64_t lo = b, hi = e;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (col[mid] < c) lo = mid + 1;
        else hi = mid;
    }
    return (lo < e && col[lo] == c) ? lo : -1;
}

// Philox4x32-10 (Salmon et al., SC'11): counter (c0..c3), key (k0,k1) -> 4 x 32 random bits.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
        k.x += 0x9E3779B9u;
        k.y += 0xBB67AE85u;
    }
    return c;
}
// uniform double in [0, 1) from 53 random bits
__device__ __forceinline__ double u01_from_bits(uint32_t a, uint32_t b) {
    const uint64_t x = (static_cast<uint64_t>(a) << 32) | b;
    return static_cast<double>(x >> 11) * (1.0 / 9007199254740992.0);
}
#endif

}  // namespace sphb
