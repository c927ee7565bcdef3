// TEST INFRASTRUCTURE ONLY — host stand-ins for the three CUB device primitives the graph passes use.
#pragma once
#include <cuda_runtime.h>

#include <numeric>

namespace cub {
namespace detail_emu {
template <class K>
inline K masked(K k, int begin_bit, int end_bit) {
    const int bits = static_cast<int>(sizeof(K) * 8);
    K v = begin_bit >= bits ? 0 : (k >> begin_bit);
    const int width = end_bit - begin_bit;
    if (width < bits) v &= ((static_cast<K>(1) << width) - 1);
    return v;
}
}  // namespace detail_emu
struct DeviceRadixSort {
    template <class K, class V>
    static cudaError_t SortPairs(void* tmp, size_t& bytes, const K* kin, K* kout, const V* vin, V* vout, int num, int begin_bit, int end_bit, cudaStream_t) {
        if (!tmp) { bytes = 16; return cudaSuccess; }
        std::vector<int> order(static_cast<size_t>(num));
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return detail_emu::masked(kin[a], begin_bit, end_bit) < detail_emu::masked(kin[b], begin_bit, end_bit); });
        for (int i = 0; i < num; ++i) { kout[i] = kin[order[i]]; vout[i] = vin[order[i]]; }
        return cudaSuccess;
    }
    template <class K>
    static cudaError_t SortKeys(void* tmp, size_t& bytes, const K* kin, K* kout, int num, int begin_bit, int end_bit, cudaStream_t) {
        if (!tmp) { bytes = 16; return cudaSuccess; }
        std::vector<K> v(kin, kin + num);
        std::stable_sort(v.begin(), v.end(), [&](K a, K b) { return detail_emu::masked(a, begin_bit, end_bit) < detail_emu::masked(b, begin_bit, end_bit); });
        std::copy(v.begin(), v.end(), kout);
        return cudaSuccess;
    }
};
struct DeviceScan {
    template <class I, class O>
    static cudaError_t ExclusiveSum(void* tmp, size_t& bytes, const I* in, O* out, int num, cudaStream_t) {
        if (!tmp) { bytes = 16; return cudaSuccess; }
        O acc = 0;
        for (int i = 0; i < num; ++i) { const O v = static_cast<O>(in[i]); out[i] = acc; acc += v; }
        return cudaSuccess;
    }
};
}  // namespace cub
