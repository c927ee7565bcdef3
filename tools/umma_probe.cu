// tools/umma_probe.cu — B200 microbenchmark behind the Stage-A sweep design.
//   part 1: tcgen05.mma with the A operand in TENSOR MEMORY (written by tcgen05.st, lane = row, 32-bit
//           column c = K elements 2c, 2c+1) checked against the shared-memory form and a scalar evaluation;
//   part 2: MMA rate of the sweep's tile shapes (M = 128 per half, two halves, K = 8 data k-steps + 1 norm
//           k-step) with A in shared memory (SS) or tensor memory (TS), with and without the database
//           stream landing in the shared-memory ring (cp.async.bulk) and with epilogue warps reading TMEM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/build/umma_probe tools/umma_probe.cu
#include "../sph_b200/csrc/ptx_sm100.cuh"

#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace sphb;

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__);     \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// byte offset of element (row, k) of a K-major [rows, 64] fp16 tile in the SWIZZLE_128B layout
__device__ __forceinline__ uint32_t sw128_off(int row, int k) {
    return (row >> 3) * 1024 + (row & 7) * 128 + ((((k * 2) >> 4) ^ (row & 7)) << 4) + ((k * 2) & 15);
}

__device__ __forceinline__ float a_val(int r, int k) { return static_cast<float>((r * 7 + k * 3) % 5 - 2); }
__device__ __forceinline__ float b_val(int n, int k) { return static_cast<float>((n * 5 + k) % 7 - 3); }

// ---------------------------------------------------------------- part 1: correctness of the TS form
template <int N>
__global__ void __launch_bounds__(128) probe_correct(int* out /* [0] ss mismatches, [1] ts mismatches, [2..] decode */) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sA = smem;                  // 128 x 64 fp16
    uint8_t* sB = smem + 16384;          // N x 64 fp16
    uint8_t* sB2 = sB + N * 128;         // decode tile: B2[n][k] = k + 1 (k < 16)
    uint64_t* bar = reinterpret_cast<uint64_t*>(sB2 + N * 128);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < 128 * 64; i += 128) {
        const int r = i >> 6, k = i & 63;
        *reinterpret_cast<__half*>(sA + sw128_off(r, k)) = __float2half(a_val(r, k));
    }
    for (int i = tid; i < N * 64; i += 128) {
        const int n = i >> 6, k = i & 63;
        *reinterpret_cast<__half*>(sB + sw128_off(n, k)) = __float2half(b_val(n, k));
        *reinterpret_cast<__half*>(sB2 + sw128_off(n, k)) = __float2half(k < 16 ? static_cast<float>(k + 1) : 0.f);
    }
    if (tid == 0) {
        ptx::mbar_init(ptx::smem_u32(bar), 1);
        ptx::fence_barrier_init();
    }
    if (warp == 0) {
        ptx::tmem_alloc(ptx::smem_u32(slot), 512);
        ptx::tmem_relinquish();
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> visible to the MMA (async proxy)
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = *slot;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    const int row = warp * 32 + lane;
    // A -> TMEM columns 448..479 (64 fp16 of K per lane)
    for (int c8 = 0; c8 < 4; ++c8) {
        uint32_t r[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int k = (c8 * 8 + c) * 2;
            const __half2 h = __floats2half2_rn(a_val(row, k), a_val(row, k + 1));  // .x = low half = even k
            r[c] = *reinterpret_cast<const uint32_t*>(&h);
        }
        tmem_st_x8(tb + lane_base + 448 + c8 * 8, r);
    }
    tmem_st_wait();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    constexpr uint32_t idesc = ptx::make_idesc_f16(128, N);
    uint32_t phase = 0;
    if (warp == 0) {
        if (ptx::elect_one_sync()) {
            const uint64_t ad = ptx::make_smem_desc_sw128(ptx::smem_u32(sA)), bd = ptx::make_smem_desc_sw128(ptx::smem_u32(sB));
            for (int kk = 0; kk < 4; ++kk) ptx::umma_f16(tb + 0, ad + 2 * kk, bd + 2 * kk, idesc, kk != 0);
            for (int kk = 0; kk < 4; ++kk) umma_f16_ts(tb + 256, tb + 448 + kk * 8, bd + 2 * kk, idesc, kk != 0);
            ptx::umma_commit(ptx::smem_u32(bar));
        }
        __syncwarp();
    }
    ptx::mbar_wait(ptx::smem_u32(bar), phase);
    phase ^= 1;
    ptx::tc_fence_after();
    int bad_ss = 0, bad_ts = 0;
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v0[16], v1[16];
        tmem_ld_x16(tb + lane_base + c0, v0);
        tmem_ld_x16(tb + lane_base + 256 + c0, v1);
        ptx::tmem_ld_wait();
        for (int c = 0; c < 16; ++c) {
            float e = 0.f;
            for (int k = 0; k < 64; ++k) e += a_val(row, k) * b_val(c0 + c, k);
            if (__uint_as_float(v0[c]) != e) ++bad_ss;
            if (__uint_as_float(v1[c]) != e) ++bad_ts;
            if (row < 2 && c0 == 0 && c < 4) out[64 + row * 16 + c * 2] = __float_as_int(__uint_as_float(v1[c])), out[64 + row * 16 + c * 2 + 1] = __float_as_int(e);
        }
    }
    atomicAdd(out + 0, bad_ss);
    atomicAdd(out + 1, bad_ts);
    // decode: which K element does TMEM slot (column c, half h) of a k-step feed, and which row does lane l feed?
    for (int s = 0; s < 16; ++s) {
        ptx::tc_fence_before();
        __syncthreads();
        uint32_t r[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) r[c] = (c == (s >> 1)) ? ((s & 1) ? 0x3c000000u : 0x00003c00u) : 0u;  // fp16 1.0 in the high / low half
        tmem_st_x8(tb + lane_base + 448, r);
        tmem_st_wait();
        ptx::tc_fence_before();
        __syncthreads();
        ptx::tc_fence_after();
        if (warp == 0) {
            if (ptx::elect_one_sync()) {
                umma_f16_ts(tb + 256, tb + 448, ptx::make_smem_desc_sw128(ptx::smem_u32(sB2)), idesc, 0);
                ptx::umma_commit(ptx::smem_u32(bar));
            }
            __syncwarp();
        }
        ptx::mbar_wait(ptx::smem_u32(bar), phase);
        phase ^= 1;
        ptx::tc_fence_after();
        uint32_t v[16];
        tmem_ld_x16(tb + lane_base + 256, v);
        ptx::tmem_ld_wait();
        if (tid == 0) out[2 + s] = static_cast<int>(__uint_as_float(v[0])) - 1;  // k fed by slot s
    }
    {   // lane -> row: slot 0 of lane l holds l + 1 (as fp16), B2 column 0 weight 1
        ptx::tc_fence_before();
        __syncthreads();
        uint32_t r[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const __half2 h = __floats2half2_rn(static_cast<float>(row + 1), 0.f);
        r[0] = *reinterpret_cast<const uint32_t*>(&h);
        tmem_st_x8(tb + lane_base + 448, r);
        tmem_st_wait();
        ptx::tc_fence_before();
        __syncthreads();
        ptx::tc_fence_after();
        if (warp == 0) {
            if (ptx::elect_one_sync()) {
                umma_f16_ts(tb + 256, tb + 448, ptx::make_smem_desc_sw128(ptx::smem_u32(sB2)), idesc, 0);
                ptx::umma_commit(ptx::smem_u32(bar));
            }
            __syncwarp();
        }
        ptx::mbar_wait(ptx::smem_u32(bar), phase);
        phase ^= 1;
        ptx::tc_fence_after();
        uint32_t v[16];
        tmem_ld_x16(tb + lane_base + 256, v);
        ptx::tmem_ld_wait();
        // expected (row + 1) * B2[0][k(slot 0)] = (row + 1) * (k0 + 1)
        if (lane == 0 || lane == 5) out[32 + warp * 2 + (lane == 5)] = static_cast<int>(__uint_as_float(v[0]));
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc(tb, 512);
}

// ---------------------------------------------------------------- part 2: rate
struct RateParams {
    const uint8_t* gsrc;     // database stream source
    size_t gbytes;
    int tiles;               // database tiles per CTA
    int load;                // 1: the producer streams every tile into the ring (cp.async.bulk)
    int epi;                 // 1: 8 epilogue warps read the accumulators (tcgen05.ld + min tree), unsynchronised
    int stages;
    int stage_bytes;
    long long* cycles;       // [grid] MMA-warp cycles
    float* sink;
};

template <bool TS, int N>
__global__ void __launch_bounds__(320, 1) probe_rate(const RateParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    // [A: 2 halves x 2 chunks x 16 KB (SS only) | ones 4 KB (SS only) | ring]
    constexpr int A_BYTES = TS ? 0 : (4 * 16384 + 4096);
    uint8_t* ring = smem + A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + static_cast<size_t>(P.stages) * P.stage_bytes);
    const uint32_t bar_full = ptx::smem_u32(bars), bar_empty = ptx::smem_u32(bars + 8), bar_done = ptx::smem_u32(bars + 16);
    uint32_t* slot = reinterpret_cast<uint32_t*>(bars + 17);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid * 4; i < A_BYTES + P.stages * P.stage_bytes; i += blockDim.x * 4) *reinterpret_cast<uint32_t*>(smem + i) = 0x3c003800u;  // 1.0, 0.5
    if (tid == 0) {
        for (int s = 0; s < P.stages; ++s) {
            ptx::mbar_init(bar_full + 8 * s, 1);
            ptx::mbar_init(bar_empty + 8 * s, 1);
        }
        ptx::mbar_init(bar_done, 1);
        ptx::fence_barrier_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc(ptx::smem_u32(slot), 512);
        ptx::tmem_relinquish();
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tb = *slot;
    constexpr int ACC1 = (2 * N + 136 <= 512) ? N : 0;  // column of the second half's accumulator
    constexpr int ACOL = 512 - 136;                      // TS: A halves at ACOL, ACOL + 64; ones at ACOL + 128
    if (TS && warp >= 2 && warp < 6) {
        const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
        uint32_t r[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) r[c] = 0x3c003800u;
        for (int c8 = 0; c8 < 17; ++c8) tmem_st_x8(tb + lane_base + ACOL + c8 * 8, r);
        tmem_st_wait();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();

    if (warp == 0) {
        if (P.load) {
            int stage = 0;
            uint32_t phase = 0;
            const size_t tile_bytes = static_cast<size_t>(N) * 288;
            size_t off = (static_cast<size_t>(blockIdx.x) * 4 * tile_bytes) % (P.gbytes - tile_bytes);
            for (int t = 0; t < P.tiles; ++t) {
                ptx::mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                if (ptx::elect_one_sync()) {
                    const uint32_t dst = ptx::smem_u32(ring + static_cast<size_t>(stage) * P.stage_bytes);
                    ptx::mbar_expect_tx(bar_full + 8 * stage, N * 288);
                    bulk_g2s(dst, P.gsrc + off, N * 128, bar_full + 8 * stage);
                    bulk_g2s(dst + N * 128, P.gsrc + off + N * 128, N * 128, bar_full + 8 * stage);
                    bulk_g2s(dst + N * 256, P.gsrc + off + N * 256, N * 32, bar_full + 8 * stage);
                }
                __syncwarp();
                off += tile_bytes;
                if (off + tile_bytes > P.gbytes) off = 0;
                if (++stage == P.stages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc = ptx::make_idesc_f16(128, N);
        int stage = 0;
        uint32_t phase = 0;
        const long long t0 = clock64();
        for (int t = 0; t < P.tiles; ++t) {
            if (P.load) {
                ptx::mbar_wait(bar_full + 8 * stage, phase);
                ptx::tc_fence_after();
            }
            if (ptx::elect_one_sync()) {
                const uint32_t sb = ptx::smem_u32(ring + static_cast<size_t>(stage) * P.stage_bytes);
                const uint64_t b0 = ptx::make_smem_desc_sw128(sb), b1 = ptx::make_smem_desc_sw128(sb + N * 128), bn = ptx::make_smem_desc_sw32(sb + N * 256);
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const uint32_t d = tb + (m ? ACC1 : 0);
                    if constexpr (TS) {
                        const uint32_t a = tb + ACOL + m * 64;
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) umma_f16_ts(d, a + kk * 8, b0 + 2 * kk, idesc, kk != 0);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) umma_f16_ts(d, a + 32 + kk * 8, b1 + 2 * kk, idesc, 1);
                        umma_f16_ts(d, tb + ACOL + 128, bn, idesc, 1);
                    } else {
                        const uint32_t sa = ptx::smem_u32(smem) + m * 32768;
                        const uint64_t a0 = ptx::make_smem_desc_sw128(sa), a1 = ptx::make_smem_desc_sw128(sa + 16384);
                        const uint64_t ao = ptx::make_smem_desc_sw32(ptx::smem_u32(smem) + 65536);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) ptx::umma_f16(d, a0 + 2 * kk, b0 + 2 * kk, idesc, kk != 0);
#pragma unroll
                        for (int kk = 0; kk < 4; ++kk) ptx::umma_f16(d, a1 + 2 * kk, b1 + 2 * kk, idesc, 1);
                        ptx::umma_f16(d, ao, bn, idesc, 1);
                    }
                }
                if (P.load) ptx::umma_commit(bar_empty + 8 * stage);
                if (t == P.tiles - 1) ptx::umma_commit(bar_done);
            }
            __syncwarp();
            if (P.load) {
                if (++stage == P.stages) { stage = 0; phase ^= 1; }
            }
        }
        ptx::mbar_wait(bar_done, 0);
        if (lane == 0) P.cycles[blockIdx.x] = clock64() - t0;
    } else if (P.epi) {
        const int ew = warp - 2, m = ew >> 2;
        const uint32_t taddr = tb + (static_cast<uint32_t>((warp & 3) * 32) << 16) + (m ? ACC1 : 0);
        float acc = 3.0e38f;
        for (int t = 0; t < P.tiles; ++t) {
#pragma unroll 1
            for (int c0 = 0; c0 < N; c0 += 32) {
                uint32_t v0[16], v1[16];
                tmem_ld_x16(taddr + c0, v0);
                tmem_ld_x16(taddr + (c0 + 16 < N ? c0 + 16 : c0), v1);
                ptx::tmem_ld_wait();
                float mn = 3.0e38f;
#pragma unroll
                for (int c = 0; c < 16; ++c) mn = fminf(mn, fminf(__uint_as_float(v0[c]), __uint_as_float(v1[c])));
                acc = fminf(acc, mn);
            }
        }
        if (acc == 12345.f) P.sink[tid] = acc;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) ptx::tmem_dealloc(tb, 512);
}

template <bool TS, int N>
static void run_rate(const char* name, const uint8_t* gsrc, size_t gbytes, int tiles, int load, int epi, long long* d_cycles, float* d_sink, int sms) {
    constexpr int A_BYTES = TS ? 0 : (4 * 16384 + 4096);
    const int stage_bytes = (N * 288 + 1023) / 1024 * 1024;
    int stages = std::min(6, (226 * 1024 - A_BYTES - 1024 - 512) / stage_bytes);
    if (stages < 1) { printf("%-28s does not fit\n", name); return; }
    if (load && stages < 2) { printf("%-28s ring too small\n", name); return; }
    const size_t smem = A_BYTES + static_cast<size_t>(stages) * stage_bytes + 1024 + 512;
    CK(cudaFuncSetAttribute(probe_rate<TS, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    RateParams P{gsrc, gbytes, tiles, load, epi, stages, stage_bytes, d_cycles, d_sink};
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    probe_rate<TS, N><<<sms, 320, smem>>>(P);  // warm-up
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    probe_rate<TS, N><<<sms, 320, smem>>>(P);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    std::vector<long long> h(sms);
    CK(cudaMemcpy(h.data(), d_cycles, sms * sizeof(long long), cudaMemcpyDeviceToHost));
    double mean = 0, mx = 0;
    for (long long c : h) { mean += static_cast<double>(c) / sms; mx = std::max(mx, static_cast<double>(c)); }
    const double floor_cyc = 2.0 * 9.0 * (N / 2.0);  // per tile: 2 halves x 9 k-steps x (128 * N / 256) cycles
    const double flops = 2.0 * 256.0 * N * 144.0 * tiles * sms;
    printf("%-28s N=%3d load=%d epi=%d stages=%d | cycles/tile mean %.0f max %.0f | floor %.0f -> tensor util %.1f %% (mean) | %.2f ms, %.0f TFLOP/s executed, clock %.2f GHz\n",
           name, N, load, epi, stages, mean / tiles, mx / tiles, floor_cyc, 100.0 * floor_cyc * tiles / mean, ms, flops / (ms * 1e-3) / 1e12,
           mx / (ms * 1e-3) / 1e9);
    fflush(stdout);
}

int main(int argc, char** argv) {
    const int tiles = argc > 1 ? atoi(argv[1]) : 4096;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    printf("device %s, %d SMs\n", prop.name, sms);
    int* d_out;
    CK(cudaMalloc(&d_out, 256 * sizeof(int)));
    auto correct = [&](auto kern, int N) {
        CK(cudaMemset(d_out, 0xff, 256 * sizeof(int)));
        CK(cudaMemset(d_out, 0, 2 * sizeof(int)));
        const size_t smem = 16384 + 2 * N * 128 + 1024 + 256;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
        kern<<<1, 128, smem>>>(d_out);
        CK(cudaDeviceSynchronize());
        int h[256];
        CK(cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost));
        printf("correct N=%d: SS mismatches %d, TS mismatches %d (of %d)\n  slot->k:", N, h[0], h[1], 128 * N);
        for (int s = 0; s < 16; ++s) printf(" %d", h[2 + s]);
        printf("\n  lane->row value (warp w lane 0, lane 5; expect (row+1)*(k0+1)):");
        for (int i = 0; i < 8; ++i) printf(" %d", h[32 + i]);
        printf("\n  TS samples (got, expected):");
        for (int i = 0; i < 8; ++i) { const int o = 64 + (i >> 2) * 16 + (i & 3) * 2; float a, b; memcpy(&a, &h[o], 4); memcpy(&b, &h[o + 1], 4); printf(" (%g,%g)", a, b); }
        printf("\n");
        fflush(stdout);
    };
    correct(probe_correct<128>, 128);
    correct(probe_correct<80>, 80);
    correct(probe_correct<96>, 96);

    const size_t gbytes = 512ull << 20;
    uint8_t* gsrc;
    CK(cudaMalloc(&gsrc, gbytes));
    CK(cudaMemset(gsrc, 0x38, gbytes));  // fp16 0x3838 = 0.527
    long long* d_cycles;
    float* d_sink;
    CK(cudaMalloc(&d_cycles, sms * sizeof(long long)));
    CK(cudaMalloc(&d_sink, 1024 * sizeof(float)));
    for (int epi = 0; epi < 2; ++epi) {
        for (int load = 0; load < 2; ++load) {
            run_rate<false, 128>("SS M128xN128 (current)", gsrc, gbytes, tiles, load, epi, d_cycles, d_sink, sms);
            run_rate<false, 256>("SS M128xN256", gsrc, gbytes, tiles / 2, load, epi, d_cycles, d_sink, sms);
            run_rate<false, 64>("SS M128xN64", gsrc, gbytes, tiles * 2, load, epi, d_cycles, d_sink, sms);
            run_rate<true, 128>("TS M128xN128", gsrc, gbytes, tiles, load, epi, d_cycles, d_sink, sms);
            run_rate<true, 96>("TS M128xN96", gsrc, gbytes, tiles, load, epi, d_cycles, d_sink, sms);
            run_rate<true, 80>("TS M128xN80", gsrc, gbytes, tiles, load, epi, d_cycles, d_sink, sms);
            run_rate<true, 64>("TS M128xN64", gsrc, gbytes, tiles * 2, load, epi, d_cycles, d_sink, sms);
            run_rate<true, 176>("TS M128xN176", gsrc, gbytes, tiles / 2, load, epi, d_cycles, d_sink, sms);
        }
    }
    printf("done\n");
    return 0;
}
