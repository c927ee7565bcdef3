// TEST INFRASTRUCTURE ONLY — a minimal CPU stand-in for the CUDA runtime, used to execute the repo's graph kernels
// functionally on a machine without a GPU (tests/emu/README.md).  "Device" memory is host memory; a kernel launch runs
// the blocks one after the other, the threads of a block as cooperatively scheduled fibers, warp collectives and
// __syncthreads as rendezvous points.  It checks kernel LOGIC (indexing, collectives, barriers that every thread reaches);
// it says nothing about performance, memory-model races or sm_100a code generation.
#pragma once
#include <ucontext.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static

typedef void* cudaStream_t;
enum cudaError_t { cudaSuccess = 0, cudaErrorUnknown = 1 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize };
inline const char* cudaGetErrorString(cudaError_t) { return "emulated runtime error"; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) {
    if (n) std::memmove(d, s, n);
    return cudaSuccess;
}
inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t) {
    if (n) std::memset(d, v, n);
    return cudaSuccess;
}
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
typedef void* cudaEvent_t;
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = nullptr; return cudaSuccess; }
inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
template <class F>
inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }

struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
struct float4 { float x, y, z, w; };
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

namespace emu {

struct Idx { unsigned x = 0, y = 0, z = 0; };

struct Fiber {
    ucontext_t ctx;
    std::vector<char> stack;
    bool done = false;
    unsigned tid = 0;
};

struct WarpState {
    uint64_t buf[2][32];
    int arrived = 0;
    unsigned gen = 0;
};

struct BlockState {
    std::vector<Fiber> fibers;
    std::vector<WarpState> warps;
    int bar_arrived = 0;
    unsigned bar_gen = 0;
    unsigned nthreads = 0;
    std::vector<char> dyn_smem;
    ucontext_t sched;
    Fiber* cur = nullptr;
    uint64_t progress = 0;  // bumped whenever a rendezvous completes or a fiber finishes
    const std::function<void()>* body = nullptr;
};

inline BlockState& B() {
    static BlockState b;
    return b;
}
inline Idx& g_threadIdx() { static Idx v; return v; }
inline Idx& g_blockIdx() { static Idx v; return v; }
inline Idx& g_blockDim() { static Idx v; return v; }
inline Idx& g_gridDim() { static Idx v; return v; }

inline void yield() {
    BlockState& b = B();
    Fiber* f = b.cur;
    swapcontext(&f->ctx, &b.sched);
    g_threadIdx().x = f->tid;
}

inline void trampoline() {
    BlockState& b = B();
    (*b.body)();
    b.cur->done = true;
    ++b.progress;
    swapcontext(&b.cur->ctx, &b.sched);
}

inline void* dyn_smem() { return B().dyn_smem.data(); }

// Runs `body` once per thread of a grid x block launch.
inline void launch(unsigned grid, unsigned block, size_t smem, const std::function<void()>& body) {
    BlockState& b = B();
    if (block == 0 || block % 32 != 0 && block > 32) {
        std::fprintf(stderr, "emu: block size %u is not a multiple of 32\n", block);
        std::abort();
    }
    g_gridDim().x = grid;
    g_blockDim().x = block;
    b.nthreads = block;
    if (b.fibers.size() < block) b.fibers.resize(block);
    b.warps.assign((block + 31) / 32, WarpState());
    b.dyn_smem.assign(std::max<size_t>(smem, 16), 0);
    b.body = &body;
    constexpr size_t STACK = 256 << 10;
    for (unsigned blk = 0; blk < grid; ++blk) {
        g_blockIdx().x = blk;
        b.bar_arrived = 0;
        for (auto& w : b.warps) w.arrived = 0;
        for (unsigned t = 0; t < block; ++t) {
            Fiber& f = b.fibers[t];
            if (f.stack.size() < STACK) f.stack.resize(STACK);
            f.done = false;
            f.tid = t;
            getcontext(&f.ctx);
            f.ctx.uc_stack.ss_sp = f.stack.data();
            f.ctx.uc_stack.ss_size = f.stack.size();
            f.ctx.uc_link = &b.sched;
            makecontext(&f.ctx, reinterpret_cast<void (*)()>(&trampoline), 0);
        }
        unsigned remaining = block;
        while (remaining) {
            const uint64_t before = b.progress;
            remaining = 0;
            for (unsigned t = 0; t < block; ++t) {
                Fiber& f = b.fibers[t];
                if (f.done) continue;
                b.cur = &f;
                g_threadIdx().x = t;
                swapcontext(&b.sched, &f.ctx);
                if (!f.done) ++remaining;
            }
            if (remaining && b.progress == before) {
                std::fprintf(stderr, "emu: deadlock in block %u — some threads wait at a barrier / warp collective that the others never reach\n", blk);
                std::abort();
            }
        }
    }
}

// warp-wide exchange: every lane contributes a 64-bit value and sees all 32
inline void warp_exchange(uint64_t mine, uint64_t out[32]) {
    BlockState& b = B();
    const unsigned tid = b.cur->tid, lane = tid & 31;
    WarpState& w = b.warps[tid >> 5];
    const unsigned g = w.gen;
    const int lanes = static_cast<int>(std::min(32u, b.nthreads - (tid & ~31u)));
    w.buf[g & 1][lane] = mine;
    if (++w.arrived == lanes) {
        w.arrived = 0;
        ++w.gen;
        ++b.progress;
    } else {
        while (w.gen == g) yield();
    }
    std::memcpy(out, w.buf[g & 1], sizeof(uint64_t) * 32);
}
inline void block_barrier() {
    BlockState& b = B();
    const unsigned g = b.bar_gen;
    int alive = 0;
    for (unsigned t = 0; t < b.nthreads; ++t) alive += b.fibers[t].done ? 0 : 1;
    if (++b.bar_arrived >= alive) {
        b.bar_arrived = 0;
        ++b.bar_gen;
        ++b.progress;
    } else {
        while (b.bar_gen == g) yield();
    }
}
template <class T>
inline uint64_t to_bits(T v) {
    static_assert(sizeof(T) <= 8, "shuffle of a wide type");
    uint64_t u = 0;
    std::memcpy(&u, &v, sizeof(T));
    return u;
}
template <class T>
inline T from_bits(uint64_t u) {
    T v;
    std::memcpy(&v, &u, sizeof(T));
    return v;
}
inline unsigned lane_id() { return B().cur->tid & 31; }

}  // namespace emu

#define threadIdx (emu::g_threadIdx())
#define blockIdx (emu::g_blockIdx())
#define blockDim (emu::g_blockDim())
#define gridDim (emu::g_gridDim())

inline void __syncthreads() { emu::block_barrier(); }
inline void __syncwarp(unsigned = 0xffffffffu) {
    uint64_t o[32];
    emu::warp_exchange(0, o);
}
template <class T>
inline T __shfl_xor_sync(unsigned, T v, int lane_mask) {
    uint64_t o[32];
    emu::warp_exchange(emu::to_bits(v), o);
    return emu::from_bits<T>(o[(emu::lane_id() ^ static_cast<unsigned>(lane_mask)) & 31]);
}
template <class T>
inline T __shfl_sync(unsigned, T v, int src) {
    uint64_t o[32];
    emu::warp_exchange(emu::to_bits(v), o);
    return emu::from_bits<T>(o[src & 31]);
}
template <class T>
inline T __shfl_up_sync(unsigned, T v, unsigned delta) {
    uint64_t o[32];
    emu::warp_exchange(emu::to_bits(v), o);
    const unsigned l = emu::lane_id();
    return l >= delta ? emu::from_bits<T>(o[l - delta]) : v;
}
inline unsigned __ballot_sync(unsigned, int pred) {
    uint64_t o[32];
    emu::warp_exchange(pred ? 1 : 0, o);
    unsigned m = 0;
    for (int l = 0; l < 32; ++l) m |= o[l] ? (1u << l) : 0u;
    // lanes beyond a partial last warp contribute stale slots: mask them off
    const unsigned tid = emu::B().cur->tid, lanes = std::min(32u, emu::B().nthreads - (tid & ~31u));
    return lanes == 32 ? m : (m & ((1u << lanes) - 1u));
}
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
inline int __all_sync(unsigned m, int pred) {
    const unsigned tid = emu::B().cur->tid, lanes = std::min(32u, emu::B().nthreads - (tid & ~31u));
    const unsigned full = lanes == 32 ? 0xffffffffu : ((1u << lanes) - 1u);
    return __ballot_sync(m, pred) == full;
}

// fast-math intrinsics: glibc declares (but does not export) these names
#define __logf logf
#define __expf expf
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(unsigned v) { return __builtin_ffs(static_cast<int>(v)); }
inline unsigned __float_as_uint(float f) { return emu::from_bits<unsigned>(emu::to_bits(f)); }
inline float __uint_as_float(unsigned u) { return emu::from_bits<float>(u); }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fdiv_rn(float a, float b) { return a / b; }
inline unsigned __umulhi(unsigned a, unsigned b) { return static_cast<unsigned>((static_cast<uint64_t>(a) * b) >> 32); }
template <class T>
inline T __ldg(const T* p) { return *p; }
template <class T>
inline T __ldcg(const T* p) { return *p; }
inline void __threadfence() {}
inline void __threadfence_block() {}
template <class T>
inline unsigned __match_any_sync(unsigned, T v) {
    uint64_t o[32];
    emu::warp_exchange(emu::to_bits(v), o);
    const unsigned tid = emu::B().cur->tid, lanes = std::min(32u, emu::B().nthreads - (tid & ~31u));
    unsigned m = 0;
    for (unsigned l = 0; l < lanes; ++l) m |= (o[l] == o[emu::lane_id()]) ? (1u << l) : 0u;
    return m;
}

// single OS thread: atomics are plain read-modify-writes
template <class T>
inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { unsigned long long o = *p; *p = o + v; return o; }
template <class T>
inline T atomicMax(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <class T>
inline T atomicMin(T* p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <class T>
inline T atomicExch(T* p, T v) { T o = *p; *p = v; return o; }
template <class T>
inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }
template <class T>
inline T atomicAnd(T* p, T v) { T o = *p; *p = o & v; return o; }

using std::max;
using std::min;
