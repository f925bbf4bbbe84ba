// cuda_emu.h -- TEST INFRASTRUCTURE ONLY: a tiny SIMT emulation shim so that the CUDA source of the
// simulator (quadruped_rl_locomotion_b200/csrc/go1sim.cu) can be compiled with g++ and executed on the
// CPU, one host thread per CUDA thread of a block, blocks run one after another.  Warp shuffles and
// __syncthreads are implemented with a spin barrier over the block's threads.
//
// Purpose: this container has no GPU; the emulator lets `pytest -m "not gpu"` run the SAME kernel
// source through the SAME C ABI against the oracle, so that GPU minutes are spent on measurement, not
// on finding indexing bugs.  It is never built into, loaded by, or reachable from the product package
// (quadruped_rl_locomotion_b200 only ever dlopens libgo1sim.so built by nvcc) -- it is not a CPU
// fallback.  Restrictions: blockDim <= 32 (a block is a single "warp"), shuffles use lane ids
// relative to the block, every thread of the block must reach every barrier (true for k_step with
// GO1_BLOCK=4: one 4-lane tile per block, tile-uniform control flow).
#pragma once
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __launch_bounds__(...)
#define __align__(x) alignas(x)
#define __shared__ static

struct EmuDim { unsigned x = 0, y = 0, z = 0; };
inline thread_local EmuDim threadIdx, blockIdx, blockDim, gridDim;

struct EmuBlock {
    int n = 1;
    std::atomic<int> count{0};
    std::atomic<int> gen{0};
    alignas(16) unsigned char slots[32][16];
    alignas(16) unsigned char dyn_smem[1 << 17];
    void wait() {
        if (n == 1) return;
        int g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
            count.store(0, std::memory_order_relaxed);
            gen.fetch_add(1, std::memory_order_release);
        } else {
            int spins = 0;
            while (gen.load(std::memory_order_acquire) == g) {
                if (++spins > 2000) { std::this_thread::yield(); spins = 0; }
            }
        }
    }
};
inline EmuBlock g_emu_block;

template <typename T> inline T __shfl_xor_sync(unsigned, T v, int lane_mask) {
    static_assert(sizeof(T) <= 16, "emu shuffle payload");
    EmuBlock& b = g_emu_block;
    if (b.n == 1) return v;
    std::memcpy(b.slots[threadIdx.x], &v, sizeof(T));
    b.wait();
    T r;
    std::memcpy(&r, b.slots[(threadIdx.x ^ (unsigned)lane_mask) % (unsigned)b.n], sizeof(T));
    b.wait();
    return r;
}
inline void __syncthreads() { g_emu_block.wait(); }
inline void __syncwarp(unsigned = 0xffffffffu) { g_emu_block.wait(); }
inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
inline void sincospif(float x, float* s, float* c) { *s = std::sin(3.14159265358979323846f * x); *c = std::cos(3.14159265358979323846f * x); }
inline void sincosf(float x, float* s, float* c) { *s = std::sin(x); *c = std::cos(x); }
inline void sincos(double x, double* s, double* c) { *s = std::sin(x); *c = std::cos(x); }
inline std::mutex g_emu_atomic_mutex;
inline double atomicAdd(double* p, double v) { std::lock_guard<std::mutex> l(g_emu_atomic_mutex); double o = *p; *p = o + v; return o; }

// ---- runtime API subset (all "device" memory is host memory)
typedef int cudaError_t;
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
enum { cudaSuccess = 0 };
enum { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaStreamNonBlocking = 1, cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
inline cudaError_t cudaSetDevice(int) { return 0; }
inline cudaError_t cudaGetLastError() { return 0; }
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, int) { *s = (void*)1; return 0; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = (void*)1; return 0; }
inline cudaError_t cudaEventDestroy(cudaEvent_t) { return 0; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return 0; }
inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0; return 0; }
template <typename T> inline cudaError_t cudaMalloc(T** p, size_t n) { *p = (T*)std::malloc(n ? n : 1); return *p ? 0 : 2; }
template <typename T> inline cudaError_t cudaMallocHost(T** p, size_t n) { *p = (T*)std::malloc(n ? n : 1); return *p ? 0 : 2; }
inline cudaError_t cudaFree(void* p) { std::free(p); return 0; }
inline cudaError_t cudaFreeHost(void* p) { std::free(p); return 0; }
inline cudaError_t cudaMemset(void* p, int v, size_t n) { std::memset(p, v, n); return 0; }
inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return 0; }
inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, int) { std::memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, int, cudaStream_t) { std::memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpy2D(void* d, size_t dp, const void* s, size_t sp, size_t w, size_t h, int) {
    for (size_t r = 0; r < h; r++) std::memcpy((char*)d + r * dp, (const char*)s + r * sp, w);
    return 0;
}
template <typename F> inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }

template <typename K, typename... Args>
inline void emu_launch(K kernel, int grid, int block, Args... args) {
    if (block > 32) std::abort();
    for (int b = 0; b < grid; b++) {
        g_emu_block.n = block;
        g_emu_block.count.store(0);
        std::vector<std::thread> th;
        auto body = [&](int t) {
            threadIdx.x = (unsigned)t; blockIdx.x = (unsigned)b; blockDim.x = (unsigned)block; gridDim.x = (unsigned)grid;
            kernel(args...);
        };
        for (int t = 1; t < block; t++) th.emplace_back(body, t);
        body(0);
        for (auto& x : th) x.join();
    }
}
