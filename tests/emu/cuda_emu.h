// cuda_emu.h -- TEST INFRASTRUCTURE ONLY: a tiny SIMT emulation shim so that the CUDA source of the
// simulator (quadruped_rl_locomotion_b200/csrc/go1sim.cu) can be compiled with g++ and executed on the
// CPU.  Every CUDA thread of a block is a FIBER (hand-rolled x86-64 context switch) on one host thread;
// warp collectives (__shfl*_sync, __ballot_sync, __any_sync, __syncwarp) and __syncthreads are a
// cooperative barrier over the block's fibers.  Blocks are distributed over a pool of host threads.
//
// Purpose: this container has no GPU; the emulator lets `pytest -m "not gpu"` run the SAME kernel
// source through the SAME C ABI against the oracle, so that GPU minutes are spent on measurement, not
// on finding indexing bugs.  It is never built into, loaded by, or reachable from the product package
// (quadruped_rl_locomotion_b200 only ever dlopens libgo1sim.so built by nvcc) -- it is not a CPU
// fallback.  Restrictions: blockDim <= 32 (a block is a single warp), collectives are executed by ALL
// live threads of the block (warp-uniform control flow at every collective; true for k_step).
#pragma once
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#if !defined(__x86_64__)
#error "cuda_emu.h: the fiber switch is written for x86-64"
#endif

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __launch_bounds__(...)
#define __grid_constant__
#define __align__(x) alignas(x)
#define __shared__ static thread_local

struct EmuDim { unsigned x = 0, y = 0, z = 0; };
inline thread_local EmuDim threadIdx, blockIdx, blockDim, gridDim;

// ---- fiber switch: saves callee-saved registers on the current stack, swaps stack pointers
extern "C" void go1_emu_switch(void** save_sp, void* load_sp);
asm(R"(
.text
.globl go1_emu_switch
.type go1_emu_switch,@function
go1_emu_switch:
    pushq %rbp
    pushq %rbx
    pushq %r12
    pushq %r13
    pushq %r14
    pushq %r15
    movq %rsp, (%rdi)
    movq %rsi, %rsp
    popq %r15
    popq %r14
    popq %r13
    popq %r12
    popq %rbx
    popq %rbp
    ret
.size go1_emu_switch,.-go1_emu_switch
)");

struct EmuBlock {
    static constexpr int MAXT = 32;
    static constexpr size_t STACK = 512 * 1024;
    int n = 0, live = 0, arrived = 0, cur = -1;
    unsigned gen = 0;
    void* sp[MAXT] = {};
    void* main_sp = nullptr;
    bool done[MAXT] = {};
    unsigned char* stacks = nullptr;
    std::function<void()> body;
    alignas(16) unsigned char slots[MAXT][16];
    alignas(16) unsigned char* dyn_smem = nullptr;
    EmuBlock() {
        stacks = (unsigned char*)std::aligned_alloc(64, STACK * MAXT);
        dyn_smem = (unsigned char*)std::aligned_alloc(64, 1 << 18);
    }
    ~EmuBlock() { std::free(stacks); std::free(dyn_smem); }
};
inline thread_local EmuBlock* g_emu_blk = nullptr;
#define g_emu_block (*g_emu_blk)

inline int emu_next_live(EmuBlock& b, int from) {
    for (int k = 1; k <= b.n; k++) { int t = (from + k) % b.n; if (!b.done[t]) return t; }
    return -1;
}
inline void emu_switch_to(EmuBlock& b, int to) {
    const int me = b.cur;
    const unsigned my_tid = threadIdx.x;
    b.cur = to;
    go1_emu_switch(&b.sp[me], b.sp[to]);
    threadIdx.x = my_tid;      // fibers share the host thread's thread_local threadIdx
}
// cooperative barrier over all live fibers of the block
inline void emu_barrier() {
    EmuBlock& b = g_emu_block;
    if (b.live <= 1) return;
    if (++b.arrived == b.live) { b.arrived = 0; b.gen++; return; }
    const unsigned g = b.gen;
    while (b.gen == g) {
        int to = emu_next_live(b, b.cur);
        if (to < 0 || to == b.cur) { std::fprintf(stderr, "cuda_emu: barrier deadlock\n"); std::abort(); }
        emu_switch_to(b, to);
    }
}
extern "C" inline void go1_emu_fiber_main() {
    EmuBlock& b = g_emu_block;
    threadIdx.x = (unsigned)b.cur;
    b.body();
    b.done[b.cur] = true;
    b.live--;
    if (b.live > 0 && b.arrived == b.live) { b.arrived = 0; b.gen++; }   // others were waiting for this one
    int to = emu_next_live(b, b.cur);
    void* dummy;
    if (to < 0) go1_emu_switch(&dummy, b.main_sp);
    else { b.cur = to; go1_emu_switch(&dummy, b.sp[to]); }
    std::abort();   // a finished fiber is never resumed
}

// ---- warp collectives (all live fibers of the block participate)
template <typename T> inline T emu_exchange(T v, int src) {
    static_assert(sizeof(T) <= 16, "emu shuffle payload");
    EmuBlock& b = g_emu_block;
    if (b.n == 1) return v;
    std::memcpy(b.slots[threadIdx.x], &v, sizeof(T));
    emu_barrier();
    T r = v;
    if (src >= 0 && src < b.n) std::memcpy(&r, b.slots[src], sizeof(T));
    emu_barrier();
    return r;
}
template <typename T> inline T __shfl_xor_sync(unsigned, T v, int lane_mask, int width = 32) {
    return emu_exchange(v, (int)(threadIdx.x ^ (unsigned)lane_mask));
}
template <typename T> inline T __shfl_sync(unsigned, T v, int src_lane, int width = 32) {
    const int base = ((int)threadIdx.x / width) * width;
    return emu_exchange(v, base + (src_lane % width + width) % width);
}
template <typename T> inline T __shfl_down_sync(unsigned, T v, unsigned delta, int width = 32) {
    const int src = (int)threadIdx.x + (int)delta;
    return emu_exchange(v, (src / width == (int)threadIdx.x / width) ? src : -1);
}
template <typename T> inline T __shfl_up_sync(unsigned, T v, unsigned delta, int width = 32) {
    const int src = (int)threadIdx.x - (int)delta;
    return emu_exchange(v, (src >= 0 && src / width == (int)threadIdx.x / width) ? src : -1);
}
inline unsigned __ballot_sync(unsigned, int pred) {
    EmuBlock& b = g_emu_block;
    int p = pred ? 1 : 0;
    std::memcpy(b.slots[threadIdx.x], &p, sizeof(int));
    emu_barrier();
    unsigned r = 0;
    for (int t = 0; t < b.n; t++) {
        if (b.done[t]) continue;
        int q; std::memcpy(&q, b.slots[t], sizeof(int));
        if (q) r |= 1u << t;
    }
    emu_barrier();
    return r;
}
inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
inline int __all_sync(unsigned m, int pred) {
    EmuBlock& b = g_emu_block;
    unsigned livemask = 0;
    for (int t = 0; t < b.n; t++) if (!b.done[t]) livemask |= 1u << t;
    return (__ballot_sync(m, pred) & livemask) == livemask;
}
inline void __syncthreads() { emu_barrier(); }
inline int __syncthreads_and(int pred) { return __all_sync(0xffffffffu, pred); }   // a block is one warp here
inline int __syncthreads_or(int pred) { return __any_sync(0xffffffffu, pred); }
inline void __syncwarp(unsigned = 0xffffffffu) { emu_barrier(); }
inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline int __clz(int v) { return v == 0 ? 32 : __builtin_clz((unsigned)v); }
inline int __float_as_int(float f) { int i; std::memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
inline void sincospif(float x, float* s, float* c) { *s = std::sin(3.14159265358979323846f * x); *c = std::cos(3.14159265358979323846f * x); }
inline void sincosf(float x, float* s, float* c) { *s = std::sin(x); *c = std::cos(x); }
inline void sincos(double x, double* s, double* c) { *s = std::sin(x); *c = std::cos(x); }
inline std::mutex g_emu_atomic_mutex;
inline double atomicAdd(double* p, double v) { std::lock_guard<std::mutex> l(g_emu_atomic_mutex); double o = *p; *p = o + v; return o; }
inline float atomicAdd(float* p, float v) { std::lock_guard<std::mutex> l(g_emu_atomic_mutex); float o = *p; *p = o + v; return o; }
inline int atomicAdd(int* p, int v) { std::lock_guard<std::mutex> l(g_emu_atomic_mutex); int o = *p; *p = o + v; return o; }

// ---- runtime API subset (all "device" memory is host memory)
typedef int cudaError_t;
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
enum { cudaSuccess = 0 };
enum { cudaHostRegisterDefault = 0 };
enum { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaStreamNonBlocking = 1,
       cudaFuncAttributeMaxDynamicSharedMemorySize = 8, cudaFuncAttributePreferredSharedMemoryCarveout = 9 };
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
inline cudaError_t cudaHostRegister(void*, size_t, unsigned) { return 0; }
inline cudaError_t cudaHostUnregister(void*) { return 0; }
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

// runs one block: `block` fibers on the calling host thread
inline void emu_run_block(EmuBlock& B, int b, int grid, int block, const std::function<void()>& body) {
    g_emu_blk = &B;
    B.n = block; B.live = block; B.arrived = 0; B.gen = 0; B.body = body;
    blockIdx.x = (unsigned)b; blockDim.x = (unsigned)block; gridDim.x = (unsigned)grid;
    for (int t = 0; t < block; t++) {
        B.done[t] = false;
        unsigned char* top = B.stacks + (size_t)(t + 1) * EmuBlock::STACK;
        void** sp = (void**)(top - 64);
        for (int k = 0; k < 6; k++) sp[k] = nullptr;
        sp[6] = (void*)&go1_emu_fiber_main;
        sp[7] = nullptr;
        B.sp[t] = sp;
    }
    B.cur = 0;
    go1_emu_switch(&B.main_sp, B.sp[0]);
}

template <typename K, typename... Args>
inline void emu_launch(K kernel, int grid, int block, Args... args) {
    if (block > EmuBlock::MAXT) std::abort();
    const std::function<void()> body = [=]() { kernel(args...); };
    int nth = (int)std::thread::hardware_concurrency();
    if (nth < 1) nth = 1;
    if (nth > grid) nth = grid;
    if (const char* e = std::getenv("GO1_EMU_THREADS")) nth = std::max(1, std::min(grid, std::atoi(e)));
    std::atomic<int> next{0};
    auto worker = [&]() {
        EmuBlock* B = new EmuBlock();
        for (;;) {
            int b = next.fetch_add(1);
            if (b >= grid) break;
            emu_run_block(*B, b, grid, block, body);
        }
        g_emu_blk = nullptr;
        delete B;
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nth; t++) th.emplace_back(worker);
    worker();
    for (auto& x : th) x.join();
}
