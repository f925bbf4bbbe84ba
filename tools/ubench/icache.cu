// Instruction-fetch microbenchmark (B200): cycles per warp-instruction of a straight-line FFMA loop body of S KB,
// with all warps of a block in lockstep (one fetch stream) or staggered (one stream per warp).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o icache icache.cu && ./icache
#include <cstdio>
#include <cuda_runtime.h>

template <int N> __device__ __forceinline__ void body(float& a, float& b, float& c, float& d, float k, float m) {
#pragma unroll
    for (int i = 0; i < N; i++) { a = fmaf(a, k, m); b = fmaf(b, k, m); c = fmaf(c, k, m); d = fmaf(d, k, m); }
}

// N*4 FFMA per trip = N*64 bytes of code
template <int N> __global__ void __launch_bounds__(768, 1) k(float* out, int trips, int stagger, long long* cyc) {
    float a = threadIdx.x, b = a + 1, c = a + 2, d = a + 3;
    const float kk = 1.0000001f, m = 1e-9f;
    const int warp = threadIdx.x >> 5;
    if (stagger) {   // spread the warps over the loop body: each starts `warp * stagger` cycles later
        const long long t0 = clock64();
        while (clock64() - t0 < (long long)warp * stagger) { }
    }
    const long long t0 = clock64();
#pragma unroll 1
    for (int t = 0; t < trips; t++) body<N>(a, b, c, d, kk, m);
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
    if ((threadIdx.x & 31) == 0) cyc[blockIdx.x * (blockDim.x >> 5) + warp] = t1 - t0;
}

template <int N> void run(int warps, int stagger) {
    float* out; long long* cyc;
    const int blocks = 148;
    cudaMalloc(&out, sizeof(float) * blocks * 768); cudaMalloc(&cyc, sizeof(long long) * blocks * 24);
    const int trips = (1 << 22) / (N * 4) + 1;     // ~4M instructions per warp
    k<N><<<blocks, warps * 32>>>(out, 2, stagger, cyc);
    k<N><<<blocks, warps * 32>>>(out, trips, stagger, cyc);
    cudaDeviceSynchronize();
    static long long h[148 * 24];
    cudaMemcpy(h, cyc, sizeof(long long) * blocks * warps, cudaMemcpyDeviceToHost);
    double s = 0; for (int i = 0; i < blocks * warps; i++) s += (double)h[i];
    const double per_warp = s / (blocks * warps) / ((double)trips * N * 4);
    printf("code %4d KB  warps/SM %2d  stagger %5d : %.3f cycles per warp-instruction -> SM IPC %.2f\n", N * 64 / 1024, warps, stagger,
           per_warp, warps / per_warp);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int stagger : {0, 3000}) {
        for (int warps : {4, 24}) {
            run<128>(warps, stagger); run<256>(warps, stagger); run<512>(warps, stagger); run<768>(warps, stagger);
            run<1024>(warps, stagger); run<1536>(warps, stagger); run<2048>(warps, stagger); run<4096>(warps, stagger);
        }
    }
    return 0;
}
