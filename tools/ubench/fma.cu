// FP32 FMA throughput of the CUDA-core pipe (the roof SURVEY 8(d) names for this path): independent FMA chains in registers.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fma tools/ubench/fma.cu && build/fma
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void __launch_bounds__(256) k_fma(float* out, int iters, float a, float b) {
    float x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = threadIdx.x * 1e-3f + c;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) x[c] = fmaf(x[c], a, b);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += x[c];
    if (s == 123.456f) out[0] = s;      // never true: keeps the chains alive
}

template <int CHAINS>
static double run(int blocks, int iters) {
    float* d; cudaMalloc(&d, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_fma<CHAINS><<<blocks, 256>>>(d, iters, 0.999f, 0.001f);      // warm-up
    cudaDeviceSynchronize();
    double best = 0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k_fma<CHAINS><<<blocks, 256>>>(d, iters, 0.999f, 0.001f);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * CHAINS * (double)iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaFree(d);
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 8;
    printf("{\"device\": \"%s\", \"sms\": %d, \"fp32_fma_tflops\": {\"chains4\": %.2f, \"chains8\": %.2f, \"chains16\": %.2f}, "
           "\"nominal_tflops_at_1965MHz\": %.2f}\n", p.name, p.multiProcessorCount, run<4>(blocks, 1 << 16), run<8>(blocks, 1 << 16),
           run<16>(blocks, 1 << 15), p.multiProcessorCount * 128 * 2 * 1.965e9 / 1e12);
    return 0;
}
