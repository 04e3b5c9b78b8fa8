// microbenchmark: FFMA vs FFMA2 (fma.rn.f32x2) issue throughput on sm_100a, plus MUFU rsqrt and LDS/STS mix
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k_ffma(float* out, int iters) {
    float a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    float b = out[0], c = out[1];
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            a0 = fmaf(a0, b, c); a1 = fmaf(a1, b, c); a2 = fmaf(a2, b, c); a3 = fmaf(a3, b, c);
            a4 = fmaf(a4, b, c); a5 = fmaf(a5, b, c); a6 = fmaf(a6, b, c); a7 = fmaf(a7, b, c);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__global__ void k_ffma2(float* out, int iters) {
    unsigned long long a[8];
    for (int j = 0; j < 8; ++j) a[j] = ((unsigned long long)__float_as_uint((float)threadIdx.x + j) << 32) | __float_as_uint((float)j);
    unsigned long long b = ((unsigned long long)__float_as_uint(out[0]) << 32) | __float_as_uint(out[0]);
    unsigned long long c = ((unsigned long long)__float_as_uint(out[1]) << 32) | __float_as_uint(out[1]);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int q = 0; q < 8; ++q) a[q] = fma2(a[q], b, c);
        }
    }
    unsigned long long s = 0; for (int j = 0; j < 8; ++j) s ^= a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = __uint_as_float((unsigned)s) + __uint_as_float((unsigned)(s >> 32));
}
__global__ void k_mufu(float* out, int iters) {
    float a0 = threadIdx.x + 1.f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a0)); asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a1));
            asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a2)); asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a3));
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}
template <typename F> float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 1024 * 4 * 4); cudaMemset(d, 0, 148 * 8 * 1024 * 4 * 4);
    const int iters = 4096, blocks = 148 * 4, threads = 512;
    for (int w = 0; w < 2; ++w) {
        float t1 = timeit([&] { k_ffma<<<blocks, threads>>>(d, iters); });
        float t2 = timeit([&] { k_ffma2<<<blocks, threads>>>(d, iters); });
        float t3 = timeit([&] { k_mufu<<<blocks, threads>>>(d, iters); });
        double n = (double)blocks * threads / 32 * iters * 64;   // warp instructions
        printf("FFMA  : %.3f ms  %.1f G warp-instr/s  (%.2f per clk per SM @1.965GHz)\n", t1, n / t1 * 1e-6, n / t1 * 1e-6 / 148 / 1.965);
        printf("FFMA2 : %.3f ms  %.1f G warp-instr/s  (%.2f per clk per SM)\n", t2, n / t2 * 1e-6, n / t2 * 1e-6 / 148 / 1.965);
        printf("MUFU  : %.3f ms  %.1f G warp-instr/s  (%.2f per clk per SM)\n", t3, n / t3 * 1e-6, n / t3 * 1e-6 / 148 / 1.965);
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
