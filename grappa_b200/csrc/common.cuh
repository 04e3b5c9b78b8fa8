// Shared helpers for libgrappa_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <algorithm>
#include <stdlib.h>
#include "../../include/grappa_b200.h"

namespace grappa {

extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;

inline int fail(const char* fmt, const char* a = "", long long b = 0, long long c = 0) {
    snprintf(g_err, sizeof(g_err), fmt, a, b, c);
    return 1;
}

#define GRAPPA_CHECK_ARG(cond, msg)                                                         \
    do {                                                                                    \
        if (!(cond)) {                                                                      \
            snprintf(grappa::g_err, sizeof(grappa::g_err), "%s: %s", __func__, msg);       \
            return 2;                                                                       \
        }                                                                                   \
    } while (0)

// Checks the launch (not the execution: calls are asynchronous).
#define GRAPPA_CHECK_LAUNCH()                                                               \
    do {                                                                                    \
        grappa::g_launches.fetch_add(1, std::memory_order_relaxed);                        \
        cudaError_t e__ = cudaGetLastError();                                               \
        if (e__ != cudaSuccess) {                                                           \
            snprintf(grappa::g_err, sizeof(grappa::g_err), "%s: CUDA launch failed: %s",   \
                     __func__, cudaGetErrorString(e__));                                    \
            return 3;                                                                       \
        }                                                                                   \
    } while (0)

#define GRAPPA_CUDA(call)                                                                   \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            snprintf(grappa::g_err, sizeof(grappa::g_err), "%s: %s failed: %s", __func__,  \
                     #call, cudaGetErrorString(e__));                                       \
            return 3;                                                                       \
        }                                                                                   \
    } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Counter-based keep mask for dropout: splitmix64 finaliser of (seed, element index); uniform in [0,1).
__host__ __device__ __forceinline__ float hash_uniform(uint64_t seed, uint64_t idx) {
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (idx + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float elu1(float x) { return x > 0.f ? x : expm1f(x); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace grappa
