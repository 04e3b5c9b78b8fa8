// Shared helpers for libgrappa_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <algorithm>
#include <stdlib.h>
#include <utility>
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

// Programmatic dependent launch: a kernel launched through launch_pdl may start while its predecessor in the stream is
// still draining (as soon as every CTA of the predecessor has called pdl_trigger() or exited); it must call pdl_wait()
// before its first access to memory the predecessor touches.  What overlaps is the launch latency and the prologue
// (barrier init, TMEM allocation, descriptor prefetch).  OFF by default (GRAPPA_PDL=1 turns the attribute on): measured on
// the cfg-2 step, triggering at kernel start costs 8 % (early CTAs park 197 KB of shared memory on SMs that the other four
// streams' kernels need), triggering before the write-out is neutral (3.524 vs 3.514 ms).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("GRAPPA_PDL"); on = (e && atoi(e) != 0) ? 1 : 0; }
    return on != 0;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

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
