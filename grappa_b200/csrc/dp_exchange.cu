// Data-parallel gradient exchange fused with the optimiser step, over NVSwitch multicast memory.
//
// Replaces, for world_size > 1 on one NVSwitch domain, "NCCL all-reduce of the gradients, then AdamW on every rank"
// (the reference: Lightning's DDP all-reduce + torch.optim.AdamW, training/lightning_model.py:368-370) by ONE kernel
// per parameter stage:
//     every rank owns 1/world of the stage's elements;
//     it reads ITS elements' gradients already summed over all ranks by the switch (multimem.ld_reduce.add.f32 on the
//     multicast address of the flat gradient buffer), applies AdamW to them with its own first/second-moment shard,
//     and writes the new parameters to EVERY rank's flat parameter buffer with one multicast store (multimem.st).
// Per GPU and step the links carry the gradients once up and the parameters once down (what an in-switch all-reduce
// moves), AdamW touches 1/world of the optimiser state, and no rank ever holds a reduced gradient in memory.
// All ranks get bit-identical parameters: each element is computed once, by its owner.
//
// Cross-GPU ordering: a per-CTA flag exchange before (all ranks' backward passes have finished writing this stage's
// gradients) and after (all owners' parameter stores have landed; nobody still reads my gradients).  CTA b of rank r
// raises flag [b][r] in every peer's signal pad (system-scope release CAS 0 -> 1) and consumes the flags [b][*] in its
// own pad (acquire CAS 1 -> 0): the flags reset themselves, so successive launches reuse a slot range.  Every rank must
// launch the same grid.  A wait that lasts longer than spin_seconds traps (a lost peer fails the step instead of
// hanging the GPU).
#include "common.cuh"

namespace grappa {

constexpr int kMaxWorld = 16;
struct DpPeers {
    uint32_t* pad[kMaxWorld];
};

__device__ __forceinline__ void flag_raise(uint32_t* addr, long long limit) {
    uint32_t old;
    const long long t0 = clock64();
    do {
        asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
        if (old != 0 && limit > 0 && clock64() - t0 > limit) __trap();
    } while (old != 0);
}

__device__ __forceinline__ void flag_consume(uint32_t* addr, long long limit) {
    uint32_t old;
    const long long t0 = clock64();
    do {
        asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
        if (old != 1 && limit > 0 && clock64() - t0 > limit) __trap();
    } while (old != 1);
}

__device__ __forceinline__ void cross_gpu_barrier(const DpPeers& peers, int slot_base, int rank, int world, long long limit) {
    __syncthreads();
    if ((int)threadIdx.x < world) {
        const int peer = threadIdx.x;
        flag_raise(peers.pad[peer] + slot_base + (int)blockIdx.x * world + rank, limit);
        flag_consume(peers.pad[rank] + slot_base + (int)blockIdx.x * world + peer, limit);
    }
    __syncthreads();
}

__device__ __forceinline__ float4 switch_sum(const float4* mc) {          // sum over all ranks' copies, done in the switch
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
    return r;
}

__device__ __forceinline__ void switch_store(float4* mc, float4 v) {       // one store, delivered to every rank's copy
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

struct AdamConsts {
    float lr, b1, b2, eps, wd, bc1, bc2_sqrt, gscale;
};

__device__ __forceinline__ void adamw_elem(float& p, float g, float& m, float& v, const AdamConsts& c) {
    g *= c.gscale;
    float pi = p * (1.0f - c.lr * c.wd);
    m = c.b1 * m + (1.0f - c.b1) * g;
    v = c.b2 * v + (1.0f - c.b2) * g * g;
    const float denom = sqrtf(v) / c.bc2_sqrt + c.eps;
    p = pi - (c.lr / c.bc1) * (m / denom);
}

template <int UNROLL>
__global__ void __launch_bounds__(512) dp_exchange_adamw_kernel(const __grid_constant__ DpPeers peers, int slot_base, int rank, int world,
                                                                 const float4* __restrict__ grads_mc, float4* __restrict__ params_mc,
                                                                 const float4* __restrict__ params, float4* __restrict__ m4,
                                                                 float4* __restrict__ v4, int64_t first4, int64_t n4,
                                                                 AdamConsts c, long long limit) {
    cross_gpu_barrier(peers, slot_base, rank, world, limit);          // every rank's gradients of this stage are complete
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += stride * UNROLL) {
        float4 g[UNROLL], p[UNROLL], m[UNROLL], v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) g[u] = switch_sum(grads_mc + first4 + i);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) {
                p[u] = params[first4 + i];
                m[u] = m4[first4 + i];
                v[u] = v4[first4 + i];
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            if (i < n4) {
                adamw_elem(p[u].x, g[u].x, m[u].x, v[u].x, c);
                adamw_elem(p[u].y, g[u].y, m[u].y, v[u].y, c);
                adamw_elem(p[u].z, g[u].z, m[u].z, v[u].z, c);
                adamw_elem(p[u].w, g[u].w, m[u].w, v[u].w, c);
                m4[first4 + i] = m[u];
                v4[first4 + i] = v[u];
                switch_store(params_mc + first4 + i, p[u]);
            }
        }
    }
    __threadfence_system();
    cross_gpu_barrier(peers, slot_base, rank, world, limit);          // all owners' parameters have landed on this rank
}

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_dp_shard(int64_t lo, int64_t hi, int rank, int world, int64_t* first_out, int64_t* count_out) {
    GRAPPA_CHECK_ARG(lo >= 0 && hi >= lo && lo % 4 == 0 && (hi - lo) % 4 == 0, "the range must be a multiple of 4 elements");
    GRAPPA_CHECK_ARG(world >= 1 && rank >= 0 && rank < world && first_out && count_out, "bad arguments");
    const int64_t t4 = (hi - lo) / 4;
    const int64_t a = t4 * rank / world, b = t4 * (rank + 1) / world;
    *first_out = lo + 4 * a;
    *count_out = 4 * (b - a);
    return 0;
}

extern "C" int grappa_dp_exchange_adamw(int64_t lo, int64_t hi, int rank, int world, const float* grads_mc, float* params_mc,
                                        float* params, float* exp_avg, float* exp_avg_sq, const uint64_t* peer_signal_pads,
                                        int64_t pad_words, int64_t slot_base, int n_ctas, float lr, float beta1, float beta2,
                                        float eps, float weight_decay, int64_t step, float grad_scale, float spin_seconds,
                                        void* stream) {
    GRAPPA_CHECK_ARG(world >= 2 && world <= kMaxWorld && rank >= 0 && rank < world, "world size must be 2..16");
    GRAPPA_CHECK_ARG(grads_mc && params_mc && params && exp_avg && exp_avg_sq && peer_signal_pads && step >= 1, "bad arguments");
    GRAPPA_CHECK_ARG(n_ctas >= 1 && n_ctas <= 1024 && slot_base >= 0 && slot_base + (int64_t)n_ctas * world <= pad_words,
                     "the signal pad is too small for this grid");
    int64_t first = 0, count = 0;
    if (int rc = grappa_dp_shard(lo, hi, rank, world, &first, &count)) return rc;
    GRAPPA_CHECK_ARG((reinterpret_cast<uintptr_t>(grads_mc) | reinterpret_cast<uintptr_t>(params_mc) | reinterpret_cast<uintptr_t>(params) |
                      reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) % 16 == 0, "buffers must be 16-byte aligned");
    DpPeers peers{};
    for (int r = 0; r < world; ++r) {
        GRAPPA_CHECK_ARG(peer_signal_pads[r] != 0, "missing signal pad");
        peers.pad[r] = reinterpret_cast<uint32_t*>(peer_signal_pads[r]);
    }
    AdamConsts c;
    c.lr = lr; c.b1 = beta1; c.b2 = beta2; c.eps = eps; c.wd = weight_decay;
    c.bc1 = 1.0f - powf(beta1, (float)step);
    c.bc2_sqrt = sqrtf(1.0f - powf(beta2, (float)step));
    c.gscale = grad_scale;
    static std::atomic<int> khz_of[64];          // the clock-rate query is slow: once per device
    int dev = 0;
    GRAPPA_CUDA(cudaGetDevice(&dev));
    int khz = khz_of[dev & 63].load(std::memory_order_relaxed);
    if (khz == 0) {
        GRAPPA_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
        khz_of[dev & 63].store(khz, std::memory_order_relaxed);
    }
    const long long limit = spin_seconds > 0 ? (long long)(spin_seconds * 1e3 * (double)khz) : 0;
    dp_exchange_adamw_kernel<4><<<n_ctas, 512, 0, (cudaStream_t)stream>>>(
        peers, (int)slot_base, rank, world, reinterpret_cast<const float4*>(grads_mc), reinterpret_cast<float4*>(params_mc),
        reinterpret_cast<const float4*>(params), reinterpret_cast<float4*>(exp_avg), reinterpret_cast<float4*>(exp_avg_sq),
        first / 4, count / 4, c, limit);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}
