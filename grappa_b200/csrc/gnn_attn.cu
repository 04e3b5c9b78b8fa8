// K1: dot-product graph attention over the bond graph, CSR-sorted, one warp per node, no atomics.
// Forward: for destination v, one pass over its incoming edges with an online softmax (running max / sum), so the
// per-edge scores and messages of DGL's SDDMM -> edge_softmax -> SpMM chain are never materialised.
// Backward: one warp per node u accumulates dft[u] from BOTH roles of u: as destination (its incoming edges) and as
// source (its outgoing edges, via the CSR-by-source), recomputing the softmax weights from the saved log-sum-exp.
// Algorithmic HBM bytes per layer, forward: (2N + E) * H*D*4  (read ft[v], gather ft[u] per edge, write out[v]).
#include "common.cuh"

namespace grappa {

__device__ __forceinline__ float dot4(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }

// sum over the D/4 consecutive lanes that hold one head
template <int D>
__device__ __forceinline__ float head_sum(float v) {
#pragma unroll
    for (int o = D / 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int D, int NI>
__global__ void __launch_bounds__(256) gnn_attn_fwd_kernel(int64_t N, int H, const float* __restrict__ ft,
                                                           const int32_t* __restrict__ in_ptr, const int32_t* __restrict__ in_src,
                                                           float* __restrict__ out, float* __restrict__ lse) {
    const int64_t v = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (v >= N) return;
    constexpr int HD4 = NI * 32;          // float4s per row
    const float scale = rsqrtf((float)D);
    const float4* fv_p = reinterpret_cast<const float4*>(ft) + v * HD4;
    float4 fv[NI], acc[NI];
    float m[NI], l[NI];
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        fv[i] = __ldg(fv_p + lane + 32 * i);
        acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        m[i] = -INFINITY; l[i] = 0.f;
    }
    const int e0 = __ldg(in_ptr + v), e1 = __ldg(in_ptr + v + 1);
    for (int e = e0; e < e1; ++e) {
        const int64_t u = __ldg(in_src + e);
        const float4* fu_p = reinterpret_cast<const float4*>(ft) + u * HD4;
        float4 fu[NI];
#pragma unroll
        for (int i = 0; i < NI; ++i) fu[i] = __ldg(fu_p + lane + 32 * i);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const float s = head_sum<D>(dot4(fu[i], fv[i])) * scale;
            const float mn = fmaxf(m[i], s);
            const float corr = expf(m[i] - mn);     // exp(-inf) = 0 on the first edge
            const float p = expf(s - mn);
            l[i] = l[i] * corr + p;
            acc[i].x = acc[i].x * corr + p * fu[i].x;
            acc[i].y = acc[i].y * corr + p * fu[i].y;
            acc[i].z = acc[i].z * corr + p * fu[i].z;
            acc[i].w = acc[i].w * corr + p * fu[i].w;
            m[i] = mn;
        }
    }
    float4* o = reinterpret_cast<float4*>(out) + v * HD4;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const float inv = l[i] > 0.f ? 1.0f / l[i] : 0.f;
        o[lane + 32 * i] = make_float4(acc[i].x * inv, acc[i].y * inv, acc[i].z * inv, acc[i].w * inv);
        if ((lane % (D / 4)) == 0) {
            const int head = i * (128 / D) + lane / (D / 4);
            lse[v * H + head] = l[i] > 0.f ? m[i] + logf(l[i]) : 0.f;
        }
    }
}

// delta[v,h] = <dOut[v,h,:], out[v,h,:]>
template <int D, int NI>
__global__ void __launch_bounds__(256) gnn_attn_delta_kernel(int64_t N, int H, const float* __restrict__ out,
                                                             const float* __restrict__ dOut, float* __restrict__ delta) {
    const int64_t v = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (v >= N) return;
    constexpr int HD4 = NI * 32;
    const float4* a = reinterpret_cast<const float4*>(out) + v * HD4;
    const float4* b = reinterpret_cast<const float4*>(dOut) + v * HD4;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const float s = head_sum<D>(dot4(__ldg(a + lane + 32 * i), __ldg(b + lane + 32 * i)));
        if ((lane % (D / 4)) == 0) delta[v * H + i * (128 / D) + lane / (D / 4)] = s;
    }
}

template <int D, int NI>
__global__ void __launch_bounds__(256) gnn_attn_bwd_kernel(int64_t N, int H, const float* __restrict__ ft,
                                                           const float* __restrict__ lse, const float* __restrict__ delta,
                                                           const float* __restrict__ dOut,
                                                           const int32_t* __restrict__ in_ptr, const int32_t* __restrict__ in_src,
                                                           const int32_t* __restrict__ out_ptr, const int32_t* __restrict__ out_dst,
                                                           float* __restrict__ dft) {
    const int64_t u = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (u >= N) return;
    constexpr int HD4 = NI * 32;
    const float scale = rsqrtf((float)D);
    const float4* F = reinterpret_cast<const float4*>(ft);
    const float4* G = reinterpret_cast<const float4*>(dOut);
    float4 fu[NI], gu[NI], acc[NI];
    float lse_u[NI], del_u[NI];
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        fu[i] = __ldg(F + u * HD4 + lane + 32 * i);
        gu[i] = __ldg(G + u * HD4 + lane + 32 * i);
        acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int head = i * (128 / D) + lane / (D / 4);
        lse_u[i] = __ldg(lse + u * H + head);
        del_u[i] = __ldg(delta + u * H + head);
    }
    // u as destination: edges w -> u.  d ft[u] += scale * ds_wu * ft[w]
    for (int e = __ldg(in_ptr + u), e1 = __ldg(in_ptr + u + 1); e < e1; ++e) {
        const int64_t w = __ldg(in_src + e);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const float4 fw = __ldg(F + w * HD4 + lane + 32 * i);
            const float s = head_sum<D>(dot4(fw, fu[i])) * scale;
            const float da = head_sum<D>(dot4(gu[i], fw));
            const float a = expf(s - lse_u[i]);
            const float ds = a * (da - del_u[i]) * scale;
            acc[i].x = fmaf(ds, fw.x, acc[i].x); acc[i].y = fmaf(ds, fw.y, acc[i].y);
            acc[i].z = fmaf(ds, fw.z, acc[i].z); acc[i].w = fmaf(ds, fw.w, acc[i].w);
        }
    }
    // u as source: edges u -> v.  d ft[u] += a_uv * dOut[v] + scale * ds_uv * ft[v]
    for (int e = __ldg(out_ptr + u), e1 = __ldg(out_ptr + u + 1); e < e1; ++e) {
        const int64_t v = __ldg(out_dst + e);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const float4 fv = __ldg(F + v * HD4 + lane + 32 * i);
            const float4 gv = __ldg(G + v * HD4 + lane + 32 * i);
            const int head = i * (128 / D) + lane / (D / 4);
            const float s = head_sum<D>(dot4(fu[i], fv)) * scale;
            const float da = head_sum<D>(dot4(gv, fu[i]));
            const float a = expf(s - __ldg(lse + v * H + head));
            const float ds = a * (da - __ldg(delta + v * H + head)) * scale;
            acc[i].x += a * gv.x + ds * fv.x; acc[i].y += a * gv.y + ds * fv.y;
            acc[i].z += a * gv.z + ds * fv.z; acc[i].w += a * gv.w + ds * fv.w;
        }
    }
    float4* o = reinterpret_cast<float4*>(dft) + u * HD4;
#pragma unroll
    for (int i = 0; i < NI; ++i) o[lane + 32 * i] = acc[i];
}

template <int D, int NI>
static int attn_fwd_launch(int64_t N, int H, const float* ft, const int32_t* in_ptr, const int32_t* in_src, float* out,
                           float* lse, cudaStream_t st) {
    gnn_attn_fwd_kernel<D, NI><<<(unsigned)ceil_div(N * 32, 256), 256, 0, st>>>(N, H, ft, in_ptr, in_src, out, lse);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

template <int D, int NI>
static int attn_bwd_launch(int64_t N, int H, const float* ft, const float* out, const float* lse, const float* dOut,
                           const int32_t* in_ptr, const int32_t* in_src, const int32_t* out_ptr, const int32_t* out_dst,
                           float* delta, float* dft, cudaStream_t st) {
    const unsigned grid = (unsigned)ceil_div(N * 32, 256);
    gnn_attn_delta_kernel<D, NI><<<grid, 256, 0, st>>>(N, H, out, dOut, delta);
    GRAPPA_CHECK_LAUNCH();
    gnn_attn_bwd_kernel<D, NI><<<grid, 256, 0, st>>>(N, H, ft, lse, delta, dOut, in_ptr, in_src, out_ptr, out_dst, dft);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

}  // namespace grappa

using namespace grappa;

#define GNN_DISPATCH(FN, ...)                                                                      \
    do {                                                                                           \
        const int ni = (H * D) / 128;                                                              \
        if (D == 32 && ni == 4) return FN<32, 4>(__VA_ARGS__);                                     \
        if (D == 32 && ni == 2) return FN<32, 2>(__VA_ARGS__);                                     \
        if (D == 32 && ni == 1) return FN<32, 1>(__VA_ARGS__);                                     \
        if (D == 32 && ni == 8) return FN<32, 8>(__VA_ARGS__);                                     \
        if (D == 64 && ni == 4) return FN<64, 4>(__VA_ARGS__);                                     \
        if (D == 64 && ni == 2) return FN<64, 2>(__VA_ARGS__);                                     \
        if (D == 64 && ni == 1) return FN<64, 1>(__VA_ARGS__);                                     \
        if (D == 64 && ni == 8) return FN<64, 8>(__VA_ARGS__);                                     \
        if (D == 16 && ni == 4) return FN<16, 4>(__VA_ARGS__);                                     \
        if (D == 16 && ni == 2) return FN<16, 2>(__VA_ARGS__);                                     \
        if (D == 16 && ni == 1) return FN<16, 1>(__VA_ARGS__);                                     \
        GRAPPA_CHECK_ARG(false, "unsupported (heads, head_dim): need head_dim in {16,32,64} and heads*head_dim in {128,256,512,1024}"); \
    } while (0)

extern "C" int grappa_gnn_attn_fwd(int64_t N, int H, int D, const float* ft, const int32_t* in_ptr, const int32_t* in_src,
                                   float* out, float* lse, void* stream) {
    GRAPPA_CHECK_ARG(N >= 0 && H > 0 && D > 0, "bad sizes");
    if (N == 0) return 0;
    GRAPPA_CHECK_ARG(ft && in_ptr && in_src && out && lse, "null pointer");
    GRAPPA_CHECK_ARG((H * D) % 128 == 0, "heads*head_dim must be a multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    GNN_DISPATCH(attn_fwd_launch, N, H, ft, in_ptr, in_src, out, lse, st);
    return 0;
}

extern "C" int grappa_gnn_attn_bwd(int64_t N, int H, int D, const float* ft, const float* out, const float* lse,
                                   const float* dOut, const int32_t* in_ptr, const int32_t* in_src,
                                   const int32_t* out_ptr, const int32_t* out_dst, float* delta, float* dft, void* stream) {
    GRAPPA_CHECK_ARG(N >= 0 && H > 0 && D > 0, "bad sizes");
    if (N == 0) return 0;
    GRAPPA_CHECK_ARG(ft && out && lse && dOut && in_ptr && in_src && out_ptr && out_dst && delta && dft, "null pointer");
    GRAPPA_CHECK_ARG((H * D) % 128 == 0, "heads*head_dim must be a multiple of 128");
    cudaStream_t st = (cudaStream_t)stream;
    GNN_DISPATCH(attn_bwd_launch, N, H, ft, out, lse, dOut, in_ptr, in_src, out_ptr, out_dst, delta, dft, st);
    return 0;
}
