// K2 (skinny shapes): GEMMs in which one matrix dimension is <= 8 -- the last layer of every symmetriser
// (256 -> 2 or 6 outputs, perm_equiv_transformer.py:283-290 via network_utils.py:44-54) forward, its dX and its dW.
// On the 64x64 FFMA tiles these waste 7/8 of the arithmetic and occupy 20-95 CTAs for 15-20 us each; they are
// byte-bound, so each gets a streaming kernel: one pass over the big operand, the small one in shared memory.
//   layout 0, N <= 8:  C[M,N]   = A[M,K] . B[N,K]^T (+ bias)                         warp per row
//   layout 1, K <= 8:  C[M,N]   = A[M,K] . B[K,N]   (* ELU'(elu_out), optional)      thread per 4 output columns
//   layout 2, M <= 8:  C[M,N]   = A[K,M]^T . B[K,N]                                  two-stage fixed-order reduction over K
// Deterministic (no atomics).  Anything else about the epilogue (activation, dropout, residuals) falls back to the tiles.
#include "gemm_common.cuh"

namespace grappa {

constexpr int kSkinny = 8;

// ---- layout 0 ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) skinny_nt_kernel(GemmParams p) {
    extern __shared__ float w_s[];                       // [N][K]
    const int N = (int)p.N, K = (int)p.K;
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) w_s[i] = __ldg(p.B + (int64_t)(i / K) * p.ldb + (i % K));
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int K4 = K >> 2;
    for (int64_t m = warp; m < p.M; m += n_warps) {
        float acc[kSkinny];
#pragma unroll
        for (int n = 0; n < kSkinny; ++n) acc[n] = 0.f;
        const float4* a4 = reinterpret_cast<const float4*>(p.A + m * p.lda);
        for (int k4 = lane; k4 < K4; k4 += 32) {
            const float4 a = __ldg(a4 + k4);
#pragma unroll
            for (int n = 0; n < kSkinny; ++n) {
                if (n < N) {
                    const float4 w = *reinterpret_cast<const float4*>(w_s + n * K + k4 * 4);
                    acc[n] = fmaf(a.x, w.x, fmaf(a.y, w.y, fmaf(a.z, w.z, fmaf(a.w, w.w, acc[n]))));
                }
            }
        }
#pragma unroll
        for (int n = 0; n < kSkinny; ++n)
            if (n < N) acc[n] = warp_sum(acc[n]);
        if (lane == 0) {
#pragma unroll
            for (int n = 0; n < kSkinny; ++n)
                if (n < N) p.C[m * p.ldc + n] = acc[n] + (p.bias ? __ldg(p.bias + n) : 0.f);
        }
    }
}

// ---- layout 1 ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) skinny_nn_kernel(GemmParams p) {
    extern __shared__ float w_s[];                       // [K][N]
    const int N = (int)p.N, K = (int)p.K;
    for (int i = threadIdx.x; i < K * N; i += blockDim.x) w_s[i] = __ldg(p.B + (int64_t)(i / N) * p.ldb + (i % N));
    __syncthreads();
    const int N4 = N >> 2;
    const int64_t total = p.M * N4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t m = i / N4;
        const int c = (int)(i % N4) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < kSkinny; ++k) {
            if (k < K) {
                const float a = __ldg(p.A + m * p.lda + k);
                const float4 w = *reinterpret_cast<const float4*>(w_s + k * N + c);
                v.x = fmaf(a, w.x, v.x); v.y = fmaf(a, w.y, v.y); v.z = fmaf(a, w.z, v.z); v.w = fmaf(a, w.w, v.w);
            }
        }
        if (p.elu_out) {
            const float4 e = __ldg(reinterpret_cast<const float4*>(p.elu_out + m * p.ld_eo + c));
            v.x *= elu_grad_from_out(e.x); v.y *= elu_grad_from_out(e.y); v.z *= elu_grad_from_out(e.z); v.w *= elu_grad_from_out(e.w);
        }
        *reinterpret_cast<float4*>(p.C + m * p.ldc + c) = v;
    }
}

// ---- layout 2 ---------------------------------------------------------------------------------------------------
// stage 1: CTA = (chunk of rows of the reduction dimension, 256 output columns); thread per column, M accumulators
__global__ void __launch_bounds__(256) skinny_tn_stage1_kernel(GemmParams p, int rows_per_chunk, float* __restrict__ part) {
    const int M = (int)p.M;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t k0 = (int64_t)blockIdx.y * rows_per_chunk, k1 = min(p.K, k0 + rows_per_chunk);
    __shared__ float a_s[64 * kSkinny];
    float acc[kSkinny];
#pragma unroll
    for (int m = 0; m < kSkinny; ++m) acc[m] = 0.f;
    for (int64_t kb = k0; kb < k1; kb += 64) {
        const int nr = (int)min((int64_t)64, k1 - kb);
        __syncthreads();
        for (int i = threadIdx.x; i < nr * M; i += blockDim.x) a_s[(i / M) * kSkinny + (i % M)] = __ldg(p.A + (kb + i / M) * p.lda + (i % M));
        __syncthreads();
        if (n < p.N) {
            for (int r = 0; r < nr; ++r) {
                const float b = __ldg(p.B + (kb + r) * p.ldb + n);
#pragma unroll
                for (int m = 0; m < kSkinny; ++m)
                    if (m < M) acc[m] = fmaf(a_s[r * kSkinny + m], b, acc[m]);
            }
        }
    }
    if (n < p.N) {
#pragma unroll
        for (int m = 0; m < kSkinny; ++m)
            if (m < M) part[((int64_t)blockIdx.y * M + m) * p.N + n] = acc[m];
    }
}

__global__ void __launch_bounds__(256) skinny_tn_stage2_kernel(GemmParams p, int chunks, const float* __restrict__ part) {
    const int64_t total = p.M * p.N;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int c = 0; c < chunks; ++c) s += __ldcg(part + (int64_t)c * total + i);
        p.C[(i / p.N) * p.ldc + (i % p.N)] = s;
    }
}

// Returns handled = true when one of the streaming kernels took the call.
int gemm_skinny_dispatch(int layout, const GemmParams& p, cudaStream_t st, float* workspace, int64_t workspace_floats, bool* handled) {
    *handled = false;
    const bool plain = p.act == 0 && p.drop_p == 0.f && !p.res1 && !p.res2;
    if (layout == 0 && p.N <= kSkinny && plain && !p.elu_out && p.a_vec && p.K % 4 == 0 && p.N * p.K * 4 <= 32 * 1024 && p.M >= 64) {
        const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(p.M, 8), 148 * 8);
        skinny_nt_kernel<<<grid, 256, (size_t)(p.N * p.K * 4), st>>>(p);
        GRAPPA_CHECK_LAUNCH();
        *handled = true;
    } else if (layout == 1 && p.K <= kSkinny && plain && !p.bias && p.c_vec && p.N % 4 == 0 && p.b_vec && p.K * p.N * 4 <= 32 * 1024 &&
               (!p.elu_out || (p.ld_eo % 4 == 0 && (reinterpret_cast<uintptr_t>(p.elu_out) & 15) == 0)) && p.M >= 64) {
        const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(p.M * (p.N / 4), 256), 148 * 8);
        skinny_nn_kernel<<<grid, 256, (size_t)(p.K * p.N * 4), st>>>(p);
        GRAPPA_CHECK_LAUNCH();
        *handled = true;
    } else if (layout == 2 && p.M <= kSkinny && plain && !p.bias && !p.elu_out && p.K >= 256 && workspace != nullptr) {
        const int rows_per_chunk = (int)std::max<int64_t>(128, ceil_div(ceil_div(p.K, 96), 64) * 64);
        const int chunks = (int)ceil_div(p.K, rows_per_chunk);
        if ((int64_t)chunks * p.M * p.N > workspace_floats) return 0;
        dim3 grid((unsigned)ceil_div(p.N, 256), (unsigned)chunks);
        skinny_tn_stage1_kernel<<<grid, 256, 0, st>>>(p, rows_per_chunk, workspace);
        GRAPPA_CHECK_LAUNCH();
        skinny_tn_stage2_kernel<<<(unsigned)std::min<int64_t>(ceil_div(p.M * p.N, 256), 148), 256, 0, st>>>(p, chunks, workspace);
        GRAPPA_CHECK_LAUNCH();
        *handled = true;
    }
    return 0;
}

}  // namespace grappa
