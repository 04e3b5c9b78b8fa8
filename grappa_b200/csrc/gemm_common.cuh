// Parameters shared by the FFMA (gemm.cu) and tcgen05 (gemm_tc.cu) GEMM paths.
#pragma once
#include "common.cuh"

namespace grappa {

struct GemmParams {
    const float* A; const float* B; float* C;
    const float* bias; const float* res1; const float* res2;
    int64_t M, N, K, lda, ldb, ldc, ldr1, ldr2;
    int act; float drop_p; float drop_scale; uint64_t seed; const uint64_t* seed_dev;
    int a_vec, b_vec, c_vec;   // 16-byte vector access allowed for A / B / (C, res)
    // optional last epilogue step (backward GEMMs): C *= ELU'(pre-activation) read off the saved ELU OUTPUT s:
    // ELU'(x) = 1 for x > 0, exp(x) = s + 1 otherwise
    const float* elu_out = nullptr; int64_t ld_eo = 0;
    const float* B_lo = nullptr;   // optional pre-split lo plane of B (same shape / ldb): see gemm_tc.cu B_PRE
    // bf16x3 path (gemm_tc.cu, gemm_bf16_kernel): B given as pre-split bf16 planes (N, K) with leading dimension ldb16
    const uint16_t* B_hi16 = nullptr; const uint16_t* B_lo16 = nullptr; int64_t ldb16 = 0;
    // optional row gather of res1 (tensor-core paths): row m of the residual is res1[res1_idx[m], :] -- the residual is a
    // table of distinct token contents and is never materialised per token (readout: out_proj + gathered LN(x) rows)
    const int32_t* res1_idx = nullptr;
    int cta_budget = 0;        // > 0: background GEMM (weight gradients on a side stream): split-K grids are sized for this
                               // many CTAs instead of the whole machine, so that the critical chain always finds free SMs
};

__device__ __forceinline__ float elu_grad_from_out(float s) { return s > 0.f ? 1.f : s + 1.f; }

int gemm_tc_dispatch(int layout, const GemmParams& p, cudaStream_t st, float* workspace, int64_t workspace_floats,
                     bool* handled);   // gemm_tc.cu

int gemm_bf16_launch(const GemmParams& p, cudaStream_t st, float* workspace, int64_t workspace_floats);   // gemm_tc.cu

// streaming kernels for GEMMs with one dimension <= 8 (gemm_skinny.cu)
int gemm_skinny_dispatch(int layout, const GemmParams& p, cudaStream_t st, float* workspace, int64_t workspace_floats,
                         bool* handled);

// C = epilogue(sum_z partial[z]) in fixed order; shared by both GEMM paths (gemm_tc.cu)
int splitk_reduce(const GemmParams& p, int splits, const float* partial, cudaStream_t st);

}  // namespace grappa
