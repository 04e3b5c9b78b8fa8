// Fused molecule-wise loss: replaces the Python loop over dgl.unbatch output in
// training/loss.py:51-174 (+ graph_utils.get_energies :46-74, get_gradients :78-97, dgl_utils.delete_dummy_confs
// :90-125) for the configuration the training step uses (param_weight 0).  One CTA per molecule computes the loss
// term AND its gradients w.r.t. energy, gradient and torsion force constants (the loss is a sum of means of
// squares, so the backward is closed-form); fixed-order block reductions, no atomics.
#include "common.cuh"

namespace grappa {

constexpr int kLossThreads = 256;

__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kLossThreads / 32; ++i) s += sh[i];
    return s;
}

__global__ void __launch_bounds__(kLossThreads) molwise_loss_kernel(
    int64_t n_mol, int64_t C, const int32_t* __restrict__ a_off, const int32_t* __restrict__ p_off, const int32_t* __restrict__ i_off,
    const float* __restrict__ E, const float* __restrict__ Er, const float* __restrict__ G, const float* __restrict__ Gr,
    const float* __restrict__ dummy, const float* __restrict__ k4, int P4, const float* __restrict__ k4i, int P4i,
    float ew, float gw, float preg, float ireg, const float* __restrict__ upstream, float* __restrict__ parts,
    float* __restrict__ dE, float* __restrict__ dG, float* __restrict__ dk4, float* __restrict__ dk4i) {
    __shared__ float sh[kLossThreads / 32];
    const int64_t m = blockIdx.x;
    const int tid = threadIdx.x;
    const float inv_norm = 1.0f / (float)n_mol;
    const float inv_mol = upstream ? inv_norm * (*upstream) : inv_norm;   // scales the gradient outputs only
    // valid conformations and conformation means
    float nv = 0.f, se = 0.f, ser = 0.f;
    for (int64_t c = tid; c < C; c += kLossThreads) {
        const bool ok = !dummy || dummy[m * C + c] == 0.f;
        if (ok) { nv += 1.f; se += E[m * C + c]; ser += Er[m * C + c]; }
    }
    nv = block_sum(nv, sh); se = block_sum(se, sh); ser = block_sum(ser, sh);
    const float me = nv > 0.f ? se / nv : 0.f, mer = nv > 0.f ? ser / nv : 0.f;
    float le = 0.f;
    for (int64_t c = tid; c < C; c += kLossThreads) {
        const bool ok = !dummy || dummy[m * C + c] == 0.f;
        float g = 0.f;
        if (ok) {
            const float d = (E[m * C + c] - me) - (Er[m * C + c] - mer);
            le = fmaf(d, d, le);
            g = ew * 2.0f * d / nv * inv_mol;
        }
        if (dE) dE[m * C + c] = g;
    }
    le = block_sum(le, sh);
    float term = nv > 0.f ? ew * le / nv : 0.f;
    // gradients
    const int64_t a0 = a_off[m], a1 = a_off[m + 1];
    const int64_t ng = (a1 - a0) * C * 3;
    const float gnorm = (float)(a1 - a0) * nv * 3.0f;
    float lg = 0.f;
    for (int64_t i = tid; i < ng; i += kLossThreads) {
        const int64_t c = (i / 3) % C;
        const bool ok = !dummy || dummy[m * C + c] == 0.f;
        const int64_t j = a0 * C * 3 + i;
        float g = 0.f;
        if (ok) {
            const float d = G[j] - Gr[j];
            lg = fmaf(d, d, lg);
            g = gw * 2.0f * d / gnorm * inv_mol;
        }
        if (dG) dG[j] = g;
    }
    lg = block_sum(lg, sh);
    if (gnorm > 0.f) term += gw * lg / gnorm;
    // L2 on torsion force constants (the improper term counts twice when impropers exist: loss.py:135-139)
    if (preg > 0.f) {
        const int64_t n = (int64_t)(p_off[m + 1] - p_off[m]) * P4, o = (int64_t)p_off[m] * P4;
        float s = 0.f;
        for (int64_t i = tid; i < n; i += kLossThreads) {
            const float v = k4[o + i];
            s = fmaf(v, v, s);
            if (dk4) dk4[o + i] = preg * 2.0f * v / (float)n * inv_mol;
        }
        s = block_sum(s, sh);
        if (n > 0) term += preg * s / (float)n;
    } else if (dk4) {
        const int64_t n = (int64_t)(p_off[m + 1] - p_off[m]) * P4, o = (int64_t)p_off[m] * P4;
        for (int64_t i = tid; i < n; i += kLossThreads) dk4[o + i] = 0.f;
    }
    {
        const int64_t n = (int64_t)(i_off[m + 1] - i_off[m]) * P4i, o = (int64_t)i_off[m] * P4i;
        if (ireg > 0.f) {
            float s = 0.f;
            for (int64_t i = tid; i < n; i += kLossThreads) {
                const float v = k4i[o + i];
                s = fmaf(v, v, s);
                if (dk4i) dk4i[o + i] = 2.0f * ireg * 2.0f * v / (float)n * inv_mol;
            }
            s = block_sum(s, sh);
            if (n > 0) term += 2.0f * ireg * s / (float)n;
        } else if (dk4i) {
            for (int64_t i = tid; i < n; i += kLossThreads) dk4i[o + i] = 0.f;
        }
    }
    if (tid == 0 && parts) parts[m] = term * inv_norm;
}

__global__ void loss_total_kernel(int64_t n, const float* __restrict__ parts, float* __restrict__ out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float s = 0.f;
        for (int64_t i = 0; i < n; ++i) s += parts[i];
        *out = s;
    }
}

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_molwise_loss(int64_t n_mol, int64_t C, const int32_t* mol_atom_off, const int32_t* mol_p_off,
                                   const int32_t* mol_i_off, const float* energy, const float* energy_ref,
                                   const float* gradient, const float* gradient_ref, const float* is_dummy,
                                   const float* k4, int P4, const float* k4i, int P4i,
                                   float energy_weight, float gradient_weight, float proper_reg, float improper_reg,
                                   const float* upstream, float* loss_parts, float* loss_out, float* dE, float* dG,
                                   float* dk4, float* dk4i, void* stream) {
    GRAPPA_CHECK_ARG(n_mol > 0 && C > 0, "need at least one molecule and one conformation");
    GRAPPA_CHECK_ARG(mol_atom_off && mol_p_off && mol_i_off && energy && energy_ref && gradient && gradient_ref, "null input");
    GRAPPA_CHECK_ARG((loss_parts != nullptr) == (loss_out != nullptr), "loss_parts and loss_out go together");
    GRAPPA_CHECK_ARG((proper_reg <= 0.f || k4) && (improper_reg <= 0.f || k4i), "torsion constants needed for regularisation");
    cudaStream_t st = (cudaStream_t)stream;
    molwise_loss_kernel<<<(unsigned)n_mol, kLossThreads, 0, st>>>(n_mol, C, mol_atom_off, mol_p_off, mol_i_off, energy, energy_ref,
                                                                  gradient, gradient_ref, is_dummy, k4, P4, k4i, P4i,
                                                                  energy_weight, gradient_weight, proper_reg, improper_reg,
                                                                  upstream, loss_parts, dE, dG, dk4, dk4i);
    GRAPPA_CHECK_LAUNCH();
    if (loss_out) {
        loss_total_kernel<<<1, 32, 0, st>>>(n_mol, loss_parts, loss_out);
        GRAPPA_CHECK_LAUNCH();
    }
    return 0;
}
