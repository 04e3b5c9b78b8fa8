// K3 building blocks of the interaction readout (everything between the dense projections):
// tuple gather (+ positional-encoding column) and its pull-style backward, self-attention over the <= 4 tokens of a
// tuple in registers, permuted-view construction for the symmetriser, and the permutation sum fused with the
// physical range maps.  Token layout is tuple-major (T, n_seq, W): a tuple's tokens are contiguous, so the identity
// permutation of the symmetriser input is the token buffer itself.
#include "common.cuh"

namespace grappa {

// ------------------------------------------------------------------------------------------ tuple gather ------

// one thread per float4 of X; requires W % 4 == 0 and lda % 4 == 0
__global__ void tuple_gather_fwd_kernel(int64_t rows, int n_seq, int F, int W, const float* __restrict__ A, int64_t lda,
                                        const int32_t* __restrict__ idx, const float* __restrict__ pe, float* __restrict__ X) {
    const int W4 = W / 4;
    const int64_t total = rows * W4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / W4;
        const int c = (int)(i % W4) * 4;
        const int64_t a = __ldg(idx + r);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < F) v = __ldg(reinterpret_cast<const float4*>(A + a * lda + c));
        if (c + 3 >= F) {
            const float pv = pe ? __ldg(pe + (r % n_seq)) : 0.f;
            if (c + 0 >= F) v.x = pv;
            if (c + 1 >= F) v.y = pv;
            if (c + 2 >= F) v.z = pv;
            if (c + 3 >= F) v.w = pv;
        }
        reinterpret_cast<float4*>(X)[i] = v;
    }
}

// one warp per atom: dA[a, 0:F] = sum over incident token rows of dX[row, 0:F]; columns >= F of dA are zeroed
__global__ void __launch_bounds__(256) tuple_gather_bwd_kernel(int64_t N, int F, int W, const float* __restrict__ dX,
                                                               const int32_t* __restrict__ inc_ptr, const int32_t* __restrict__ inc_tok,
                                                               float* __restrict__ dA, int64_t ldda) {
    const int64_t a = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (a >= N) return;
    const int W4 = W / 4;
    const int q0 = __ldg(inc_ptr + a), q1 = __ldg(inc_ptr + a + 1);
    for (int c4 = lane; c4 < W4; c4 += 32) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q = q0; q < q1; ++q) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(dX + (int64_t)__ldg(inc_tok + q) * W) + c4);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        const int c = c4 * 4;
        if (c + 0 >= F) s.x = 0.f;
        if (c + 1 >= F) s.y = 0.f;
        if (c + 2 >= F) s.z = 0.f;
        if (c + 3 >= F) s.w = 0.f;
        *reinterpret_cast<float4*>(dA + a * ldda + c) = s;
    }
}


// ------------------------------------------------------------------------------- token de-duplication (row tables) ---
// The first LayerNorm and the QKV projection of a writer's transformer see tokens whose content depends only on
// (atom, positional-encoding value): at most 2 N distinct rows for T * n_seq tokens (4.75x fewer for proper torsions of
// the benchmark molecules, 5.3x for a protein).  They are therefore computed once per table row and gathered per token.

// U[2a + pe, c] = c < F ? P[a, c] : pe          (pe in {0, 1}; W = padded width, multiple of 4)
__global__ void pe_table_fwd_kernel(int64_t N, int F, int W, const float* __restrict__ P, int64_t ldp, float* __restrict__ U) {
    const int W4 = W / 4;
    const int64_t total = 2 * N * W4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / W4;
        const int c = (int)(i % W4) * 4;
        const float pv = (float)(r & 1);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < F) v = __ldg(reinterpret_cast<const float4*>(P + (r >> 1) * ldp + c));
        if (c + 0 >= F) v.x = pv;
        if (c + 1 >= F) v.y = pv;
        if (c + 2 >= F) v.z = pv;
        if (c + 3 >= F) v.w = pv;
        reinterpret_cast<float4*>(U)[i] = v;
    }
}

// dP[a, c] = c < F ? dU[2a, c] + dU[2a + 1, c] : 0     (columns up to lddp are written)
__global__ void pe_table_bwd_kernel(int64_t N, int F, int W, const float* __restrict__ dU, float* __restrict__ dP, int64_t lddp) {
    const int L4 = (int)(lddp / 4);
    const int64_t total = N * L4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = i / L4;
        const int c = (int)(i % L4) * 4;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < W) {
            const float4 u = __ldg(reinterpret_cast<const float4*>(dU + (2 * a) * W + c));
            const float4 v = __ldg(reinterpret_cast<const float4*>(dU + (2 * a + 1) * W + c));
            s = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
        }
        if (c + 0 >= F) s.x = 0.f;
        if (c + 1 >= F) s.y = 0.f;
        if (c + 2 >= F) s.z = 0.f;
        if (c + 3 >= F) s.w = 0.f;
        *reinterpret_cast<float4*>(dP + a * lddp + c) = s;
    }
}

// X[r, :] = Tab[idx[r], :]
__global__ void row_gather_fwd_kernel(int64_t rows, int D, const float* __restrict__ Tab, int64_t ldt,
                                      const int32_t* __restrict__ idx, float* __restrict__ X, int64_t ldx) {
    const int D4 = D / 4;
    const int64_t total = rows * D4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / D4;
        const int c = (int)(i % D4) * 4;
        const float4 v = __ldg(reinterpret_cast<const float4*>(Tab + (int64_t)__ldg(idx + r) * ldt + c));
        *reinterpret_cast<float4*>(X + r * ldx + c) = v;
    }
}

// one warp per (table row, 128-column slab): dTab[j, :] = sum over the rows listed for j of dX[row, :], in list order
__global__ void __launch_bounds__(256) row_gather_bwd_kernel(int64_t R, int D, const float* __restrict__ dX, int64_t ldx,
                                                             const int32_t* __restrict__ ptr, const int32_t* __restrict__ tok,
                                                             float* __restrict__ dTab, int64_t ldt) {
    const int slabs = (D + 127) / 128;
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= R * slabs) return;
    const int64_t j = w / slabs;
    const int c = (int)(w % slabs) * 128 + lane * 4;
    if (c >= D) return;
    const int q0 = __ldg(ptr + j), q1 = __ldg(ptr + j + 1);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int q = q0;
    for (; q + 1 < q1; q += 2) {                   // two independent loads in flight, added in list order
        const float4 u = __ldg(reinterpret_cast<const float4*>(dX + (int64_t)__ldg(tok + q) * ldx + c));
        const float4 v = __ldg(reinterpret_cast<const float4*>(dX + (int64_t)__ldg(tok + q + 1) * ldx + c));
        s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    if (q < q1) {
        const float4 u = __ldg(reinterpret_cast<const float4*>(dX + (int64_t)__ldg(tok + q) * ldx + c));
        s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
    }
    *reinterpret_cast<float4*>(dTab + j * ldt + c) = s;
}

// ------------------------------------------------------------------------------------------ tuple attention ---

__device__ __forceinline__ float sum8(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}
__device__ __forceinline__ float dot8(const float4 (&a)[2], const float4 (&b)[2]) {
    return a[0].x * b[0].x + a[0].y * b[0].y + a[0].z * b[0].z + a[0].w * b[0].w +
           a[1].x * b[1].x + a[1].y * b[1].y + a[1].z * b[1].z + a[1].w * b[1].w;
}
__device__ __forceinline__ void axpy8(float s, const float4 (&x)[2], float4 (&y)[2]) {
#pragma unroll
    for (int r = 0; r < 2; ++r) { y[r].x = fmaf(s, x[r].x, y[r].x); y[r].y = fmaf(s, x[r].y, y[r].y); y[r].z = fmaf(s, x[r].z, y[r].z); y[r].w = fmaf(s, x[r].w, y[r].w); }
}

// head_dim 64: 8 lanes per head, each lane holds float4 #sub and #(sub+8) of the head's 16 float4s.
// One warp handles 4 heads of one tuple.
template <int S>
__global__ void __launch_bounds__(128) tuple_attn_fwd_kernel(int64_t T, int n_heads, const float* __restrict__ qkv,
                                                             const int32_t* __restrict__ row_idx, float* __restrict__ out) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int groups = n_heads / 4;
    if (w >= T * groups) return;
    const int64_t t = w / groups;
    const int head = (int)(w % groups) * 4 + lane / 8;
    const int sub = lane % 8;
    const int E = n_heads * 64;
    float4 q[S][2], k[S][2], v[S][2];
#pragma unroll
    for (int j = 0; j < S; ++j) {
        // row_idx: token -> row of a table of distinct token contents (the gather of the QKV rows is fused in here)
        const int64_t src = row_idx ? (int64_t)__ldg(row_idx + t * S + j) : t * S + j;
        const float4* row = reinterpret_cast<const float4*>(qkv + src * 3 * E + head * 64);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            q[j][r] = __ldg(row + sub + 8 * r);
            k[j][r] = __ldg(row + E / 4 + sub + 8 * r);
            v[j][r] = __ldg(row + 2 * (E / 4) + sub + 8 * r);
        }
    }
#pragma unroll
    for (int i = 0; i < S; ++i) {
        float s[S], mx = -INFINITY;
#pragma unroll
        for (int j = 0; j < S; ++j) { s[j] = sum8(dot8(q[i], k[j])) * 0.125f; mx = fmaxf(mx, s[j]); }
        float den = 0.f;
#pragma unroll
        for (int j = 0; j < S; ++j) { s[j] = expf(s[j] - mx); den += s[j]; }
        const float inv = 1.0f / den;
        float4 o[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
#pragma unroll
        for (int j = 0; j < S; ++j) axpy8(s[j] * inv, v[j], o);
        float4* orow = reinterpret_cast<float4*>(out + (t * S + i) * E + head * 64);
        orow[sub] = o[0];
        orow[sub + 8] = o[1];
    }
}

template <int S>
__global__ void __launch_bounds__(128) tuple_attn_bwd_kernel(int64_t T, int n_heads, const float* __restrict__ qkv,
                                                             const int32_t* __restrict__ row_idx,
                                                             const float* __restrict__ dOut, float* __restrict__ dqkv) {
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int groups = n_heads / 4;
    if (w >= T * groups) return;
    const int64_t t = w / groups;
    const int head = (int)(w % groups) * 4 + lane / 8;
    const int sub = lane % 8;
    const int E = n_heads * 64;
    float4 q[S][2], k[S][2], v[S][2], g[S][2];
#pragma unroll
    for (int j = 0; j < S; ++j) {
        const int64_t src = row_idx ? (int64_t)__ldg(row_idx + t * S + j) : t * S + j;
        const float4* row = reinterpret_cast<const float4*>(qkv + src * 3 * E + head * 64);
        const float4* grow = reinterpret_cast<const float4*>(dOut + (t * S + j) * E + head * 64);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            q[j][r] = __ldg(row + sub + 8 * r);
            k[j][r] = __ldg(row + E / 4 + sub + 8 * r);
            v[j][r] = __ldg(row + 2 * (E / 4) + sub + 8 * r);
            g[j][r] = __ldg(grow + sub + 8 * r);
        }
    }
    float P[S][S], dS[S][S];
#pragma unroll
    for (int i = 0; i < S; ++i) {
        float mx = -INFINITY, den = 0.f;
#pragma unroll
        for (int j = 0; j < S; ++j) { P[i][j] = sum8(dot8(q[i], k[j])) * 0.125f; mx = fmaxf(mx, P[i][j]); }
#pragma unroll
        for (int j = 0; j < S; ++j) { P[i][j] = expf(P[i][j] - mx); den += P[i][j]; }
        const float inv = 1.0f / den;
        float dsum = 0.f;
#pragma unroll
        for (int j = 0; j < S; ++j) {
            P[i][j] *= inv;
            dS[i][j] = sum8(dot8(g[i], v[j]));       // dP_ij
            dsum = fmaf(P[i][j], dS[i][j], dsum);
        }
#pragma unroll
        for (int j = 0; j < S; ++j) dS[i][j] = P[i][j] * (dS[i][j] - dsum) * 0.125f;   // includes the 1/sqrt(64) scale
    }
#pragma unroll
    for (int i = 0; i < S; ++i) {
        float4 dq[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
        float4 dk[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
        float4 dv[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
#pragma unroll
        for (int j = 0; j < S; ++j) {
            axpy8(dS[i][j], k[j], dq);      // dq_i = sum_j dS_ij k_j
            axpy8(dS[j][i], q[j], dk);      // dk_i = sum_j dS_ji q_j
            axpy8(P[j][i], g[j], dv);       // dv_i = sum_j P_ji dO_j
        }
        float4* orow = reinterpret_cast<float4*>(dqkv + (t * S + i) * 3 * E + head * 64);
        orow[sub] = dq[0]; orow[sub + 8] = dq[1];
        orow[E / 4 + sub] = dk[0]; orow[E / 4 + sub + 8] = dk[1];
        orow[2 * (E / 4) + sub] = dv[0]; orow[2 * (E / 4) + sub + 8] = dv[1];
    }
}

// ------------------------------------------------------------------------------------------ symmetriser -------

struct PermTable {
    int perm[6][4];
    int inv[6][4];
    float pref[6];
};

__global__ void symm_permute_fwd_kernel(int64_t T, int n_seq, int W, int P, PermTable pt, const float* __restrict__ X,
                                        float* __restrict__ Y) {
    const int W4 = W / 4;
    const int64_t row4 = (int64_t)n_seq * W4;
    const int64_t total = (int64_t)P * T * row4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = i / row4;
        const int rem = (int)(i % row4);
        const int pos = rem / W4, c4 = rem % W4;
        const int p = (int)(row / T);
        const int64_t t = row % T;
        reinterpret_cast<float4*>(Y)[i] = __ldg(reinterpret_cast<const float4*>(X) + (t * n_seq + pt.perm[p][pos]) * W4 + c4);
    }
}

__global__ void symm_permute_bwd_kernel(int64_t T, int n_seq, int W, int P, PermTable pt, const float* __restrict__ dY,
                                        float* __restrict__ dX) {
    const int W4 = W / 4;
    const int64_t row4 = (int64_t)n_seq * W4;
    const int64_t total = T * row4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i / row4;
        const int rem = (int)(i % row4);
        const int j = rem / W4, c4 = rem % W4;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < P; ++p) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(dY) + ((int64_t)p * T + t) * row4 + pt.inv[p][j] * W4 + c4);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        reinterpret_cast<float4*>(dX)[i] = s;
    }
}

// -------------------------------------------------------------------- fused symmetriser input (permute + LayerNorm)
// perm_equiv_transformer.py:266-279 stacks the permuted copies of every tuple's token row (P*T, n_seq*W) and the first
// FeedForwardLayer normalises each copy (network_utils.py:46).  All permutations of a tuple hold the same multiset of
// values, so mean / rstd are computed ONCE per tuple and the permuted, normalised, affine-transformed rows are written
// directly: the un-normalised permuted copy never exists and the statistics pass runs over T rows instead of P*T.
//   Y[p*T + t, i*W + c] = (X[t, perm[p][i]*W + c] - mean_t) * rstd_t * gamma[i*W + c] + beta[i*W + c]
// One warp per tuple; the row (<= 8 KB) stays L1-resident across the passes.
constexpr float kSymmLnEps = 1e-5f;

__global__ void __launch_bounds__(256) symm_ln_fwd_kernel(int64_t T, int n_seq, int W, int P, PermTable pt,
                                                          const float* __restrict__ X, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, float* __restrict__ Y,
                                                          float* __restrict__ mean, float* __restrict__ rstd) {
    const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= T) return;
    const int W4 = W / 4, D4 = n_seq * W4, D = n_seq * W;
    const float4* x4 = reinterpret_cast<const float4*>(X) + t * D4;
    float s = 0.f, v = 0.f;
    for (int i = lane; i < D4; i += 32) { const float4 q = x4[i]; s += (q.x + q.y) + (q.z + q.w); }
    const float mu = warp_sum(s) / (float)D;
    for (int i = lane; i < D4; i += 32) {
        const float4 q = x4[i];
        const float a = q.x - mu, b = q.y - mu, c = q.z - mu, d = q.w - mu;
        v += (a * a + b * b) + (c * c + d * d);
    }
    const float rs = rsqrtf(warp_sum(v) / (float)D + kSymmLnEps);
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
    for (int p = 0; p < P; ++p) {
        float4* y4 = reinterpret_cast<float4*>(Y) + ((int64_t)p * T + t) * D4;
        for (int i = lane; i < D4; i += 32) {                   // i = destination chunk
            const int pos = i / W4, off = i - pos * W4;
            const float4 q = x4[pt.perm[p][pos] * W4 + off], g = __ldg(g4 + i), bb = __ldg(b4 + i);
            y4[i] = make_float4((q.x - mu) * rs * g.x + bb.x, (q.y - mu) * rs * g.y + bb.y, (q.z - mu) * rs * g.z + bb.z,
                                (q.w - mu) * rs * g.w + bb.w);
        }
    }
    if (lane == 0) { mean[t] = mu; rstd[t] = rs; }
}

// dX[t, s] = rstd (g_s - mean(g) - xhat_s mean(g xhat)),  g_s = sum_p dY[p*T + t, dest_p(s)] * gamma[dest_p(s)]
__global__ void __launch_bounds__(256) symm_ln_bwd_dx_kernel(int64_t T, int n_seq, int W, int P, PermTable pt,
                                                             const float* __restrict__ dY, const float* __restrict__ X,
                                                             const float* __restrict__ gamma, const float* __restrict__ mean,
                                                             const float* __restrict__ rstd, float* __restrict__ dX) {
    const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= T) return;
    const int W4 = W / 4, D4 = n_seq * W4, D = n_seq * W;
    const float4* x4 = reinterpret_cast<const float4*>(X) + t * D4;
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float mu = mean[t], rs = rstd[t];
    auto grad = [&](int i) {                                     // i = source chunk
        const int blk = i / W4, off = i - blk * W4;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = 0; p < P; ++p) {
            const int d = pt.inv[p][blk] * W4 + off;
            const float4 dy = __ldg(reinterpret_cast<const float4*>(dY) + ((int64_t)p * T + t) * D4 + d), ga = __ldg(g4 + d);
            g.x = fmaf(dy.x, ga.x, g.x); g.y = fmaf(dy.y, ga.y, g.y); g.z = fmaf(dy.z, ga.z, g.z); g.w = fmaf(dy.w, ga.w, g.w);
        }
        return g;
    };
    float s1 = 0.f, s2 = 0.f;
    for (int i = lane; i < D4; i += 32) {
        const float4 g = grad(i), q = x4[i];
        s1 += (g.x + g.y) + (g.z + g.w);
        s2 += (g.x * ((q.x - mu) * rs) + g.y * ((q.y - mu) * rs)) + (g.z * ((q.z - mu) * rs) + g.w * ((q.w - mu) * rs));
    }
    s1 = warp_sum(s1) / (float)D;
    s2 = warp_sum(s2) / (float)D;
    float4* o4 = reinterpret_cast<float4*>(dX) + t * D4;
    for (int i = lane; i < D4; i += 32) {
        const float4 g = grad(i), q = x4[i];
        o4[i] = make_float4(rs * (g.x - s1 - (q.x - mu) * rs * s2), rs * (g.y - s1 - (q.y - mu) * rs * s2),
                            rs * (g.z - s1 - (q.z - mu) * rs * s2), rs * (g.w - s1 - (q.w - mu) * rs * s2));
    }
}

// dgamma[d] = sum_p sum_t dY[p*T + t, d] xhat[t, src_p(d)], dbeta[d] = sum_p sum_t dY[p*T + t, d]: partial sums over a
// chunk of tuples per CTA (warps interleave tuples, combined in warp order), then a fixed-order sum over the chunks.
__global__ void __launch_bounds__(256) symm_ln_affine_stage1_kernel(int64_t T, int n_seq, int W, int P, PermTable pt,
                                                                    const float* __restrict__ dY, const float* __restrict__ X,
                                                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                    int rows_per_chunk, float* __restrict__ part_g,
                                                                    float* __restrict__ part_b) {
    __shared__ float4 sh[2][8][32];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int W4 = W / 4, D4 = n_seq * W4;
    const int d4 = blockIdx.x * 32 + lane;
    const bool ok = d4 < D4;
    const int pos = ok ? d4 / W4 : 0, off = ok ? d4 - pos * W4 : 0;
    const int64_t t0 = (int64_t)blockIdx.y * rows_per_chunk, t1 = min(T, t0 + rows_per_chunk);
    float4 sg = make_float4(0.f, 0.f, 0.f, 0.f), sb = sg;
    if (ok) {
        for (int64_t t = t0 + wrp; t < t1; t += 8) {
            const float mu = __ldg(mean + t), rs = __ldg(rstd + t);
            for (int p = 0; p < P; ++p) {
                const float4 dy = __ldg(reinterpret_cast<const float4*>(dY) + ((int64_t)p * T + t) * D4 + d4);
                const float4 q = __ldg(reinterpret_cast<const float4*>(X) + t * D4 + pt.perm[p][pos] * W4 + off);
                sb.x += dy.x; sb.y += dy.y; sb.z += dy.z; sb.w += dy.w;
                sg.x = fmaf(dy.x, (q.x - mu) * rs, sg.x); sg.y = fmaf(dy.y, (q.y - mu) * rs, sg.y);
                sg.z = fmaf(dy.z, (q.z - mu) * rs, sg.z); sg.w = fmaf(dy.w, (q.w - mu) * rs, sg.w);
            }
        }
    }
    sh[0][wrp][lane] = sg; sh[1][wrp][lane] = sb;
    __syncthreads();
    if (wrp < 2 && ok) {
        float4 a = sh[wrp][0][lane];
#pragma unroll
        for (int j = 1; j < 8; ++j) { const float4 u = sh[wrp][j][lane]; a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w; }
        float* part = wrp == 0 ? part_g : part_b;
        reinterpret_cast<float4*>(part + (int64_t)blockIdx.y * n_seq * W)[d4] = a;
    }
}

__global__ void __launch_bounds__(256) symm_ln_affine_stage2_kernel(int64_t R, int64_t D, const float* __restrict__ part_g,
                                                                    const float* __restrict__ part_b, float* __restrict__ dgamma,
                                                                    float* __restrict__ dbeta) {
    const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= D) return;
    const float* part = blockIdx.y == 0 ? part_g : part_b;
    float s = 0.f;
    for (int64_t r = 0; r < R; ++r) s += part[r * D + col];
    (blockIdx.y == 0 ? dgamma : dbeta)[col] = s;
}

// ------------------------------------------------------------------------------------------ finalize ----------

struct FinalizeConsts {
    float eq_a, eq_b, eq_c, k_mos, k_std, k_min, cutoff;
    float pref[6];
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float to_positive(float c, float mos, float sd, float mn) { return sd * (elu1(mos + c - 1.0f) + 1.0f) + mn; }
__device__ __forceinline__ float to_positive_grad(float c, float mos, float sd) { const float z = mos + c - 1.0f; return sd * (z > 0.f ? 1.0f : expf(z)); }

constexpr int kMaxOut = 12;

__global__ void readout_finalize_fwd_kernel(int kind, int64_t T, int P, int n_out, const float* __restrict__ S, FinalizeConsts fc,
                                            const float* __restrict__ kstd, const float* __restrict__ kmean, int n_per, int col0,
                                            float* __restrict__ C, float* __restrict__ eq, float* __restrict__ k) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    float c[kMaxOut];
#pragma unroll
    for (int j = 0; j < kMaxOut; ++j) {
        c[j] = 0.f;
        if (j < n_out) {
            for (int p = 0; p < P; ++p) c[j] += fc.pref[p] * S[((int64_t)p * T + t) * n_out + j];
            C[t * n_out + j] = c[j];
        }
    }
    if (kind == 0) {
        eq[t] = to_positive(c[0], fc.eq_a, fc.eq_b, fc.eq_c);
        k[t] = to_positive(c[1], fc.k_mos, fc.k_std, fc.k_min);
    } else if (kind == 1) {
        eq[t] = fc.eq_b * sigmoidf_(fc.eq_a * c[0]);
        k[t] = to_positive(c[1], fc.k_mos, fc.k_std, fc.k_min);
    } else {
        const int nh = kind == 2 ? n_out / 2 : n_out;
        for (int j = 0; j < n_per; ++j) k[t * n_per + j] = 0.f;
#pragma unroll
        for (int j = 0; j < kMaxOut / 2; ++j) {
            if (j < nh) {
                float v = kind == 2 ? c[j] * sigmoidf_(c[nh + j]) * kstd[j] : c[j] * kstd[j] + kmean[j];
                if (fc.cutoff > 0.f && !(fabsf(v) > fc.cutoff)) v = 0.f;
                k[t * n_per + col0 + j] = v;
            }
        }
    }
}

__global__ void readout_finalize_bwd_kernel(int kind, int64_t T, int P, int n_out, const float* __restrict__ C, FinalizeConsts fc,
                                            const float* __restrict__ kstd, const float* __restrict__ kmean, int n_per, int col0,
                                            const float* __restrict__ d_eq, const float* __restrict__ d_k, float* __restrict__ dS) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    float c[kMaxOut], dc[kMaxOut];
#pragma unroll
    for (int j = 0; j < kMaxOut; ++j) { c[j] = j < n_out ? C[t * n_out + j] : 0.f; dc[j] = 0.f; }
    if (kind == 0) {
        dc[0] = (d_eq ? d_eq[t] : 0.f) * to_positive_grad(c[0], fc.eq_a, fc.eq_b);
        dc[1] = (d_k ? d_k[t] : 0.f) * to_positive_grad(c[1], fc.k_mos, fc.k_std);
    } else if (kind == 1) {
        const float s = sigmoidf_(fc.eq_a * c[0]);
        dc[0] = (d_eq ? d_eq[t] : 0.f) * fc.eq_b * fc.eq_a * s * (1.0f - s);
        dc[1] = (d_k ? d_k[t] : 0.f) * to_positive_grad(c[1], fc.k_mos, fc.k_std);
    } else {
        const int nh = kind == 2 ? n_out / 2 : n_out;
#pragma unroll
        for (int j = 0; j < kMaxOut / 2; ++j) {
            if (j < nh) {
                const float g = d_k ? d_k[t * n_per + col0 + j] : 0.f;
                if (kind == 2) {
                    const float s = sigmoidf_(c[nh + j]);
                    const float v = c[j] * s * kstd[j];
                    const bool pass = !(fc.cutoff > 0.f) || fabsf(v) > fc.cutoff;
                    dc[j] = pass ? g * s * kstd[j] : 0.f;
                    dc[nh + j] = pass ? g * c[j] * s * (1.0f - s) * kstd[j] : 0.f;
                } else {
                    const float v = c[j] * kstd[j] + kmean[j];
                    const bool pass = !(fc.cutoff > 0.f) || fabsf(v) > fc.cutoff;
                    dc[j] = pass ? g * kstd[j] : 0.f;
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kMaxOut; ++j)
        if (j < n_out)
            for (int p = 0; p < P; ++p) dS[((int64_t)p * T + t) * n_out + j] = fc.pref[p] * dc[j];
}

static inline unsigned grid_for(int64_t n, int threads = 256, int64_t cap = 148 * 16) {
    return (unsigned)std::max<int64_t>(1, std::min<int64_t>(cap, ceil_div(n, threads)));
}

static int make_perm_table(int n_seq, int P, const int32_t* perm, PermTable& pt) {
    GRAPPA_CHECK_ARG(P >= 1 && P <= 6 && n_seq >= 1 && n_seq <= 4 && perm, "need 1..6 permutations of 1..4 positions");
    for (int p = 0; p < P; ++p) {
        bool seen[4] = {false, false, false, false};
        for (int i = 0; i < n_seq; ++i) {
            const int j = perm[p * n_seq + i];
            GRAPPA_CHECK_ARG(j >= 0 && j < n_seq && !seen[j], "invalid permutation");
            seen[j] = true;
            pt.perm[p][i] = j;
            pt.inv[p][j] = i;
        }
    }
    return 0;
}

static int make_consts(int P, const float* prefactor, const float* consts, FinalizeConsts& fc) {
    GRAPPA_CHECK_ARG(P >= 1 && P <= 6 && prefactor && consts, "bad prefactor/consts");
    fc.eq_a = consts[0]; fc.eq_b = consts[1]; fc.eq_c = consts[2]; fc.k_mos = consts[3]; fc.k_std = consts[4];
    fc.k_min = consts[5]; fc.cutoff = consts[6];
    for (int p = 0; p < 6; ++p) fc.pref[p] = p < P ? prefactor[p] : 0.f;
    return 0;
}

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_tuple_gather_fwd(int64_t T, int n_seq, int F, int W, const float* A, int64_t lda, const int32_t* idx,
                                       const float* pe, float* X, void* stream) {
    GRAPPA_CHECK_ARG(T >= 0 && n_seq >= 1 && n_seq <= 4 && F > 0 && F <= W, "bad sizes");
    GRAPPA_CHECK_ARG(W % 4 == 0 && lda % 4 == 0 && lda >= ((F + 3) / 4) * 4, "W and lda must be multiples of 4 (lda >= F rounded up)");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(A && idx && X, "null pointer");
    tuple_gather_fwd_kernel<<<grid_for(T * n_seq * (W / 4)), 256, 0, (cudaStream_t)stream>>>(T * n_seq, n_seq, F, W, A, lda, idx, pe, X);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_tuple_gather_bwd(int64_t N, int F, int W, const float* dX, const int32_t* inc_ptr, const int32_t* inc_tok,
                                       float* dA, int64_t ldda, void* stream) {
    GRAPPA_CHECK_ARG(N >= 0 && F > 0 && F <= W && W % 4 == 0 && ldda % 4 == 0 && ldda >= W, "bad sizes");
    if (N == 0) return 0;
    GRAPPA_CHECK_ARG(inc_ptr && inc_tok && dA, "null pointer");
    tuple_gather_bwd_kernel<<<(unsigned)ceil_div(N * 32, 256), 256, 0, (cudaStream_t)stream>>>(N, F, W, dX, inc_ptr, inc_tok, dA, ldda);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_pe_table_fwd(int64_t N, int F, int W, const float* P, int64_t ldp, float* U, void* stream) {
    GRAPPA_CHECK_ARG(N >= 0 && F > 0 && F <= W && W % 4 == 0 && ldp % 4 == 0 && ldp >= ((F + 3) / 4) * 4, "bad sizes");
    if (N == 0) return 0;
    GRAPPA_CHECK_ARG(P && U, "null pointer");
    pe_table_fwd_kernel<<<grid_for(2 * N * (W / 4)), 256, 0, (cudaStream_t)stream>>>(N, F, W, P, ldp, U);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_pe_table_bwd(int64_t N, int F, int W, const float* dU, float* dP, int64_t lddp, void* stream) {
    GRAPPA_CHECK_ARG(N >= 0 && F > 0 && F <= W && W % 4 == 0 && lddp % 4 == 0 && lddp >= W, "bad sizes");
    if (N == 0) return 0;
    GRAPPA_CHECK_ARG(dU && dP, "null pointer");
    pe_table_bwd_kernel<<<grid_for(N * (lddp / 4)), 256, 0, (cudaStream_t)stream>>>(N, F, W, dU, dP, lddp);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_row_gather_fwd(int64_t rows, int D, const float* table, int64_t ldt, const int32_t* idx, float* X,
                                     int64_t ldx, void* stream) {
    GRAPPA_CHECK_ARG(rows >= 0 && D > 0 && D % 4 == 0 && ldt % 4 == 0 && ldx % 4 == 0 && ldt >= D && ldx >= D, "bad sizes");
    if (rows == 0) return 0;
    GRAPPA_CHECK_ARG(table && idx && X, "null pointer");
    row_gather_fwd_kernel<<<grid_for(rows * (D / 4)), 256, 0, (cudaStream_t)stream>>>(rows, D, table, ldt, idx, X, ldx);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_row_gather_bwd(int64_t table_rows, int D, const float* dX, int64_t ldx, const int32_t* ptr,
                                     const int32_t* tok, float* dTable, int64_t ldt, void* stream) {
    GRAPPA_CHECK_ARG(table_rows >= 0 && D > 0 && D % 4 == 0 && ldt % 4 == 0 && ldx % 4 == 0 && ldt >= D && ldx >= D, "bad sizes");
    if (table_rows == 0) return 0;
    GRAPPA_CHECK_ARG(dX && ptr && tok && dTable, "null pointer");
    const int64_t warps = table_rows * ((D + 127) / 128);
    row_gather_bwd_kernel<<<(unsigned)ceil_div(warps * 32, 256), 256, 0, (cudaStream_t)stream>>>(table_rows, D, dX, ldx, ptr, tok, dTable, ldt);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_tuple_attn_fwd(int64_t T, int n_seq, int n_heads, int head_dim, const float* qkv, const int32_t* row_idx,
                                     float* out, void* stream) {
    GRAPPA_CHECK_ARG(head_dim == 64 && n_heads % 4 == 0 && n_heads > 0, "tuple attention needs head_dim 64 and n_heads % 4 == 0");
    GRAPPA_CHECK_ARG(n_seq >= 1 && n_seq <= 4 && T >= 0, "n_seq must be 1..4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(qkv && out, "null pointer");
    const int64_t warps = T * (n_heads / 4);
    const unsigned grid = (unsigned)ceil_div(warps * 32, 128);
    cudaStream_t st = (cudaStream_t)stream;
    switch (n_seq) {
        case 1: tuple_attn_fwd_kernel<1><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, out); break;
        case 2: tuple_attn_fwd_kernel<2><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, out); break;
        case 3: tuple_attn_fwd_kernel<3><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, out); break;
        default: tuple_attn_fwd_kernel<4><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, out); break;
    }
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_tuple_attn_bwd(int64_t T, int n_seq, int n_heads, int head_dim, const float* qkv, const int32_t* row_idx,
                                     const float* dOut, float* dqkv, void* stream) {
    GRAPPA_CHECK_ARG(head_dim == 64 && n_heads % 4 == 0 && n_heads > 0, "tuple attention needs head_dim 64 and n_heads % 4 == 0");
    GRAPPA_CHECK_ARG(n_seq >= 1 && n_seq <= 4 && T >= 0, "n_seq must be 1..4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(qkv && dOut && dqkv, "null pointer");
    const int64_t warps = T * (n_heads / 4);
    const unsigned grid = (unsigned)ceil_div(warps * 32, 128);
    cudaStream_t st = (cudaStream_t)stream;
    switch (n_seq) {
        case 1: tuple_attn_bwd_kernel<1><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, dOut, dqkv); break;
        case 2: tuple_attn_bwd_kernel<2><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, dOut, dqkv); break;
        case 3: tuple_attn_bwd_kernel<3><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, dOut, dqkv); break;
        default: tuple_attn_bwd_kernel<4><<<grid, 128, 0, st>>>(T, n_heads, qkv, row_idx, dOut, dqkv); break;
    }
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_symm_permute_fwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* X, float* Y, void* stream) {
    PermTable pt;
    if (int rc = make_perm_table(n_seq, P, perm, pt)) return rc;
    GRAPPA_CHECK_ARG(T >= 0 && W > 0 && W % 4 == 0, "W must be a multiple of 4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(X && Y, "null pointer");
    symm_permute_fwd_kernel<<<grid_for((int64_t)P * T * n_seq * (W / 4)), 256, 0, (cudaStream_t)stream>>>(T, n_seq, W, P, pt, X, Y);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_symm_permute_bwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* dY, float* dX, void* stream) {
    PermTable pt;
    if (int rc = make_perm_table(n_seq, P, perm, pt)) return rc;
    GRAPPA_CHECK_ARG(T >= 0 && W > 0 && W % 4 == 0, "W must be a multiple of 4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(dY && dX, "null pointer");
    symm_permute_bwd_kernel<<<grid_for(T * n_seq * (W / 4)), 256, 0, (cudaStream_t)stream>>>(T, n_seq, W, P, pt, dY, dX);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

static inline int symm_ln_rows_per_chunk(int64_t T) { return (int)std::max<int64_t>(8, ceil_div(ceil_div(T, 37), 8) * 8); }   // ~37 chunks x 16 column blocks

extern "C" int64_t grappa_symm_ln_bwd_scratch_floats(int64_t T, int n_seq, int W) {
    return 2 * ceil_div(std::max<int64_t>(T, 1), symm_ln_rows_per_chunk(T)) * n_seq * W;
}

extern "C" int grappa_symm_ln_fwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* X, const float* gamma,
                                  const float* beta, float* Y, float* mean, float* rstd, void* stream) {
    PermTable pt;
    if (int rc = make_perm_table(n_seq, P, perm, pt)) return rc;
    GRAPPA_CHECK_ARG(T >= 0 && W > 0 && W % 4 == 0, "W must be a multiple of 4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(X && gamma && beta && Y && mean && rstd, "null pointer");
    symm_ln_fwd_kernel<<<(unsigned)ceil_div(T * 32, 256), 256, 0, (cudaStream_t)stream>>>(T, n_seq, W, P, pt, X, gamma, beta, Y, mean, rstd);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_symm_ln_bwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* dY, const float* X,
                                  const float* gamma, const float* mean, const float* rstd, float* dX, float* dgamma, float* dbeta,
                                  float* scratch, void* stream) {
    PermTable pt;
    if (int rc = make_perm_table(n_seq, P, perm, pt)) return rc;
    GRAPPA_CHECK_ARG(T >= 0 && W > 0 && W % 4 == 0, "W must be a multiple of 4");
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(dY && X && gamma && mean && rstd && (dX || dgamma), "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (dX) {
        symm_ln_bwd_dx_kernel<<<(unsigned)ceil_div(T * 32, 256), 256, 0, st>>>(T, n_seq, W, P, pt, dY, X, gamma, mean, rstd, dX);
        GRAPPA_CHECK_LAUNCH();
    }
    if (dgamma || dbeta) {
        GRAPPA_CHECK_ARG(dgamma && dbeta && scratch, "dgamma, dbeta and scratch must be given together");
        const int rpc = symm_ln_rows_per_chunk(T);
        const int64_t R = ceil_div(T, rpc), D = (int64_t)n_seq * W;
        float* pg = scratch;
        float* pb = scratch + R * D;
        symm_ln_affine_stage1_kernel<<<dim3((unsigned)ceil_div(D / 4, 32), (unsigned)R), 256, 0, st>>>(T, n_seq, W, P, pt, dY, X, mean, rstd, rpc, pg, pb);
        GRAPPA_CHECK_LAUNCH();
        symm_ln_affine_stage2_kernel<<<dim3((unsigned)ceil_div(D, 256), 2), 256, 0, st>>>(R, D, pg, pb, dgamma, dbeta);
        GRAPPA_CHECK_LAUNCH();
    }
    return 0;
}

extern "C" int grappa_readout_finalize_fwd(int kind, int64_t T, int P, int n_out, const float* S, const float* prefactor,
                                           const float* consts, const float* kstd, const float* kmean, int n_per, int col0,
                                           float* C, float* eq, float* k, void* stream) {
    GRAPPA_CHECK_ARG(kind >= 0 && kind <= 3 && n_out >= 1 && n_out <= kMaxOut, "bad kind / n_out");
    FinalizeConsts fc;
    if (int rc = make_consts(P, prefactor, consts, fc)) return rc;
    if (kind <= 1) GRAPPA_CHECK_ARG(n_out >= 2 && eq, "harmonic writers need n_out >= 2 and an eq output");
    if (kind >= 2) {
        const int nh = kind == 2 ? n_out / 2 : n_out;
        GRAPPA_CHECK_ARG(kstd && (kind == 2 || kmean) && (kind != 2 || n_out % 2 == 0), "torsion writer needs k_std (and k_mean when ungated)");
        GRAPPA_CHECK_ARG(col0 >= 0 && col0 + nh <= n_per && nh <= kMaxOut / 2, "torsion columns out of range");
    }
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(S && C && k, "null pointer");
    readout_finalize_fwd_kernel<<<(unsigned)ceil_div(T, 128), 128, 0, (cudaStream_t)stream>>>(kind, T, P, n_out, S, fc, kstd, kmean, n_per, col0, C, eq, k);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_readout_finalize_bwd(int kind, int64_t T, int P, int n_out, const float* C, const float* prefactor,
                                           const float* consts, const float* kstd, const float* kmean, int n_per, int col0,
                                           const float* d_eq, const float* d_k, float* dS, void* stream) {
    GRAPPA_CHECK_ARG(kind >= 0 && kind <= 3 && n_out >= 1 && n_out <= kMaxOut, "bad kind / n_out");
    FinalizeConsts fc;
    if (int rc = make_consts(P, prefactor, consts, fc)) return rc;
    if (T == 0) return 0;
    GRAPPA_CHECK_ARG(C && dS, "null pointer");
    readout_finalize_bwd_kernel<<<(unsigned)ceil_div(T, 128), 128, 0, (cudaStream_t)stream>>>(kind, T, P, n_out, C, fc, kstd, kmean, n_per, col0, d_eq, d_k, dS);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}
