// Row-wise / element-wise kernels around the GEMMs: LayerNorm fwd+bwd, epilogue backward (dropout mask + ELU'),
// deterministic column sums (bias / LayerNorm-affine gradients), atom featurisation, token residuals, AdamW.
// All of these are HBM-bound streaming kernels: coalesced row-major access, one pass over the data where possible,
// two-stage fixed-order reductions instead of atomics.
#include "common.cuh"
#include <mutex>

namespace grappa {

// ------------------------------------------------------------------------------------------- LayerNorm --------

constexpr float kLnEps = 1e-5f;

// one warp per row; the row (<= 8 KB) stays L1-resident across the three passes.  VEC: float4 accesses
// (D % 4 == 0, leading dimensions % 4 == 0, 16-byte aligned bases), 512 contiguous bytes per warp instruction.
template <bool VEC>
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(int64_t M, int D, const float* __restrict__ X, int64_t ldx,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            float* __restrict__ Y, int64_t ldy, float* __restrict__ mean,
                                                            float* __restrict__ rstd) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= M) return;
    const float* x = X + row * ldx;
    float* y = Y + row * ldy;
    float s = 0.f, v = 0.f;
    if (VEC) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const int D4 = D >> 2;
        for (int i = lane; i < D4; i += 32) { const float4 t = x4[i]; s += (t.x + t.y) + (t.z + t.w); }
        const float mu = warp_sum(s) / (float)D;
        for (int i = lane; i < D4; i += 32) {
            const float4 t = x4[i];
            const float a = t.x - mu, b = t.y - mu, c = t.z - mu, d = t.w - mu;
            v += (a * a + b * b) + (c * c + d * d);
        }
        const float rs = rsqrtf(warp_sum(v) / (float)D + kLnEps);
        const float4* g4 = reinterpret_cast<const float4*>(gamma);
        const float4* b4 = reinterpret_cast<const float4*>(beta);
        float4* y4 = reinterpret_cast<float4*>(y);
        for (int i = lane; i < D4; i += 32) {
            const float4 t = x4[i], g = __ldg(g4 + i), bb = __ldg(b4 + i);
            y4[i] = make_float4((t.x - mu) * rs * g.x + bb.x, (t.y - mu) * rs * g.y + bb.y, (t.z - mu) * rs * g.z + bb.z,
                                (t.w - mu) * rs * g.w + bb.w);
        }
        if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
    } else {
        for (int i = lane; i < D; i += 32) s += x[i];
        const float mu = warp_sum(s) / (float)D;
        for (int i = lane; i < D; i += 32) { const float d = x[i] - mu; v = fmaf(d, d, v); }
        const float rs = rsqrtf(warp_sum(v) / (float)D + kLnEps);
        for (int i = lane; i < D; i += 32) y[i] = (x[i] - mu) * rs * __ldg(gamma + i) + __ldg(beta + i);
        if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256) layernorm_bwd_dx_kernel(int64_t M, int D, const float* __restrict__ dY, int64_t lddy,
                                                               const float* __restrict__ X, int64_t ldx,
                                                               const float* __restrict__ gamma, const float* __restrict__ mean,
                                                               const float* __restrict__ rstd, float* __restrict__ dX,
                                                               int64_t lddx) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= M) return;
    const float* x = X + row * ldx;
    const float* dy = dY + row * lddy;
    float* dx = dX + row * lddx;
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
    if (VEC) {
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const float4* d4 = reinterpret_cast<const float4*>(dy);
        const float4* g4 = reinterpret_cast<const float4*>(gamma);
        const int D4 = D >> 2;
        for (int i = lane; i < D4; i += 32) {
            const float4 t = x4[i], d = d4[i], g = __ldg(g4 + i);
            const float a = d.x * g.x, b = d.y * g.y, c = d.z * g.z, e = d.w * g.w;
            s1 += (a + b) + (c + e);
            s2 += (a * ((t.x - mu) * rs) + b * ((t.y - mu) * rs)) + (c * ((t.z - mu) * rs) + e * ((t.w - mu) * rs));
        }
        s1 = warp_sum(s1) / (float)D;
        s2 = warp_sum(s2) / (float)D;
        float4* o4 = reinterpret_cast<float4*>(dx);
        for (int i = lane; i < D4; i += 32) {
            const float4 t = x4[i], d = d4[i], g = __ldg(g4 + i);
            o4[i] = make_float4(rs * (d.x * g.x - s1 - (t.x - mu) * rs * s2), rs * (d.y * g.y - s1 - (t.y - mu) * rs * s2),
                                rs * (d.z * g.z - s1 - (t.z - mu) * rs * s2), rs * (d.w * g.w - s1 - (t.w - mu) * rs * s2));
        }
    } else {
        for (int i = lane; i < D; i += 32) {
            const float g = dy[i] * __ldg(gamma + i);
            s1 += g;
            s2 = fmaf(g, (x[i] - mu) * rs, s2);
        }
        s1 = warp_sum(s1) / (float)D;
        s2 = warp_sum(s2) / (float)D;
        for (int i = lane; i < D; i += 32) {
            const float g = dy[i] * __ldg(gamma + i);
            dx[i] = rs * (g - s1 - (x[i] - mu) * rs * s2);
        }
    }
}

// ------------------------------------------------------------------------------------ column reductions ------
// Vectorised two-stage column reduction (fixed order, no atomics).  A CTA covers 128 columns (one float4 per lane)
// and a chunk of rows (8 warps interleaved over the rows); stage 2 adds the per-chunk partials.
//   MODE 0: out0 = sum_rows X                                   (bias gradient of a layer without activation)
//   MODE 1: out0 = sum_rows dY * xhat, out1 = sum_rows dY       (LayerNorm affine gradients)
//   MODE 2: dZ = dY * keepmask/(1-p) * act'(.), out0 = sum_rows dZ   (epilogue backward fused with the bias gradient)
struct ColArgs {
    int64_t M, N;
    const float* X; int64_t ldx;         // MODE 0: X, MODE 1/2: dY
    const float* X2; int64_t ldx2;       // MODE 1: layer input X, MODE 2: forward output Y
    const float* mean; const float* rstd;
    const float* res1; int64_t ldr1; const float* res2; int64_t ldr2;
    int act; float drop_p; uint64_t seed; const uint64_t* seed_dev;
    float* dZ; int64_t lddz;
    float* part0; float* part1;
    int rows_per_chunk;
};

__device__ __forceinline__ float4 ld4(const float* p, int nvalid) {
    if (nvalid == 4) return *reinterpret_cast<const float4*>(p);
    float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nvalid > 0) r.x = p[0];
    if (nvalid > 1) r.y = p[1];
    if (nvalid > 2) r.z = p[2];
    return r;
}

template <int MODE>
__global__ void __launch_bounds__(256) colreduce_vec_kernel(ColArgs a) {
    __shared__ float4 sh0[8][32];
    __shared__ float4 sh1[8][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t col = (int64_t)blockIdx.x * 128 + lane * 4;
    const int nvalid = (int)max((int64_t)0, min((int64_t)4, a.N - col));
    const int64_t r0 = (int64_t)blockIdx.y * a.rows_per_chunk;
    const int64_t r1 = min(a.M, r0 + a.rows_per_chunk);
    float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = make_float4(0.f, 0.f, 0.f, 0.f);
    const float keep_scale = 1.0f / (1.0f - a.drop_p);
    const uint64_t seed = (MODE == 2 && a.drop_p > 0.f) ? a.seed + (a.seed_dev ? *a.seed_dev : 0ull) : 0ull;
    if (nvalid > 0) {
        for (int64_t r = r0 + w; r < r1; r += 8) {
            float4 v = ld4(a.X + r * a.ldx + col, nvalid);
            if (MODE == 0) {
                s0.x += v.x; s0.y += v.y; s0.z += v.z; s0.w += v.w;
            } else if (MODE == 1) {
                const float4 x = ld4(a.X2 + r * a.ldx2 + col, nvalid);
                const float mu = a.mean[r], rs = a.rstd[r];
                s0.x = fmaf(v.x, (x.x - mu) * rs, s0.x); s0.y = fmaf(v.y, (x.y - mu) * rs, s0.y);
                s0.z = fmaf(v.z, (x.z - mu) * rs, s0.z); s0.w = fmaf(v.w, (x.w - mu) * rs, s0.w);
                s1.x += v.x; s1.y += v.y; s1.z += v.z; s1.w += v.w;
            } else {
                float g[4] = {v.x, v.y, v.z, v.w};
                bool keep[4] = {true, true, true, true};
                if (a.drop_p > 0.f) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        keep[j] = hash_uniform(seed, (uint64_t)(r * a.N + col + j)) >= a.drop_p;
                        g[j] = keep[j] ? g[j] * keep_scale : 0.f;
                    }
                }
                if (a.act == 1) {
                    const float4 y = ld4(a.X2 + r * a.ldx2 + col, nvalid);
                    float o[4] = {y.x, y.y, y.z, y.w};
                    if (a.res1) { const float4 q = ld4(a.res1 + r * a.ldr1 + col, nvalid); o[0] -= q.x; o[1] -= q.y; o[2] -= q.z; o[3] -= q.w; }
                    if (a.res2) { const float4 q = ld4(a.res2 + r * a.ldr2 + col, nvalid); o[0] -= q.x; o[1] -= q.y; o[2] -= q.z; o[3] -= q.w; }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (keep[j]) {
                            float av = o[j];
                            if (a.drop_p > 0.f) av *= (1.0f - a.drop_p);
                            g[j] *= (av > 0.f) ? 1.0f : (av + 1.0f);
                        }
                    }
                }
                float* dz = a.dZ + r * a.lddz + col;
                if (nvalid == 4) *reinterpret_cast<float4*>(dz) = make_float4(g[0], g[1], g[2], g[3]);
                else { if (nvalid > 0) dz[0] = g[0]; if (nvalid > 1) dz[1] = g[1]; if (nvalid > 2) dz[2] = g[2]; }
                s0.x += g[0]; s0.y += g[1]; s0.z += g[2]; s0.w += g[3];
            }
        }
    }
    if (a.part0 == nullptr) return;
    sh0[w][lane] = s0;
    if (MODE == 1) sh1[w][lane] = s1;
    __syncthreads();
    if (w == 0 && nvalid > 0) {
        float4 t0 = sh0[0][lane], t1 = sh1[0][lane];
#pragma unroll
        for (int j = 1; j < 8; ++j) {
            const float4 u = sh0[j][lane];
            t0.x += u.x; t0.y += u.y; t0.z += u.z; t0.w += u.w;
            if (MODE == 1) { const float4 q = sh1[j][lane]; t1.x += q.x; t1.y += q.y; t1.z += q.z; t1.w += q.w; }
        }
        float* p0 = a.part0 + (int64_t)blockIdx.y * a.N + col;
        const float o0[4] = {t0.x, t0.y, t0.z, t0.w};
        for (int j = 0; j < nvalid; ++j) p0[j] = o0[j];
        if (MODE == 1) {
            float* p1 = a.part1 + (int64_t)blockIdx.y * a.N + col;
            const float o1[4] = {t1.x, t1.y, t1.z, t1.w};
            for (int j = 0; j < nvalid; ++j) p1[j] = o1[j];
        }
    }
}

// stage 2: out_k[col] = sum_chunks part_k[chunk][col]; blockIdx.y selects k.  32 columns x 8 chunk-lanes per CTA:
// the chunk loop is 8-way parallel (the partials are L2-resident, the loop is latency-bound), the 8 lane sums are
// combined in lane order, so the result is deterministic.
__global__ void __launch_bounds__(256) colreduce_stage2_vec_kernel(int64_t R, int64_t N, const float* __restrict__ part0,
                                                                   const float* __restrict__ part1, float* __restrict__ out0,
                                                                   float* __restrict__ out1) {
    __shared__ float sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t col = (int64_t)blockIdx.x * 32 + tx;
    const float* part = blockIdx.y == 0 ? part0 : part1;
    float* out = blockIdx.y == 0 ? out0 : out1;
    float s = 0.f;
    if (col < N)
        for (int64_t r = ty; r < R; r += 8) s += part[r * N + col];
    sh[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && col < N) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += sh[j][tx];
        out[col] = t;
    }
}

// rows per chunk so that the stage-1 grid is about two waves of 148 SMs; multiple of 8 (the warps interleave rows)
static inline int col_rows_per_chunk(int64_t M, int64_t N) {
    const int64_t col_blocks = ceil_div(N, 128);
    int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(ceil_div(296, col_blocks), ceil_div(M, 32)));
    int64_t rpc = ceil_div(ceil_div(std::max<int64_t>(M, 1), chunks), 8) * 8;
    return (int)rpc;
}
static inline int64_t col_chunks(int64_t M, int64_t N) { return ceil_div(std::max<int64_t>(M, 1), col_rows_per_chunk(M, N)); }
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Column reductions, stage 1: CTA (32 columns x 8 row-lanes) reduces a chunk of kRowsPerChunk rows.
constexpr int kRowsPerChunk = 256;

// MODE 0: out0 = sum_rows X.   MODE 1: out0 = sum_rows dY * xhat (dgamma), out1 = sum_rows dY (dbeta).
template <int MODE>
__global__ void __launch_bounds__(256) colreduce_stage1_kernel(int64_t M, int64_t N, const float* __restrict__ X, int64_t ldx,
                                                               const float* __restrict__ X2, int64_t ldx2,
                                                               const float* __restrict__ mean, const float* __restrict__ rstd,
                                                               float* __restrict__ part0, float* __restrict__ part1) {
    __shared__ float sh0[8][33];
    __shared__ float sh1[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t col = (int64_t)blockIdx.x * 32 + tx;
    const int64_t r0 = (int64_t)blockIdx.y * kRowsPerChunk;
    const int64_t r1 = min(M, r0 + kRowsPerChunk);
    float a0 = 0.f, a1 = 0.f;
    if (col < N) {
        for (int64_t r = r0 + ty; r < r1; r += 8) {
            const float v = X[r * ldx + col];
            if (MODE == 0) {
                a0 += v;
            } else {
                const float xh = (X2[r * ldx2 + col] - mean[r]) * rstd[r];
                a0 = fmaf(v, xh, a0);
                a1 += v;
            }
        }
    }
    sh0[ty][tx] = a0;
    if (MODE == 1) sh1[ty][tx] = a1;
    __syncthreads();
    if (ty == 0 && col < N) {
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s0 += sh0[j][tx]; if (MODE == 1) s1 += sh1[j][tx]; }
        part0[(int64_t)blockIdx.y * N + col] = s0;
        if (MODE == 1) part1[(int64_t)blockIdx.y * N + col] = s1;
    }
}

// ------------------------------------------------------------------------------------- epilogue backward ------

__global__ void epilogue_bwd_kernel(int64_t M, int64_t N, const float* __restrict__ dY, int64_t lddy,
                                    const float* __restrict__ Y, int64_t ldy, const float* __restrict__ res1, int64_t ldr1,
                                    const float* __restrict__ res2, int64_t ldr2, int act, float drop_p, uint64_t seed0,
                                    const uint64_t* __restrict__ seed_dev, float* __restrict__ dZ, int64_t lddz) {
    const uint64_t seed = seed0 + ((drop_p > 0.f && seed_dev) ? *seed_dev : 0ull);
    const int64_t total = M * N;
    const float scale = 1.0f / (1.0f - drop_p);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t m = i / N, n = i % N;
        float g = dY[m * lddy + n];
        bool keep = true;
        if (drop_p > 0.f) {
            keep = hash_uniform(seed, (uint64_t)i) >= drop_p;
            g = keep ? g * scale : 0.f;
        }
        if (act == 1 && keep) {
            float a = Y[m * ldy + n];
            if (res1) a -= res1[m * ldr1 + n];
            if (res2) a -= res2[m * ldr2 + n];
            if (drop_p > 0.f) a *= (1.0f - drop_p);
            g *= (a > 0.f) ? 1.0f : (a + 1.0f);
        }
        dZ[m * lddz + n] = g;
    }
}

// ------------------------------------------------------------------------------------------- featurize --------

struct FeatBlocks {
    const float* ptr[8];
    int width[8];
    int ld[8];
    int start[9];
    int n;
};

__global__ void featurize_kernel(int64_t n_atoms, FeatBlocks fb, const float* __restrict__ charge, float* __restrict__ out,
                                 int64_t ldo) {
    const int W = fb.start[fb.n] + 16;
    const int64_t total = n_atoms * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = i / W;
        const int c = (int)(i % W);
        float v = 0.f;
        if (c < fb.start[fb.n]) {
            int b = 0;
            while (c >= fb.start[b + 1]) ++b;
            if (fb.ptr[b]) v = fb.ptr[b][a * fb.ld[b] + (c - fb.start[b])];
        } else {
            // ChargeEncoding (graph_attention.py:446-464): clamp to [-2,2], normalise to [0,1], 15 bins
            const float q = fminf(fmaxf(charge[a], -2.f), 2.f);
            const float nv = (q + 2.f) / 4.f;
            const int j = c - fb.start[fb.n];
            if (j == 0) {
                v = nv;
            } else {
                int bin = (int)floorf(nv * 15.f);
                bin = min(max(bin, 0), 14);
                v = (j - 1 == bin) ? 1.f : 0.f;
            }
        }
        out[a * ldo + c] = v;
    }
}

// ------------------------------------------------------------------------------------- token residual ---------

__global__ void token_residual_kernel(int64_t M, int W, int F, const float* __restrict__ X3, const float* __restrict__ X,
                                      const int32_t* __restrict__ row_idx, float* __restrict__ R) {
    const int64_t total = M * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % W);
        // row_idx: token -> row of the table of distinct token contents (the gather of the transformer input is fused in)
        const float x = row_idx ? X[(int64_t)__ldg(row_idx + i / W) * W + c] : X[i];
        R[i] = X3[i] + (c < F ? x + x : x);
    }
}

__global__ void token_residual_bwd_kernel(int64_t M, int W, int F, const float* __restrict__ dR, float* __restrict__ dX) {
    const int64_t total = M * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % W);
        const float g = dR[i];
        dX[i] = c < F ? g + g : g;
    }
}

// ----------------------------------------------------------------------------------------------- AdamW --------

__global__ void adamw_kernel(int64_t n, float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                             float* __restrict__ v, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt,
                             float gscale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i] * gscale;
        float pi = p[i];
        pi *= (1.0f - lr * wd);
        const float mi = b1 * m[i] + (1.0f - b1) * gi;
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = pi - (lr / bc1) * (mi / denom);
    }
}

static inline unsigned grid_for(int64_t n, int threads = 256, int64_t cap = 148 * 16) {
    return (unsigned)std::max<int64_t>(1, std::min<int64_t>(cap, ceil_div(n, threads)));
}

}  // namespace grappa

using namespace grappa;

extern "C" int grappa_layernorm_fwd(int64_t M, int64_t D, const float* X, int64_t ldx, const float* gamma, const float* beta,
                                    float* Y, int64_t ldy, float* mean, float* rstd, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && D > 0 && D <= (1 << 20), "bad sizes");
    if (M == 0) return 0;
    GRAPPA_CHECK_ARG(X && gamma && beta && Y && mean && rstd, "null pointer");
    const bool vec = D % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0 && al16(X) && al16(Y) && al16(gamma) && al16(beta);
    const unsigned grid = (unsigned)ceil_div(M * 32, 256);
    if (vec) layernorm_fwd_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(M, (int)D, X, ldx, gamma, beta, Y, ldy, mean, rstd);
    else layernorm_fwd_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(M, (int)D, X, ldx, gamma, beta, Y, ldy, mean, rstd);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

// dX and the stage-1 partial sums of dgamma / dbeta in ONE pass over dY and X (D = 32 * 4 * NV <= 512): a CTA owns a
// chunk of rows, its 8 warps interleave them; every lane keeps the row's values in registers (dY and X are read once),
// accumulates its columns' dgamma / dbeta partials over the warp's rows, and the 8 warps' partials are combined in warp
// order through shared memory.  The separate reduction pass re-read both tensors (2 of the 5 tensor passes of the op).
template <int NV>
__global__ void __launch_bounds__(256) layernorm_bwd_fused_kernel(int64_t M, const float* __restrict__ dY, int64_t lddy,
                                                                  const float* __restrict__ X, int64_t ldx,
                                                                  const float* __restrict__ gamma, const float* __restrict__ mean,
                                                                  const float* __restrict__ rstd, float* __restrict__ dX, int64_t lddx,
                                                                  int rows_per_chunk, float* __restrict__ part_g,
                                                                  float* __restrict__ part_b) {
    constexpr int D = 128 * NV;
    __shared__ float4 sh[8][32 * NV];
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
    float4 g[NV], ag[NV], ab[NV];
#pragma unroll
    for (int m = 0; m < NV; ++m) {
        g[m] = __ldg(reinterpret_cast<const float4*>(gamma) + lane + 32 * m);
        ag[m] = make_float4(0.f, 0.f, 0.f, 0.f); ab[m] = ag[m];
    }
    for (int64_t row = r0 + wrp; row < r1; row += 8) {
        const float4* x4 = reinterpret_cast<const float4*>(X + row * ldx);
        const float4* d4 = reinterpret_cast<const float4*>(dY + row * lddy);
        const float mu = mean[row], rs = rstd[row];
        float4 xh[NV], dy[NV];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int m = 0; m < NV; ++m) {
            const float4 t = x4[lane + 32 * m];
            dy[m] = d4[lane + 32 * m];
            xh[m] = make_float4((t.x - mu) * rs, (t.y - mu) * rs, (t.z - mu) * rs, (t.w - mu) * rs);
            const float a = dy[m].x * g[m].x, b = dy[m].y * g[m].y, c = dy[m].z * g[m].z, e = dy[m].w * g[m].w;
            s1 += (a + b) + (c + e);
            s2 += (a * xh[m].x + b * xh[m].y) + (c * xh[m].z + e * xh[m].w);
            ab[m].x += dy[m].x; ab[m].y += dy[m].y; ab[m].z += dy[m].z; ab[m].w += dy[m].w;
            ag[m].x = fmaf(dy[m].x, xh[m].x, ag[m].x); ag[m].y = fmaf(dy[m].y, xh[m].y, ag[m].y);
            ag[m].z = fmaf(dy[m].z, xh[m].z, ag[m].z); ag[m].w = fmaf(dy[m].w, xh[m].w, ag[m].w);
        }
        s1 = warp_sum(s1) / (float)D;
        s2 = warp_sum(s2) / (float)D;
        float4* o4 = reinterpret_cast<float4*>(dX + row * lddx);
#pragma unroll
        for (int m = 0; m < NV; ++m)
            o4[lane + 32 * m] = make_float4(rs * (dy[m].x * g[m].x - s1 - xh[m].x * s2), rs * (dy[m].y * g[m].y - s1 - xh[m].y * s2),
                                            rs * (dy[m].z * g[m].z - s1 - xh[m].z * s2), rs * (dy[m].w * g[m].w - s1 - xh[m].w * s2));
    }
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
#pragma unroll
        for (int m = 0; m < NV; ++m) sh[wrp][lane + 32 * m] = which == 0 ? ag[m] : ab[m];
        __syncthreads();
        float* part = (which == 0 ? part_g : part_b) + (int64_t)blockIdx.x * D;
        for (int i = threadIdx.x; i < 32 * NV; i += 256) {
            float4 a = sh[0][i];
#pragma unroll
            for (int j = 1; j < 8; ++j) { const float4 u = sh[j][i]; a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w; }
            reinterpret_cast<float4*>(part)[i] = a;
        }
        __syncthreads();
    }
}

// rows per CTA of the fused kernel: about two waves of CTAs, multiple of 8 (the warps interleave rows)
static inline int ln_fused_rows_per_chunk(int64_t M) {
    const int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(296, ceil_div(M, 8)));
    return (int)(ceil_div(ceil_div(std::max<int64_t>(M, 1), chunks), 8) * 8);
}
static inline bool ln_fused_ok(int64_t D) { return D == 128 || D == 256 || D == 384 || D == 512; }

extern "C" int64_t grappa_layernorm_bwd_scratch_floats(int64_t M, int64_t D) {
    int64_t R = std::max(col_chunks(M, D), ceil_div(std::max<int64_t>(M, 1), kRowsPerChunk));
    if (ln_fused_ok(D)) R = std::max(R, ceil_div(std::max<int64_t>(M, 1), ln_fused_rows_per_chunk(M)));
    return 2 * R * D;
}

extern "C" int grappa_layernorm_bwd(int64_t M, int64_t D, const float* dY, int64_t lddy, const float* X, int64_t ldx,
                                    const float* gamma, const float* mean, const float* rstd,
                                    float* dX, int64_t lddx, float* dgamma, float* dbeta, float* scratch, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && D > 0, "bad sizes");
    GRAPPA_CHECK_ARG(dY && X && gamma && mean && rstd && (dX || dgamma), "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = D % 4 == 0 && ldx % 4 == 0 && lddy % 4 == 0 && lddx % 4 == 0 && al16(X) && al16(dY) && al16(dX) && al16(gamma);
    if (M > 0 && dX != nullptr && dgamma != nullptr && vec && ln_fused_ok(D)) {      // everything in one pass over dY and X
        GRAPPA_CHECK_ARG(dbeta && scratch, "dgamma, dbeta and scratch must be given together");
        const int rpc = ln_fused_rows_per_chunk(M);
        const int64_t R = ceil_div(M, rpc);
        float* pg = scratch;
        float* pb = scratch + R * D;
        switch (D / 128) {
            case 1: layernorm_bwd_fused_kernel<1><<<(unsigned)R, 256, 0, st>>>(M, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx, rpc, pg, pb); break;
            case 2: layernorm_bwd_fused_kernel<2><<<(unsigned)R, 256, 0, st>>>(M, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx, rpc, pg, pb); break;
            case 3: layernorm_bwd_fused_kernel<3><<<(unsigned)R, 256, 0, st>>>(M, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx, rpc, pg, pb); break;
            default: layernorm_bwd_fused_kernel<4><<<(unsigned)R, 256, 0, st>>>(M, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx, rpc, pg, pb); break;
        }
        GRAPPA_CHECK_LAUNCH();
        colreduce_stage2_vec_kernel<<<dim3((unsigned)ceil_div(D, 32), 2), 256, 0, st>>>(R, D, pg, pb, dgamma, dbeta);
        GRAPPA_CHECK_LAUNCH();
        return 0;
    }
    if (M > 0 && dX != nullptr) {          // dX == NULL: only the affine reductions (issued on a side stream by the caller)
        const unsigned grid = (unsigned)ceil_div(M * 32, 256);
        if (vec) layernorm_bwd_dx_kernel<true><<<grid, 256, 0, st>>>(M, (int)D, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx);
        else layernorm_bwd_dx_kernel<false><<<grid, 256, 0, st>>>(M, (int)D, dY, lddy, X, ldx, gamma, mean, rstd, dX, lddx);
        GRAPPA_CHECK_LAUNCH();
    }
    if (dgamma || dbeta) {
        GRAPPA_CHECK_ARG(dgamma && dbeta && scratch, "dgamma, dbeta and scratch must be given together");
        if (vec) {
            const int64_t R = col_chunks(M, D);
            ColArgs a = {};
            a.M = M; a.N = D; a.X = dY; a.ldx = lddy; a.X2 = X; a.ldx2 = ldx; a.mean = mean; a.rstd = rstd;
            a.part0 = scratch; a.part1 = scratch + R * D; a.rows_per_chunk = col_rows_per_chunk(M, D);
            dim3 grid((unsigned)ceil_div(D, 128), (unsigned)R);
            colreduce_vec_kernel<1><<<grid, 256, 0, st>>>(a);
            GRAPPA_CHECK_LAUNCH();
            colreduce_stage2_vec_kernel<<<dim3((unsigned)ceil_div(D, 32), 2), 256, 0, st>>>(R, D, a.part0, a.part1, dgamma, dbeta);
            GRAPPA_CHECK_LAUNCH();
        } else {
            const int64_t R = ceil_div(std::max<int64_t>(M, 1), kRowsPerChunk);
            float* p0 = scratch;
            float* p1 = scratch + R * D;
            dim3 grid((unsigned)ceil_div(D, 32), (unsigned)R);
            colreduce_stage1_kernel<1><<<grid, 256, 0, st>>>(M, D, dY, lddy, X, ldx, mean, rstd, p0, p1);
            GRAPPA_CHECK_LAUNCH();
            colreduce_stage2_vec_kernel<<<dim3((unsigned)ceil_div(D, 32), 2), 256, 0, st>>>(R, D, p0, p1, dgamma, dbeta);
            GRAPPA_CHECK_LAUNCH();
        }
    }
    return 0;
}

// The one place that decides how a column sum is laid out: which stage-1 kernel runs, into how many row chunks the
// rows are cut, and therefore how many floats of partial sums stage 1 writes.  grappa_colsum launches from it,
// grappa_colsum_noalloc checks its fixed array against it, grappa_colsum_scratch_floats reports its worst case.
struct ColsumPlan {
    bool vec;          // float4 path (16-byte aligned input, ldx % 4 == 0)
    int64_t R;         // row chunks = rows of partial sums
    int64_t floats;    // scratch floats touched: R * N
};
static ColsumPlan colsum_plan(int64_t M, int64_t N, const float* X, int64_t ldx) {
    ColsumPlan p;
    p.vec = ldx % 4 == 0 && al16(X);
    p.R = p.vec ? col_chunks(M, N) : ceil_div(std::max<int64_t>(M, 1), kRowsPerChunk);
    p.floats = p.R * N;
    return p;
}

extern "C" int64_t grappa_colsum_scratch_floats(int64_t M, int64_t N) {
    return std::max(col_chunks(M, N), ceil_div(std::max<int64_t>(M, 1), kRowsPerChunk)) * N;    // either path of colsum_plan
}

static int colsum_launch(const ColsumPlan& cp, int64_t M, int64_t N, const float* X, int64_t ldx, float* out, float* scratch,
                         cudaStream_t st) {
    GRAPPA_CHECK_ARG(cp.R <= 65535, "too many rows for one launch");
    if (cp.vec) {
        ColArgs a = {};
        a.M = M; a.N = N; a.X = X; a.ldx = ldx; a.part0 = scratch; a.rows_per_chunk = col_rows_per_chunk(M, N);
        colreduce_vec_kernel<0><<<dim3((unsigned)ceil_div(N, 128), (unsigned)cp.R), 256, 0, st>>>(a);
        GRAPPA_CHECK_LAUNCH();
    } else {
        dim3 grid((unsigned)ceil_div(N, 32), (unsigned)cp.R);
        colreduce_stage1_kernel<0><<<grid, 256, 0, st>>>(M, N, X, ldx, nullptr, 0, nullptr, nullptr, scratch, nullptr);
        GRAPPA_CHECK_LAUNCH();
    }
    colreduce_stage2_vec_kernel<<<dim3((unsigned)ceil_div(N, 32), 1), 256, 0, st>>>(cp.R, N, scratch, nullptr, out, nullptr);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_colsum(int64_t M, int64_t N, const float* X, int64_t ldx, float* out, float* scratch, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && N > 0, "bad sizes");
    GRAPPA_CHECK_ARG(ldx >= N, "ldx must be >= N");
    GRAPPA_CHECK_ARG(X && out && scratch, "null pointer");
    return colsum_launch(colsum_plan(M, N, X, ldx), M, N, X, ldx, out, scratch, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------- allocation-free column sum ---
// Scratch for grappa_colsum_noalloc.  It lives in the library image: every GPU that loads libgrappa_b200.so carries
// these 256 KiB until the process exits.  Host code may only use this name to look the array up (cudaGetSymbolAddress);
// the device address that kernels need is fetched once per device below.
constexpr int64_t kNoAllocScratchFloats = 64 * 1024;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

static std::mutex g_noalloc_mutex;
static float* g_noalloc_ptr[64] = {nullptr};

// Device address of the static scratch on the CURRENT device; resolved on the first call per device.
static int noalloc_scratch(float** out) {
    int dev = -1;
    GRAPPA_CUDA(cudaGetDevice(&dev));
    GRAPPA_CHECK_ARG(dev >= 0 && dev < 64, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        float* ptr = nullptr;
        GRAPPA_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

extern "C" int grappa_colsum_noalloc(int64_t M, int64_t N, const float* X, int64_t ldx, float* out, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && N > 0, "bad sizes");
    GRAPPA_CHECK_ARG(X && out, "null pointer");
    GRAPPA_CHECK_ARG(ldx >= N, "ldx must be >= N");
    const ColsumPlan cp = colsum_plan(M, N, X, ldx);      // the same plan grappa_colsum launches from
    const int64_t need = cp.floats;
    if (need > kNoAllocScratchFloats) {
        snprintf(g_err, sizeof(g_err),
                 "grappa_colsum_noalloc: a %lld x %lld column sum needs %lld floats of scratch but the built-in array holds %lld; "
                 "use grappa_colsum with caller-provided scratch", (long long)M, (long long)N, (long long)need,
                 (long long)kNoAllocScratchFloats);
        return 2;
    }
    float* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    // Nothing has to be cleared: stage 1 writes every partial that stage 2 reads, so leftovers of earlier calls are dead.
    return colsum_launch(cp, M, N, X, ldx, out, scratch, (cudaStream_t)stream);
}

extern "C" int grappa_epilogue_bwd(int64_t M, int64_t N, const float* dY, int64_t lddy, const float* Y, int64_t ldy,
                                   const float* res1, int64_t ldr1, const float* res2, int64_t ldr2, int act,
                                   float drop_p, uint64_t seed, const uint64_t* seed_dev, float* dZ, int64_t lddz,
                                   float* colsum_out, float* scratch, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && N >= 0, "bad sizes");
    GRAPPA_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "drop_p must be in [0,1)");
    GRAPPA_CHECK_ARG(colsum_out == nullptr || scratch != nullptr, "column sums need scratch (grappa_colsum_scratch_floats)");
    if (M * N == 0) return 0;
    GRAPPA_CHECK_ARG(dY && dZ && (act == 0 || Y), "null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec = lddy % 4 == 0 && lddz % 4 == 0 && al16(dY) && al16(dZ) && (act == 0 || (ldy % 4 == 0 && al16(Y))) &&
                     (!res1 || (ldr1 % 4 == 0 && al16(res1))) && (!res2 || (ldr2 % 4 == 0 && al16(res2)));
    if (vec) {
        const int64_t R = col_chunks(M, N);
        GRAPPA_CHECK_ARG(R <= 65535, "too many rows for one launch");
        ColArgs a = {};
        a.M = M; a.N = N; a.X = dY; a.ldx = lddy; a.X2 = Y; a.ldx2 = ldy; a.res1 = res1; a.ldr1 = ldr1; a.res2 = res2; a.ldr2 = ldr2;
        a.act = act; a.drop_p = drop_p; a.seed = seed; a.seed_dev = seed_dev; a.dZ = dZ; a.lddz = lddz;
        a.part0 = colsum_out ? scratch : nullptr; a.rows_per_chunk = col_rows_per_chunk(M, N);
        colreduce_vec_kernel<2><<<dim3((unsigned)ceil_div(N, 128), (unsigned)R), 256, 0, st>>>(a);
        GRAPPA_CHECK_LAUNCH();
        if (colsum_out) {
            colreduce_stage2_vec_kernel<<<dim3((unsigned)ceil_div(N, 32), 1), 256, 0, st>>>(R, N, scratch, nullptr, colsum_out, nullptr);
            GRAPPA_CHECK_LAUNCH();
        }
        return 0;
    }
    epilogue_bwd_kernel<<<grid_for(M * N), 256, 0, st>>>(M, N, dY, lddy, Y, ldy, res1, ldr1, res2, ldr2, act, drop_p, seed, seed_dev, dZ, lddz);
    GRAPPA_CHECK_LAUNCH();
    if (colsum_out) return grappa_colsum(M, N, dZ, lddz, colsum_out, scratch, stream);
    return 0;
}

extern "C" int grappa_featurize(int64_t n_atoms, int n_blocks, const float* const* block_ptr, const int32_t* block_width,
                                const int32_t* block_ld, const float* charge, float* out, int64_t ldo, void* stream) {
    GRAPPA_CHECK_ARG(n_blocks >= 0 && n_blocks <= 8, "at most 8 feature blocks");
    GRAPPA_CHECK_ARG(n_atoms >= 0, "bad size");
    if (n_atoms == 0) return 0;
    GRAPPA_CHECK_ARG(charge && out, "null pointer");
    FeatBlocks fb;
    fb.n = n_blocks;
    fb.start[0] = 0;
    for (int i = 0; i < 8; ++i) {
        fb.ptr[i] = i < n_blocks ? block_ptr[i] : nullptr;
        fb.width[i] = i < n_blocks ? block_width[i] : 0;
        fb.ld[i] = i < n_blocks ? block_ld[i] : 0;
        fb.start[i + 1] = fb.start[i] + fb.width[i];
    }
    GRAPPA_CHECK_ARG(ldo >= fb.start[n_blocks] + 16, "output leading dimension too small");
    featurize_kernel<<<grid_for(n_atoms * (fb.start[n_blocks] + 16)), 256, 0, (cudaStream_t)stream>>>(n_atoms, fb, charge, out, ldo);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_token_residual(int64_t M, int W, int F, const float* X3, const float* X, const int32_t* row_idx, float* R,
                                     void* stream) {
    if (M == 0) return 0;
    GRAPPA_CHECK_ARG(X3 && X && R && W > 0, "bad arguments");
    token_residual_kernel<<<grid_for(M * W), 256, 0, (cudaStream_t)stream>>>(M, W, F, X3, X, row_idx, R);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_token_residual_bwd(int64_t M, int W, int F, const float* dR, float* dX, void* stream) {
    if (M == 0) return 0;
    GRAPPA_CHECK_ARG(dR && dX && W > 0, "bad arguments");
    token_residual_bwd_kernel<<<grid_for(M * W), 256, 0, (cudaStream_t)stream>>>(M, W, F, dR, dX);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}

extern "C" int grappa_adamw(int64_t n, float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                            float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step, float grad_scale,
                            void* stream) {
    if (n == 0) return 0;
    GRAPPA_CHECK_ARG(param && grad && exp_avg && exp_avg_sq && step >= 1, "bad arguments");
    const float bc1 = 1.0f - powf(beta1, (float)step);
    const float bc2 = 1.0f - powf(beta2, (float)step);
    adamw_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(n, param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps,
                                                                   weight_decay, bc1, sqrtf(bc2), grad_scale);
    GRAPPA_CHECK_LAUNCH();
    return 0;
}
