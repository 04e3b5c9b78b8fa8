// K2 (fp32 path): tiled SGEMM with a fused epilogue  C = dropout(act(A.op(B) + bias)) + res1 + res2.
// This is the exact-fp32 FFMA path (math_mode 0); the tcgen05 3xTF32 path lives in gemm_tc.cu (math_mode 1).
// Tiling: BMxBN CTA tile, BK-deep k-slices, 256 threads, TMxTN register tile per thread, shared tiles stored
// k-major ([k][m] / [k][n]) so that the inner product reads float4s without bank conflicts, register-staged
// double buffering of the global loads.  Algorithmic flops 2MNK.
#include "gemm_common.cuh"

namespace grappa {


// Element (row r of the tile operand, k) lives at  base[r*ld + k]  (K_CONTIG)  or  base[k*ld + r]  (!K_CONTIG).
template <int BR, int BK, bool K_CONTIG>
struct TileLoader {
    // 256 threads move BR*BK floats = BR*BK/4 float4s; NV float4 per thread
    static constexpr int NV = BR * BK / 4 / 256;
    static_assert(BR * BK % (4 * 256) == 0, "tile must be a multiple of 1024 floats");
    float4 v[NV];

    __device__ __forceinline__ void load(const float* __restrict__ base, int64_t ld, int64_t r0, int64_t R, int64_t k0,
                                         int64_t K, bool vec_ok, int tid) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = tid + i * 256;
            int64_t r, k;
            const float* ptr;
            int nvalid;   // how many of the 4 consecutive elements are in range
            bool row_ok;
            if (K_CONTIG) {
                r = r0 + f / (BK / 4); k = k0 + (f % (BK / 4)) * 4;
                ptr = base + r * ld + k;
                row_ok = r < R;
                nvalid = (int)min((int64_t)4, K - k);
            } else {
                k = k0 + f / (BR / 4); r = r0 + (f % (BR / 4)) * 4;
                ptr = base + k * ld + r;
                row_ok = k < K;
                nvalid = (int)min((int64_t)4, R - r);
            }
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row_ok && nvalid > 0) {
                if (vec_ok && nvalid == 4) {
                    t = __ldg(reinterpret_cast<const float4*>(ptr));
                } else {
                    t.x = __ldg(ptr);
                    if (nvalid > 1) t.y = __ldg(ptr + 1);
                    if (nvalid > 2) t.z = __ldg(ptr + 2);
                    if (nvalid > 3) t.w = __ldg(ptr + 3);
                }
            }
            v[i] = t;
        }
    }

    // shared tile layout: s[k][BR + PAD]
    template <int LDS>
    __device__ __forceinline__ void store(float* s, int tid) const {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int f = tid + i * 256;
            if (K_CONTIG) {
                const int r = f / (BK / 4), k = (f % (BK / 4)) * 4;
                s[(k + 0) * LDS + r] = v[i].x;
                s[(k + 1) * LDS + r] = v[i].y;
                s[(k + 2) * LDS + r] = v[i].z;
                s[(k + 3) * LDS + r] = v[i].w;
            } else {
                const int k = f / (BR / 4), r = (f % (BR / 4)) * 4;
                *reinterpret_cast<float4*>(&s[k * LDS + r]) = v[i];
            }
        }
    }
};

template <int BM, int BN, int BK, int TM, int TN, bool A_KC, bool B_KC>
__global__ void __launch_bounds__(256) sgemm_kernel(GemmParams p, int64_t k_per_split, float* __restrict__ partial) {
    constexpr int TX = BN / TN, TY = BM / TM;
    static_assert(TX * TY == 256, "256 threads per CTA");
    constexpr int HM = TM / 4, HN = TN / 4;        // number of 4-wide strips per thread
    constexpr int LDA = BM + 4, LDB = BN + 4;      // +4 keeps float4 alignment and skews the transposing stores
    __shared__ __align__(16) float As[2][BK * LDA];
    __shared__ __align__(16) float Bs[2][BK * LDB];

    const int tid = threadIdx.x;
    const int tx = tid % TX, ty = tid / TX;
    const int64_t m0 = (int64_t)blockIdx.y * BM, n0 = (int64_t)blockIdx.x * BN;

    TileLoader<BM, BK, A_KC> la;
    TileLoader<BN, BK, B_KC> lb;
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    // split-K: blockIdx.z owns k in [kbeg, kend); partial tiles are written raw to `partial`
    const int64_t kbeg = (int64_t)blockIdx.z * k_per_split;
    const int64_t kend = min(p.K, kbeg + k_per_split);
    const int64_t nk = (kend - kbeg + BK - 1) / BK;
    la.load(p.A, p.lda, m0, p.M, kbeg, kend, p.a_vec, tid);
    lb.load(p.B, p.ldb, n0, p.N, kbeg, kend, p.b_vec, tid);
    la.template store<LDA>(As[0], tid);
    lb.template store<LDB>(Bs[0], tid);
    __syncthreads();

    for (int64_t kt = 0; kt < nk; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < nk) {
            la.load(p.A, p.lda, m0, p.M, kbeg + (kt + 1) * BK, kend, p.a_vec, tid);
            lb.load(p.B, p.ldb, n0, p.N, kbeg + (kt + 1) * BK, kend, p.b_vec, tid);
        }
        const float* as = As[cur];
        const float* bs = Bs[cur];
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[TM], b[TN];
#pragma unroll
            for (int h = 0; h < HM; ++h) {
                const float4 t = *reinterpret_cast<const float4*>(&as[k * LDA + h * (BM / HM) + ty * 4]);
                a[h * 4 + 0] = t.x; a[h * 4 + 1] = t.y; a[h * 4 + 2] = t.z; a[h * 4 + 3] = t.w;
            }
#pragma unroll
            for (int h = 0; h < HN; ++h) {
                const float4 t = *reinterpret_cast<const float4*>(&bs[k * LDB + h * (BN / HN) + tx * 4]);
                b[h * 4 + 0] = t.x; b[h * 4 + 1] = t.y; b[h * 4 + 2] = t.z; b[h * 4 + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            la.template store<LDA>(As[cur ^ 1], tid);
            lb.template store<LDB>(Bs[cur ^ 1], tid);
        }
        __syncthreads();
    }

    if (partial != nullptr) {
        float* dst = partial + (int64_t)blockIdx.z * p.M * p.N;
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int64_t m = m0 + (i / 4) * (BM / HM) + ty * 4 + (i % 4);
            if (m >= p.M) continue;
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                const int64_t n = n0 + (j / 4) * (BN / HN) + tx * 4 + (j % 4);
                if (n < p.N) dst[m * p.N + n] = acc[i][j];
            }
        }
        return;
    }
    // epilogue
    const bool has_drop = p.drop_p > 0.f;
    const uint64_t seed = has_drop ? p.seed + (p.seed_dev ? *p.seed_dev : 0ull) : 0ull;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int64_t m = m0 + (i / 4) * (BM / HM) + ty * 4 + (i % 4);
        if (m >= p.M) continue;
#pragma unroll
        for (int h = 0; h < HN; ++h) {
            const int64_t n = n0 + h * (BN / HN) + tx * 4;
            if (n >= p.N) continue;
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = acc[i][h * 4 + j];
            const int nvalid = (int)min((int64_t)4, p.N - n);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (j < nvalid) {
                    float z = v[j];
                    if (p.bias) z += __ldg(p.bias + n + j);
                    if (p.act == 1) z = elu1(z);
                    if (has_drop) z = hash_uniform(seed, (uint64_t)(m * p.N + n + j)) >= p.drop_p ? z * p.drop_scale : 0.f;
                    if (p.res1) z += __ldg(p.res1 + m * p.ldr1 + n + j);
                    if (p.res2) z += __ldg(p.res2 + m * p.ldr2 + n + j);
                    if (p.elu_out) z *= elu_grad_from_out(__ldg(p.elu_out + m * p.ld_eo + n + j));
                    v[j] = z;
                }
            }
            float* c = p.C + m * p.ldc + n;
            if (p.c_vec && nvalid == 4) {
                *reinterpret_cast<float4*>(c) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < nvalid) c[j] = v[j];
            }
        }
    }
}

template <int BM, int BN, int BK, int TM, int TN>
static int launch_layout(int layout, const GemmParams& p, cudaStream_t st, float* workspace, int64_t workspace_floats) {
    const int64_t tiles = ceil_div(p.N, BN) * ceil_div(p.M, BM);
    // few output tiles and a long reduction (weight gradients of the narrow output layers): split K over CTAs
    int splits = 1;
    int64_t kps = p.K;
    if (workspace != nullptr && tiles <= 37 && p.K >= 512) {
        const int64_t want = std::min<int64_t>(ceil_div(148, tiles), p.K / 128);
        if (want > 1 && want * p.M * p.N <= workspace_floats) {
            kps = ceil_div(ceil_div(p.K, want), BK) * BK;
            splits = (int)ceil_div(p.K, kps);
        }
    }
    float* partial = splits > 1 ? workspace : nullptr;
    dim3 grid((unsigned)ceil_div(p.N, BN), (unsigned)ceil_div(p.M, BM), (unsigned)splits);
    if (layout == 0) sgemm_kernel<BM, BN, BK, TM, TN, true, true><<<grid, 256, 0, st>>>(p, kps, partial);
    else if (layout == 1) sgemm_kernel<BM, BN, BK, TM, TN, true, false><<<grid, 256, 0, st>>>(p, kps, partial);
    else sgemm_kernel<BM, BN, BK, TM, TN, false, false><<<grid, 256, 0, st>>>(p, kps, partial);
    GRAPPA_CHECK_LAUNCH();
    if (splits > 1) return splitk_reduce(p, splits, workspace, st);
    return 0;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace grappa

using namespace grappa;

static int gemm_impl(int layout, int64_t M, int64_t N, int64_t K,
                           const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                           const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                           float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                           float* workspace, int64_t workspace_floats, void* stream,
                           const float* elu_out, int64_t ld_eo, const float* B_lo = nullptr) {
    GRAPPA_CHECK_ARG(layout >= 0 && layout <= 2, "layout must be 0 (NT), 1 (NN) or 2 (TN)");
    GRAPPA_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "negative size");
    GRAPPA_CHECK_ARG(act == 0 || act == 1, "act must be 0 (none) or 1 (ELU)");
    GRAPPA_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "drop_p must be in [0,1)");
    if (M == 0 || N == 0) return 0;
    GRAPPA_CHECK_ARG(A && B && C, "null matrix pointer");
    GRAPPA_CHECK_ARG(ceil_div(M, 64) <= 65535, "M too large for one launch");
    GemmParams p;
    p.A = A; p.B = B; p.C = C; p.bias = bias; p.res1 = res1; p.res2 = res2;
    p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.ldr1 = ldr1; p.ldr2 = ldr2;
    p.act = act; p.drop_p = drop_p; p.drop_scale = 1.0f / (1.0f - drop_p); p.seed = seed; p.seed_dev = seed_dev;
    p.elu_out = elu_out; p.ld_eo = ld_eo; p.B_lo = B_lo;
    p.cta_budget = (math_mode >> 8) & 0xFF;     // optional: bits 8..15 of math_mode carry the CTA budget of a background GEMM
    math_mode &= 0xFF;
    p.a_vec = aligned16(A) && (lda % 4 == 0);
    p.b_vec = aligned16(B) && (ldb % 4 == 0);
    p.c_vec = aligned16(C) && (ldc % 4 == 0);
    cudaStream_t st = (cudaStream_t)stream;
    {                                                           // one dimension <= 8: streaming kernels (gemm_skinny.cu)
        bool handled = false;
        int rc = gemm_skinny_dispatch(layout, p, st, workspace, workspace_floats, &handled);
        if (rc) return rc;
        if (handled) return 0;
    }
    if (math_mode == 1) {
        bool handled = false;
        int rc = gemm_tc_dispatch(layout, p, st, workspace, workspace_floats, &handled);
        if (rc) return rc;
        if (handled) return 0;
    }
    // pick the large tile only when it still fills the machine (148 SMs) reasonably
    const int64_t big_ctas = ceil_div(M, 128) * ceil_div(N, 128);
    if (big_ctas >= 120) return launch_layout<128, 128, 16, 8, 8>(layout, p, st, workspace, workspace_floats);
    return launch_layout<64, 64, 16, 4, 4>(layout, p, st, workspace, workspace_floats);
}

extern "C" int grappa_gemm(int layout, int64_t M, int64_t N, int64_t K,
                           const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                           const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                           float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                           float* workspace, int64_t workspace_floats, void* stream) {
    return gemm_impl(layout, M, N, K, A, lda, B, ldb, C, ldc, bias, act, res1, ldr1, res2, ldr2, drop_p, seed, seed_dev, math_mode,
                     workspace, workspace_floats, stream, nullptr, 0);
}

extern "C" int grappa_gemm_actgrad(int layout, int64_t M, int64_t N, int64_t K,
                                   const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                                   const float* elu_out, int64_t ld_eo, int math_mode,
                                   float* workspace, int64_t workspace_floats, void* stream) {
    GRAPPA_CHECK_ARG(elu_out != nullptr && ld_eo >= N, "elu_out (M, N) with ld_eo >= N is required");
    return gemm_impl(layout, M, N, K, A, lda, B, ldb, C, ldc, nullptr, 0, nullptr, 0, nullptr, 0, 0.f, 0, nullptr, math_mode,
                     workspace, workspace_floats, stream, elu_out, ld_eo);
}

extern "C" int grappa_gemm_ex(int layout, int64_t M, int64_t N, int64_t K,
                              const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                              const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                              float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                              float* workspace, int64_t workspace_floats,
                              const float* elu_out, int64_t ld_eo, const float* B_lo, void* stream) {
    GRAPPA_CHECK_ARG(elu_out == nullptr || ld_eo >= N, "ld_eo must be >= N");
    return gemm_impl(layout, M, N, K, A, lda, B, ldb, C, ldc, bias, act, res1, ldr1, res2, ldr2, drop_p, seed, seed_dev, math_mode,
                     workspace, workspace_floats, stream, elu_out, ld_eo, B_lo);
}

extern "C" int grappa_gemm_bf16x3(int64_t M, int64_t N, int64_t K, const float* A, int64_t lda,
                                  const uint16_t* B_hi, const uint16_t* B_lo, int64_t ldb16, float* C, int64_t ldc,
                                  const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                                  float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                                  float* workspace, int64_t workspace_floats, const float* elu_out, int64_t ld_eo,
                                  const int32_t* res1_idx, void* stream) {
    GRAPPA_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "negative size");
    GRAPPA_CHECK_ARG(res1_idx == nullptr || res1 != nullptr, "res1_idx without res1");
    GRAPPA_CHECK_ARG(act == 0 || act == 1, "act must be 0 (none) or 1 (ELU)");
    GRAPPA_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "drop_p must be in [0,1)");
    GRAPPA_CHECK_ARG(elu_out == nullptr || ld_eo >= N, "ld_eo must be >= N");
    if (M == 0 || N == 0) return 0;
    GRAPPA_CHECK_ARG(A && B_hi && B_lo && C, "null matrix pointer");
    GRAPPA_CHECK_ARG(aligned16(A) && lda % 4 == 0 && aligned16(B_hi) && aligned16(B_lo) && ldb16 % 8 == 0 && ldb16 >= K,
                     "A must be 16-byte aligned with lda % 4 == 0; the planes 16-byte aligned with ldb16 % 8 == 0, ldb16 >= K");
    GRAPPA_CHECK_ARG(M >= 32 && N >= 32 && K >= 32, "shapes below 32 belong on grappa_gemm");
    GRAPPA_CHECK_ARG(ceil_div(M, 128) <= 65535 && M <= 2147483647LL / 4 && N <= 2147483647LL / 4 && K <= 2147483647LL / 4, "matrix too large");
    GemmParams p;
    p.A = A; p.B = nullptr; p.C = C; p.bias = bias; p.res1 = res1; p.res2 = res2;
    p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = 0; p.ldc = ldc; p.ldr1 = ldr1; p.ldr2 = ldr2;
    p.act = act; p.drop_p = drop_p; p.drop_scale = 1.0f / (1.0f - drop_p); p.seed = seed; p.seed_dev = seed_dev;
    p.elu_out = elu_out; p.ld_eo = ld_eo;
    p.B_hi16 = B_hi; p.B_lo16 = B_lo; p.ldb16 = ldb16; p.res1_idx = res1_idx;
    p.cta_budget = (math_mode >> 8) & 0xFF;
    p.a_vec = 1; p.b_vec = 1;
    p.c_vec = aligned16(C) && (ldc % 4 == 0);
    return gemm_bf16_launch(p, (cudaStream_t)stream, workspace, workspace_floats);
}
