/*
 * grappa_b200 — C ABI of the B200-native Grappa hot path (libgrappa_b200.so).
 *
 * The reference (graeter-group/grappa) is pure Python: its "FFI" for this path is the set of PyTorch / DGL library
 * calls made by src/grappa/models/*.py.  Every entry point below replaces one such call site (cited per function,
 * paths relative to the reference root).  The Python wrappers in grappa_b200/ bind these with ctypes; they are
 * the only callers.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  Unless a parameter is marked HOST it is a device pointer.
 *   - all floating point is fp32, indices are int32 on the device (the graph's int64 idxs are narrowed once by the
 *     host plan), row-major, leading dimensions in elements.
 *   - `stream` is a cudaStream_t passed as void*.  Calls are asynchronous on that stream and hold no global device
 *     state -- with ONE exception, grappa_colsum_noalloc, whose scratch is a fixed array inside the library image
 *     (shared by all its callers; see its declaration).
 *   - return value 0 = success; otherwise grappa_last_error() (thread-local) describes the failure.
 *   - nothing here allocates device memory: outputs and scratch are caller-provided (grappa_colsum_noalloc brings its
 *     scratch with the library instead of taking it from the caller; it does not allocate either).
 */
#ifndef GRAPPA_B200_H
#define GRAPPA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* grappa_last_error(void);
int grappa_abi_version(void);
/* number of kernel launches issued through this library by the calling process (for bench.py's gpu_launches) */
int64_t grappa_launch_count(void);

/* ------------------------------------------------------------------------------------------------------------
 * Dense layers.  Replaces torch.nn.Linear / F.linear + ELU + Dropout + residual adds at
 *   models/graph_attention.py:179,300-307,315-321,190  models/interaction_parameters.py:172
 *   models/network_utils.py:47-53,123-134  models/perm_equiv_transformer.py:282
 * C[M,N] = dropout(act(A·op(B) + bias)) + res1 + res2
 *   layout 0 ("NT"): A is (M,K) row-major, B is (N,K) row-major (a torch Linear weight)      -> forward
 *   layout 1 ("NN"): A is (M,K) row-major, B is (K,N) row-major                               -> dX = dZ·W
 *   layout 2 ("TN"): A is (K,M) row-major (used transposed), B is (K,N) row-major             -> dW = dZ^T·X
 * act: 0 none, 1 ELU(alpha=1).  bias (N) / res1 / res2 may be NULL.  drop_p in [0,1): keep-mask from a
 * counter-based hash of (seed + *seed_dev, row*N+col), kept values scaled by 1/(1-p).  seed_dev (device, may be
 * NULL) lets a captured CUDA graph draw fresh masks on every replay: the step driver bumps it between steps.
 * math_mode: low byte 0 = fp32 FFMA tiles, 1 = tcgen05 3xTF32 (split-precision tensor-core tiles, fp32-level accuracy);
 *            bits 8..15 (optional, 0 = unset): CTA budget of a BACKGROUND GEMM (weight gradients issued on a side
 *            stream): its split-K grid is sized for that many CTAs instead of all 148 SMs, so that the kernels of the
 *            critical chain always find free SMs.  Higher bits are ignored (callers may use them for bookkeeping).
 * workspace (may be NULL): caller-provided scratch of workspace_floats floats; lets the tensor-core path split the k
 * range over several CTAs when the output has few tiles (weight gradients), reduced in fixed order.
 */
int grappa_gemm(int layout, int64_t M, int64_t N, int64_t K,
                const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                float* workspace, int64_t workspace_floats, void* stream);

/* Dense layer with BF16 operand planes on tcgen05 kind::f16:  C[M,N] = epi( A[M,K] . Bw[N,K]^T ) with
 * A B ~ A_hi B_hi + A_hi B_lo + A_lo B_hi, hi = rn_bf16(x), lo = rn_bf16(x - hi) (16 mantissa bits per operand, fp32
 * accumulation; end-to-end error of the bonded parameters 3e-6 .. 3e-5 relative, profiles/r02_split_precision.md).  Same
 * call sites as grappa_gemm, for the layers whose B operand is a WEIGHT: B_hi / B_lo are its (N, K) bf16 planes, leading
 * dimension ldb16, made once per optimiser step by grappa_split_bf16 (the forward uses the planes of W, the dX GEMM the
 * planes of W^T, so both read K-major planes); A is fp32 and is split inside the kernel.  Epilogue, math_mode bits 8..15
 * (CTA budget) and workspace as grappa_gemm; elu_out / ld_eo as grappa_gemm_actgrad (may be NULL).
 * res1_idx (may be NULL): row m of the first residual is res1[res1_idx[m], :] -- the residual is a table of distinct token
 * contents gathered inside the epilogue (readout: out_proj + LN(x) rows, interaction_parameters.py:184-189 fused away). */
int grappa_gemm_bf16x3(int64_t M, int64_t N, int64_t K, const float* A, int64_t lda,
                       const uint16_t* B_hi, const uint16_t* B_lo, int64_t ldb16, float* C, int64_t ldc,
                       const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                       float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                       float* workspace, int64_t workspace_floats, const float* elu_out, int64_t ld_eo,
                       const int32_t* res1_idx, void* stream);

/* bf16 planes of a batch of weight matrices in one launch.  table (device, int64 x 4 per matrix): {offset of the matrix
 * in `params` (floats), rows, cols, offset of its planes in `planes` (uint16 elements)}; per matrix the planes are
 * [hi (rows, ldk) | lo (rows, ldk) | hi^T (cols, ldn) | lo^T (cols, ldn)] with ldk = cols and ldn = rows rounded up to 8
 * (pad columns are never read: the GEMM's tensor maps end at K); grappa_split_bf16_elems gives that size.
 * max_tiles = max over matrices of ceil(rows / 32) * ceil(cols / 32). */
int64_t grappa_split_bf16_elems(int64_t rows, int64_t cols);
int grappa_split_bf16(int64_t n_mats, const int64_t* table, int64_t max_tiles, const float* params, uint16_t* planes, void* stream);

/* Backward GEMM with the activation derivative fused into its epilogue: C = (A op B) * ELU'(pre-activation), the
 * derivative read off the saved ELU OUTPUT s (ELU'(x) = 1 for x > 0, else exp(x) = s + 1).  Used for dL/d(pre) of the
 * first layer of every Linear-ELU-Linear block (network_utils.py:44-54, graph_attention.py:304-317), where autograd in
 * the reference runs a separate elementwise ELU backward over the (rows, hidden) tensor.  Layouts, math_mode and
 * workspace as grappa_gemm; elu_out: (M, N) with leading dimension ld_eo. */
int grappa_gemm_actgrad(int layout, int64_t M, int64_t N, int64_t K,
                        const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                        const float* elu_out, int64_t ld_eo, int math_mode,
                        float* workspace, int64_t workspace_floats, void* stream);

/* Superset of the two entry points above.  elu_out / ld_eo as in grappa_gemm_actgrad (NULL: none).
 * B_lo (NULL: none): pre-split low plane of the B operand, same shape and ldb as B, as written by grappa_split_lo.  For
 * layouts 0 and 1 on the tensor-core path the kernel then TMA-loads B's lo tile instead of recomputing it in every CTA
 * and k-block (B is a weight matrix there: split once per optimiser step); results are bit-identical.  Ignored by the
 * FFMA path and for layout 2. */
int grappa_gemm_ex(int layout, int64_t M, int64_t N, int64_t K,
                   const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                   const float* bias, int act, const float* res1, int64_t ldr1, const float* res2, int64_t ldr2,
                   float drop_p, uint64_t seed, const uint64_t* seed_dev, int math_mode,
                   float* workspace, int64_t workspace_floats,
                   const float* elu_out, int64_t ld_eo, const float* B_lo, void* stream);

/* lo[i] = rn_tf32(x[i] - trunc19(x[i])): the low plane of the 3xTF32 operand split (x ~ trunc19(x) + lo), n % 4 == 0,
 * 16-byte aligned.  No counterpart in the reference. */
int grappa_split_lo(int64_t n, const float* x, float* lo, void* stream);

/* dZ = dY * keepmask/(1-p) * act'(.) where the activation output is recovered as (Y - res1 - res2)*(1-p).
 * Backward of the epilogue of grappa_gemm (autograd of the call sites above).  res1/res2 may be NULL.
 * colsum_out (N, may be NULL): also returns sum_m dZ[m,:] (the bias gradient) from the same pass; needs scratch of
 * grappa_colsum_scratch_floats(M,N) floats. */
int grappa_epilogue_bwd(int64_t M, int64_t N, const float* dY, int64_t lddy, const float* Y, int64_t ldy,
                        const float* res1, int64_t ldr1, const float* res2, int64_t ldr2, int act,
                        float drop_p, uint64_t seed, const uint64_t* seed_dev, float* dZ, int64_t lddz,
                        float* colsum_out, float* scratch, void* stream);

/* out[n] = sum_m X[m,n]  (bias gradients).  scratch: at least grappa_colsum_scratch_floats(M,N) floats. */
int64_t grappa_colsum_scratch_floats(int64_t M, int64_t N);
int grappa_colsum(int64_t M, int64_t N, const float* X, int64_t ldx, float* out, float* scratch, void* stream);

/* Column sum without any allocation on either side of the call (no counterpart in the reference: the reference never
 * manages scratch itself, PyTorch's caching allocator does).
 * Same result as grappa_colsum, but the partial-sum scratch is a 256 KiB array built into the library image instead
 * of a caller buffer.  Meant for callers that were handed device pointers by somebody else and cannot obtain device
 * memory themselves: a force-field plugin running inside an MD engine's time-step loop (the engine owns a fixed
 * memory budget and cudaMalloc would stall every stream of the device), or an FFI host (ctypes, cgo, JNI) that has
 * no device allocator at hand.  The library calls no cudaMalloc / cudaFree / allocating routine here.
 *   - limit: the partial sums must fit 65536 floats.  For a 16-byte aligned X with ldx % 4 == 0 the need is
 *     min(ceil(296 / ceil(N/128)), ceil(M/32)) * N floats, i.e. below 40 k for every layer width of the default model
 *     at any batch size; for unaligned input it is ceil(M/256) * N.  Larger problems are rejected with an error
 *     message; use grappa_colsum with your own scratch then.
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress, with that device
 *     current); make that first call before starting any stream capture on the device.
 *   - all calls on a device share the one array and leave their partial sums in it: two calls that overlap on the
 *     GPU (different streams, or concurrent host threads) race.  Calls on one stream are safe.  Nothing needs to be
 *     zeroed between calls. */
int grappa_colsum_noalloc(int64_t M, int64_t N, const float* X, int64_t ldx, float* out, void* stream);

/* LayerNorm over the last dimension (eps 1e-5).  Replaces torch.nn.LayerNorm at graph_attention.py:293,312,
 * network_utils.py:46,115,129, perm_equiv_transformer.py:80.  mean/rstd: (M) saved for backward. */
int grappa_layernorm_fwd(int64_t M, int64_t D, const float* X, int64_t ldx, const float* gamma, const float* beta,
                         float* Y, int64_t ldy, float* mean, float* rstd, void* stream);
/* dX may be NULL (only dgamma / dbeta: their column reductions on a side stream) or dgamma / dbeta / scratch NULL (only dX). */
int64_t grappa_layernorm_bwd_scratch_floats(int64_t M, int64_t D);
int grappa_layernorm_bwd(int64_t M, int64_t D, const float* dY, int64_t lddy, const float* X, int64_t ldx,
                         const float* gamma, const float* mean, const float* rstd,
                         float* dX, int64_t lddx, float* dgamma, float* dbeta, float* scratch, void* stream);

/* Atom featurisation.  Replaces the torch.cat / one_hot chain at graph_attention.py:156-177,446-464.
 * blocks: up to 8 (ptr, width, ld) feature blocks concatenated in order, followed by the 16-wide charge encoding
 * of `charge` (n_atoms) [normalised value, one-hot over 15 bins of [-2,2]].  A NULL block pointer writes zeros. */
int grappa_featurize(int64_t n_atoms, int n_blocks, const float* const* block_ptr /*HOST array*/,
                     const int32_t* block_width /*HOST*/, const int32_t* block_ld /*HOST*/,
                     const float* charge, float* out, int64_t ldo, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Graph attention message passing.  Replaces dgl.nn.pytorch.conv.DotGatConv.forward after its `fc` Linear
 * (graph_attention.py:262,296; DGL 2.4.0 dotgatconv.py: u_dot_v / sqrt(D), edge_softmax by destination, u_mul_e sum).
 * ft: (N, H*D).  in_ptr (N+1) / in_src: CSR of incoming edges by destination; out_ptr / out_dst: CSR of outgoing
 * edges by source.  lse: (N,H) log-sum-exp saved for backward.  H*D must be a multiple of 128 and D one of 16,32,64.
 */
int grappa_gnn_attn_fwd(int64_t N, int H, int D, const float* ft, const int32_t* in_ptr, const int32_t* in_src,
                        float* out, float* lse, void* stream);
int grappa_gnn_attn_bwd(int64_t N, int H, int D, const float* ft, const float* out, const float* lse,
                        const float* dOut, const int32_t* in_ptr, const int32_t* in_src,
                        const int32_t* out_ptr, const int32_t* out_dst,
                        float* delta /*scratch (N,H)*/, float* dft, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Interaction readout.
 * tuple_gather: X[t, i, 0:F] = A[idx[t,i], 0:F]; X[t, i, F:W] = pe[i]  -> tokens (T*n_seq, W), tuple-major.
 * Replaces RepProjector's advanced-index gather + transpose (interaction_parameters.py:184-189) and the
 * positional-encoding concat (perm_equiv_transformer.py:152-159).  Backward is a pull reduction over the
 * atom -> token incidence CSR (inc_ptr (N+1), inc_tok), no atomics. */
int grappa_tuple_gather_fwd(int64_t T, int n_seq, int F, int W, const float* A, int64_t lda, const int32_t* idx,
                            const float* pe /*(n_seq) or NULL*/, float* X, void* stream);
int grappa_tuple_gather_bwd(int64_t N, int F, int W, const float* dX, const int32_t* inc_ptr, const int32_t* inc_tok,
                            float* dA, int64_t ldda, void* stream);

/* Token de-duplication for the first LayerNorm + QKV projection of a writer's transformer
 * (perm_equiv_transformer.py:152-165 with network_utils.py:119-123): a token's content depends only on (atom,
 * positional-encoding value), so these two layers run once per table row (2 N rows, or N without a PE column) and
 * are gathered per token.  No counterpart in the reference, which projects every token.
 *   pe_table:   U[2a+pe, c] = c < F ? P[a, c] : pe   (U: (2N, W) dense);  bwd: dP[a, c<F] = dU[2a, c] + dU[2a+1, c]
 *   row_gather: X[r, :] = table[idx[r], :];  bwd (pull, fixed order): dTable[j, :] = sum_{q in ptr[j]..ptr[j+1]} dX[tok[q], :] */
int grappa_pe_table_fwd(int64_t N, int F, int W, const float* P, int64_t ldp, float* U, void* stream);
int grappa_pe_table_bwd(int64_t N, int F, int W, const float* dU, float* dP, int64_t lddp, void* stream);
int grappa_row_gather_fwd(int64_t rows, int D, const float* table, int64_t ldt, const int32_t* idx, float* X,
                          int64_t ldx, void* stream);
int grappa_row_gather_bwd(int64_t table_rows, int D, const float* dX, int64_t ldx, const int32_t* ptr,
                          const int32_t* tok, float* dTable, int64_t ldt, void* stream);

/* Self-attention over the n_seq <= 4 tokens of every tuple (nn.MultiheadAttention core, network_utils.py:123).
 * qkv: (T*n_seq, 3*E) = [q | k | v], E = n_heads*head_dim, head_dim 64.  out: (T*n_seq, E).
 * row_idx (may be NULL): token -> row of qkv; then qkv is the table of DISTINCT token contents (one row per (atom, pe),
 * see grappa_pe_table_fwd) and the gather of interaction_parameters.py:184-189 happens inside the kernel: the
 * (T*n_seq, 3*E) gathered tensor is never written.  dqkv is per token in both cases (T*n_seq, 3*E). */
int grappa_tuple_attn_fwd(int64_t T, int n_seq, int n_heads, int head_dim, const float* qkv, const int32_t* row_idx,
                          float* out, void* stream);
int grappa_tuple_attn_bwd(int64_t T, int n_seq, int n_heads, int head_dim, const float* qkv, const int32_t* row_idx,
                          const float* dOut, float* dqkv, void* stream);

/* R = X3 + X * (col < F ? 2 : 1): the two input residuals around the transformer
 * (perm_equiv_transformer.py:165 and :75-77).  Backward: dX = dR * (col < F ? 2 : 1).
 * row_idx (may be NULL): token -> row of X; then X is the table of distinct token contents (width W) and the gather of
 * the transformer input happens inside the kernel. */
int grappa_token_residual(int64_t M, int W, int F, const float* X3, const float* X, const int32_t* row_idx, float* R, void* stream);
int grappa_token_residual_bwd(int64_t M, int W, int F, const float* dR, float* dX, void* stream);

/* Symmetriser input: rows (P*T, n_seq*W) with row p*T+t = concat_i X[t, perm[p][i], :]
 * (perm_equiv_transformer.py:266-279).  perm: HOST (P*n_seq) ints.  Backward sums the permuted views back. */
int grappa_symm_permute_fwd(int64_t T, int n_seq, int W, int P, const int32_t* perm /*HOST*/, const float* X,
                            float* Y, void* stream);
int grappa_symm_permute_bwd(int64_t T, int n_seq, int W, int P, const int32_t* perm /*HOST*/, const float* dY,
                            float* dX, void* stream);

/* Symmetriser input with the first LayerNorm fused in (training path): the permuted copies of every tuple's token row
 * (perm_equiv_transformer.py:266-279) normalised by the first FeedForwardLayer's LayerNorm (network_utils.py:46) in ONE
 * pass.  All permutations of a tuple hold the same values, so mean / rstd are computed once per tuple (T rows instead of
 * P*T) and the un-normalised permuted copy is never written:
 *   Y[p*T + t, i*W + c] = (X[t, perm[p][i]*W + c] - mean_t) * rstd_t * gamma[i*W + c] + beta[i*W + c]
 * X: (T*n_seq, W) tuple-major tokens, gamma / beta: (n_seq*W), Y: (P*T, n_seq*W), mean / rstd: (T) saved for backward.
 * Backward: dX (T*n_seq, W), dgamma / dbeta (n_seq*W); scratch of grappa_symm_ln_bwd_scratch_floats floats; dX or
 * (dgamma, dbeta, scratch) may be NULL. */
int grappa_symm_ln_fwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* X, const float* gamma,
                       const float* beta, float* Y, float* mean, float* rstd, void* stream);
int64_t grappa_symm_ln_bwd_scratch_floats(int64_t T, int n_seq, int W);
int grappa_symm_ln_bwd(int64_t T, int n_seq, int W, int P, const int32_t* perm, const float* dY, const float* X,
                       const float* gamma, const float* mean, const float* rstd, float* dX, float* dgamma, float* dbeta,
                       float* scratch, void* stream);

/* Sum over permutations (perm_equiv_transformer.py:288-292) fused with the range maps
 * (final_layer.py:48-52, 91-97; interaction_parameters.py:276-288, 382-395, 586-611; network_utils.py:147-148).
 * S: (P*T, n_out) symmetriser outputs; prefactor: HOST (P).
 * kind 0 bond : eq = to_positive(c0; eq_mos, eq_std, eq_min)   k = to_positive(c1; k_mos, k_std, k_min)
 * kind 1 angle: eq = eq_max * sigmoid(eq_som * c0)             k = to_positive(c1; ...)
 * kind 2 torsion (gated): k[:, col0 + j] = cutoff( c[j] * sigmoid(c[n_half + j]) * k_std[j] ), other columns 0.
 * kind 3 torsion (ungated): k[:, col0 + j] = cutoff( c[j] * k_std[j] + k_mean[j] ).
 * consts: HOST float[8] = {eq_a, eq_b, eq_c, k_mos, k_std, k_min, cutoff, unused}; bond: eq_a=mean_over_std,
 * eq_b=std, eq_c=min; angle: eq_a=std_over_max, eq_b=max.  kstd / kmean: device (n_half) for kinds 2, 3.
 * C: (T, n_out) pre-map sums saved for backward. */
int grappa_readout_finalize_fwd(int kind, int64_t T, int P, int n_out, const float* S, const float* prefactor,
                                const float* consts, const float* kstd, const float* kmean, int n_per, int col0,
                                float* C, float* eq, float* k, void* stream);
int grappa_readout_finalize_bwd(int kind, int64_t T, int P, int n_out, const float* C, const float* prefactor,
                                const float* consts, const float* kstd, const float* kmean, int n_per, int col0,
                                const float* d_eq, const float* d_k, float* dS, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * MM energy + analytic dE/dxyz.  Replaces internal_coordinates (internal_coordinates.py:19-115),
 * harmonic_energy / torsion_energy / pool_energy (energy.py:8-71) and the autograd.grad call (energy.py:160-164).
 *
 * The host plan partitions the atoms of every molecule into blocks of <= block_atoms consecutive atoms and compiles,
 * per block, a PROGRAM of every tuple touching it, split into n_parts balanced parts (one per warp of the kernel's
 * CTA); tuples spanning blocks are evaluated once per block they touch, only the owner block adds their energy.
 * Record / word formats: csrc/energy.cu.  All program words are relative to the block or its molecule, so per-molecule
 * plans concatenate into a batch plan by offsetting columns 0 / 2 / 3 of the block rows (grappa_collate_int mode 1).
 * grappa_energy_plan_build fills caller-provided HOST int32 arrays; call it first with blk == NULL to obtain
 *   sizes_out[12] = {n_blocks, n_words, max_words_per_block, max_slots_per_part, max_atoms_per_block, any_multi_block,
 *                    n_parts, max tuples of one molecule at level n2, n3, n4, n4_improper, any_non_local_block}.
 * idx2/3/4/4i: HOST int64 (T,n_seq) global atom rows, rows grouped by molecule in molecule order (as dgl.batch
 * produces); mol_atom_off: HOST int64 (n_mol+1).  mol_tuple_off: (4, n_mol+1) first tuple row of every molecule per
 * level; tuple_mol (may be NULL): molecule of every tuple, levels concatenated.
 */
int grappa_energy_plan_build(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                             const int64_t* idx2, int64_t T2, const int64_t* idx3, int64_t T3,
                             const int64_t* idx4, int64_t T4, const int64_t* idx4i, int64_t T4i,
                             int block_atoms, int n_parts, int64_t* sizes_out,
                             int32_t* blk /*(n_blocks,12)*/, int32_t* prog /*(n_words)*/,
                             int32_t* mol_tuple_off /*(4,n_mol+1)*/, int32_t* tuple_mol /*(T2+T3+T4+T4i)*/);

/* blk / prog / mol_tuple_off are the device copies of the arrays above; plan_sizes is the HOST sizes_out[12] of the
 * plan (for a concatenated batch plan: sums of n_blocks / n_words, maxima of the rest).  k4: (T4,P4), k4i: (T4i,P4i).
 * Outputs: energy (n_mol,C) total; term_energy (4,n_mol,C); gradient (N,C,3) = +dE/dxyz.
 * Optional (may be NULL): x2,x3,x4,x4i (T,C) internal coordinates; e2,e3,e4,e4i (T,C) per-tuple energies.
 * partial: scratch (n_blocks,4,C) floats, used only when a molecule spans several blocks.
 * term_mask: bit l set = include level l (0 bond, 1 angle, 2 proper, 3 improper); 15 = all.  Sub-masks give the
 * per-term gradients of Energy(gradient_contributions=True) (energy.py:146-152).
 * Coordinate and gradient rows move through the TMA engine (cp.async.bulk) when C % 4 == 0 and xyz / gradient are
 * 16-byte aligned, through coalesced loads / stores otherwise; results are identical. */
int grappa_energy_fwd(int64_t n_atoms, int64_t n_mol, int64_t C, const float* xyz,
                      int64_t n_blocks, const int32_t* blk, const int32_t* prog, const int32_t* mol_tuple_off,
                      const int64_t* plan_sizes /*HOST (12)*/, int term_mask,
                      const float* k2, const float* eq2, const float* k3, const float* eq3,
                      const float* k4, int P4, const float* k4i, int P4i,
                      float* energy, float* term_energy, float* gradient,
                      float* x2, float* x3, float* x4, float* x4i, float* e2, float* e3, float* e4, float* e4i,
                      float* partial, void* stream);

/* Vector-Jacobian product of (energy, gradient) w.r.t. the MM parameters for the training loss
 * (double backward through energy.py:162 in closed form, SURVEY.md §7 "Double backward").
 * gE: (n_mol,C) dL/d energy (may be NULL), gG: (N,C,3) dL/d gradient (may be NULL).
 * idx*: device int32 (T,n_seq); tuple_mol: device int32 (T2+T3+T4+T4i), molecule of each tuple. */
int grappa_energy_vjp(int64_t C, const float* xyz, const float* gE, const float* gG, const int32_t* tuple_mol,
                      const int32_t* idx2, int64_t T2, const int32_t* idx3, int64_t T3,
                      const int32_t* idx4, int64_t T4, const int32_t* idx4i, int64_t T4i,
                      const float* k2, const float* eq2, const float* k3, const float* eq3,
                      const float* k4, int P4, const float* k4i, int P4i,
                      float* dk2, float* deq2, float* dk3, float* deq3, float* dk4, float* dk4i, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Host helpers (HOST pointers only).
 * Stable counting sort of n items by key in [0,n_keys): ptr (n_keys+1), order (n) = item ids grouped by key in
 * their original order.  Used for CSR-by-destination / by-source of the bond edges and the atom -> token incidence. */
int grappa_csr_build(int64_t n, int64_t n_keys, const int64_t* keys, int32_t* ptr, int32_t* order);

/* The whole host batch plan behind one call (HOST pointers only): idxs narrowed to int32, atom -> token incidence CSRs,
 * (atom, positional-encoding) token tables, bond-edge CSRs by destination / source, optionally the energy plan, and the
 * per-molecule offsets, as sections of ONE packed int32 buffer that reaches the device in a single copy.  Replaces what
 * the reference re-derives on every call by advanced indexing and DGL segment bookkeeping
 * (models/interaction_parameters.py:184, models/internal_coordinates.py:31-94, models/energy.py:69) and performs the
 * idxs bounds check of models/grappa.py:179-185 (return code 4, message in that wording).
 *   idx[l] (T[l], n_seq) int64 for l = n2, n3, n4, n4_improper; level_counts[l] (n_mol) tuples per molecule or NULL (level
 *   absent); pe_pattern[l] (n_seq) positional-encoding value per tuple position or NULL (no token table for that level).
 * Outputs: *n_int32_out = buffer length; sections_out[2 * i] / [2 * i + 1] = offset / length (int32 elements) of section
 * i in the order listed in csrc/plan.cu (37 sections; absent ones have length 0); energy_sizes_out[12] as
 * grappa_energy_plan_build (only with_energy); *handle_out owns the packed buffer until grappa_plan_fill copies it into
 * caller memory (pinned, typically) and grappa_plan_free releases it. */
int grappa_plan_build(int64_t n_atoms, int64_t n_mol, const int64_t* mol_atom_off,
                      const int64_t* const* idx, const int64_t* T, const int64_t* const* level_counts,
                      const int32_t* const* pe_pattern, int64_t n_edges, const int64_t* edge_src, const int64_t* edge_dst,
                      int with_energy, int block_atoms, int n_parts,
                      int64_t* n_int32_out, int64_t* sections_out, int64_t* energy_sizes_out, void** handle_out);
int grappa_plan_fill(void* handle, int32_t* out, int64_t n_int32);
void grappa_plan_free(void* handle);

/* Interaction-tuple enumeration from the bond list; replaces utils/tuple_indices.py:8-70 (get_idx_tuples) and
 * :73-110 (get_neighbor_dict).  bonds: (n_bonds,2) int64 atom rows in [0,n_atoms).  Rows come out in the reference's
 * order: outer loop over atoms in first-appearance order of the flattened bond list, sorted neighbour lists, angle
 * (a1,a2,a3) kept when a1 < a3, proper (a4,a3,a2,a1) for a4 < a1, a4 != a2.  Two calls: with angles == propers ==
 * NULL only counts[0] (angles) and counts[1] (propers) are written; with both given, counts holds their capacities
 * (rows) on entry and the row counts on return.  Self-bonds are an error (reference: assertion at :85). */
int grappa_enumerate_tuples(int64_t n_bonds, const int64_t* bonds, int64_t n_atoms, int64_t* counts,
                            int64_t* angles, int64_t* propers);

/* ------------------------------------------------------------------------------------------------------------
 * Device-side collate.  Replaces, per training step, the host collate of
 * data/graph_data_loader.py:23-73 (conformation sub-/over-sampling, is_dummy padding: utils/dgl_utils.py:137-176) and
 * utils/dgl_utils.py:11-63 (idxs offsetting + concatenation) by two gather launches over a packed per-molecule store
 * that is resident in HBM.  seg: device (n_seg,8) int64 segment rows, one CTA each:
 *   int rows   [src, dst, len, mode, a0, a1, a2, a3]   mode 0 copy+a0 | 1 energy-block rows (col0+a0, col2+a1, col3+a2)
 *              | 4 fill a0 | 5 copy from lit[src..]   (modes 2, 3 retired with the record tables of ABI version 1)
 *   float rows [src, dst, rows, inner, C_src, C_dst, conf_off, mode]   mode 0 gather conformations conf[conf_off+c]
 *              (conf_off < 0: identity) | 1 is_dummy flags (c >= C_src).
 * store / out: device buffers (src / dst are element offsets into them); lit, conf: device int32. */
int grappa_collate_int(int64_t n_seg, const int64_t* seg, const int32_t* store, const int32_t* lit, int32_t* out,
                       void* stream);
int grappa_collate_float(int64_t n_seg, const int64_t* seg, const int32_t* conf, const float* store, float* out,
                         void* stream);

/* Training-side fused kernels (next rows of SURVEY.md §8f, used by the data-parallel step driver). */
/* Molecule-wise loss (training/loss.py:51-174 with param_weight 0): forward value and gradients w.r.t. energy,
 * gradient and the torsion force constants in one pass.  loss_parts: (n_mol) per-molecule terms (already / n_mol);
 * the scalar loss is their sum in molecule order (written to loss_out). */
int grappa_molwise_loss(int64_t n_mol, int64_t C, const int32_t* mol_atom_off, const int32_t* mol_p_off,
                        const int32_t* mol_i_off, const float* energy, const float* energy_ref,
                        const float* gradient, const float* gradient_ref, const float* is_dummy,
                        const float* k4, int P4, const float* k4i, int P4i,
                        float energy_weight, float gradient_weight, float proper_reg, float improper_reg,
                        const float* upstream /*device scalar or NULL: scales dE,dG,dk4,dk4i*/,
                        float* loss_parts, float* loss_out, float* dE, float* dG, float* dk4, float* dk4i,
                        void* stream);

/* AdamW over one flat parameter buffer (torch.optim.AdamW semantics, lightning_model.py:368-370).
 * grad_scale multiplies the gradient first (1/world_size after an allreduce-sum). */
int grappa_adamw(int64_t n, float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                 float lr, float beta1, float beta2, float eps, float weight_decay, int64_t step, float grad_scale,
                 void* stream);

/* Data-parallel training on one NVSwitch domain: gradient exchange FUSED with AdamW, one launch per parameter stage
 * (replaces the NCCL all-reduce of Lightning's DDP strategy followed by torch.optim.AdamW on every rank,
 * training/lightning_model.py:368-370; csrc/dp_exchange.cu).  Flat element range [lo, hi) (multiples of 4) is cut in
 * `world` shards (grappa_dp_shard); this rank reads ITS shard's gradients summed over all ranks by the switch
 * (multimem.ld_reduce on grads_mc), updates the shard with its own exp_avg / exp_avg_sq (the optimiser state is sharded:
 * only the owner's part of those buffers is meaningful), and stores the new parameters to every rank through
 * params_mc (multimem.st).  grads_mc / params_mc: MULTICAST addresses of the flat gradient / parameter buffers
 * (element 0); params / exp_avg / exp_avg_sq: this rank's local buffers (element 0).  peer_signal_pads: HOST array of
 * `world` device addresses, entry r = rank r's signal pad (pad_words zero-initialised uint32, in peer-mapped memory);
 * the launch uses the words [slot_base, slot_base + n_ctas*world) of every pad, which launches that may overlap in
 * time must not share.  Every rank must make the same call (same range, n_ctas, slot_base) in the same order.
 * A peer that does not arrive within spin_seconds (0 = wait forever) makes the kernel trap.  step >= 1 as in
 * grappa_adamw; grad_scale = 1/world for a mean over ranks. */
int grappa_dp_shard(int64_t lo, int64_t hi, int rank, int world, int64_t* first_out, int64_t* count_out);
int grappa_dp_exchange_adamw(int64_t lo, int64_t hi, int rank, int world, const float* grads_mc, float* params_mc,
                             float* params, float* exp_avg, float* exp_avg_sq, const uint64_t* peer_signal_pads,
                             int64_t pad_words, int64_t slot_base, int n_ctas, float lr, float beta1, float beta2,
                             float eps, float weight_decay, int64_t step, float grad_scale, float spin_seconds,
                             void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRAPPA_B200_H */
