// Internal (non-ABI) declarations shared by the T5 scorer translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

enum { PN_EPI_BF16 = 0, PN_EPI_RESID_F32 = 1, PN_EPI_F32 = 2, PN_EPI_ATOMIC_F32 = 3 };

// C = A[M,K] . B[N,K]^T on the tcgen05 path (pn_t5_gemm.cu).  mode PN_EPI_BF16: bf16 out (optional
// ReLU, optional per-row scale); PN_EPI_RESID_F32: fp32 out += acc.
int pn_gemm_bf16_internal(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M,
                          int N, int K, int mode, int relu, const float* row_scale, int cta_group,
                          cudaStream_t st);

#include <cuda_bf16.h>
#include "pn_dropout.cuh"
// optional epilogue extras of the tcgen05 GEMM (all nullable)
struct PnGemmExtra {
  const float* row_scale;     // acc *= row_scale[row]
  const float* sumsq_in;      // fused RMSNorm of A: acc *= rsqrt(sum_p sumsq_in[p*M+row] / norm_dim + eps)
  int sumsq_parts;
  int norm_dim;
  float eps;
  __nv_bfloat16* copy_out;    // fp32 modes: bf16 copy of the stored values, [M, ld_copy]
  int ld_copy;
  float* sumsq_out;           // fp32 modes: [pn_gemm_sumsq_parts(N)][M] row sums of squares per n-tile
  const __nv_bfloat16* relu_mask;   // bf16 mode: out = relu_mask[row, col] > 0 ? acc : 0 (ReLU backward), [M, ld_mask]
  int ld_mask;
  int group_rows, group_b_rows;     // block-diagonal B: A rows [g*group_rows, ..) use the B rows from g*group_b_rows
  // training-mode dropout on the produced values (thresh 0 = off): element (row, col) has index row * N + col.
  //   bf16 mode: out = dropout(relu?(acc * scales))           (FFN hidden, HF:103)
  //   residual mode: out += dropout(acc * scales)              (sub-layer output before the residual add)
  PnDrop drop;
  float out_scale;                  // extra scalar factor on acc (0 = 1): 1/(1-p) of a dropped ReLU-backward mask
  // bf16 mode with relu_mask, nullable: [N / 256][M] per-n-tile row sums of out[row, col] * relu_mask[row, col]
  // (the saved activation doubles as the forward output: the row dot of the RMSNorm backward, see nb_coef)
  float* dot_out;
  // residual mode, nullable: RMSNorm backward folded into the data-gradient GEMM of a normed linear,
  //   out[row, col] += acc - nb_coef[row] * nb_hb[row, col]   (+ the bf16 copy of the updated stream in copy_out;
  //   `drop` then masks the COPY only: the dropout of the sub-layer output below, whose gradient the copy is)
  // nb_coef[row] = rs^2 * <G, h> / d, known BEFORE the product because <G, h> = <Ys, y> / rs with y the forward output
  const float* nb_coef;
  const __nv_bfloat16* nb_hb;       // [M, ld_hb] the bf16 residual stream the norm read
  int ld_hb;
};
int pn_gemm_bf16_ex(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M, int N, int K,
                    int mode, int relu, const PnGemmExtra* ex, int cta_group, cudaStream_t st);
int pn_gemm_sumsq_parts(int N);
// weight gradient on the tcgen05 path: C[M,N] (fp32, ldc) += A[K,M]^T . B[K,N], A and B bf16 row-major with K rows
// (dW = dY^T X); the K range is split across work items and accumulated with fp32 vector atomics
int pn_gemm_wgrad(const void* A, int lda, const void* B, int ldb, float* out, int ldc, int M, int N, int K,
                  int cta_group, cudaStream_t st);

// batched C_b[M,N] (fp32, accumulated into zeroed memory) = A_b[Kper,M]^T . B_b[Kper,N], batch b = rows
// [b*Kper, (b+1)*Kper) of A and B, output block out + b*M*ldc
int pn_gemm_tn_batched(const void* A, int lda, const void* B, int ldb, float* out, int ldc, int M, int N, int Kper,
                       int batches, cudaStream_t st);

int pn_make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols);

#include <cuda_bf16.h>
// T5 encoder self-attention on tcgen05 (pn_t5_attn.cu).  qkv rows are doc*S + position.
int pn_t5_self_attn(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                    int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out, float* lse_out,
                    cudaStream_t st);
// training forward: dropout of the attention probabilities (HF:331), masks indexed [doc, head, i, j] with an even
// row pitch; the logsumexp it emits is that of the un-dropped softmax
int pn_t5_self_attn_drop(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                         int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out, float* lse_out,
                         PnDrop drop, cudaStream_t st);
// Same with a key-padding mask: keys at positions >= klen[doc] (nullable, [docs]) are excluded from every row's
// softmax — HF's attention_mask for right-padded prompts (the re-rank path, eval.py:14,23 / distill/wrapper.py:36).
int pn_t5_self_attn_masked(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                           int S, int n_heads, const float* bias, const int32_t* klen, __nv_bfloat16* out,
                           int ld_out, float* lse_out, cudaStream_t st);
// Backward of pn_t5_self_attn (pn_t5_attn_bwd.cu): dqkv <- (dq | dk | dv) (each row optionally scaled by
// row_scale), dbias_diag [H, 1023] += per-diagonal sums of dS; lse2 is the forward's lse_out, dvec a [rows, H]
// workspace.
// pn_t5_rel_bias_grad folds the diagonals into the gradient of relative_attention_bias.weight.
int pn_t5_self_attn_bwd(const __nv_bfloat16* qkv, int ld, int q_col, int k_col, int v_col, int docs, int S,
                        int n_heads, const float* bias, const __nv_bfloat16* ctx, const __nv_bfloat16* dctx,
                        int ld_ctx, const float* lse2, float* dvec, const float* row_scale,
                        __nv_bfloat16* dqkv, float* dbias_diag, cudaStream_t st, PnDrop drop = PnDrop(),
                        float* dot_out = nullptr);
// dot_out (nullable): [4 * n_heads][docs * S] partial row sums of dqkv[row, :] * qkv[row, :] (plane = (pass * n_heads +
// head) * 2 + column half), the row dot the RMSNorm backward of the qkv projection needs (pn_t5_norm_coef)
// coef[row] = scale * rs[row] * sum_p part[p * rows + row] / d
int pn_t5_norm_coef(const float* part, int parts, const float* rs, int64_t rows, int d, float* coef, cudaStream_t st,
                    float scale = 1.0f);   // scale: extra factor on the sum (1 - p when the saved activation was dropped)
int pn_t5_rel_bias_grad(const float* dbias_diag, const int* bucket_of, int n_heads, int buckets, float* grad,
                        cudaStream_t st);
// decoder self / cross attention for T <= 2 query tokens per document
int pn_t5_dec_attn(const __nv_bfloat16* q, int ld_q, int q_col, const __nv_bfloat16* k,
                   const __nv_bfloat16* v, int ld_kv, int k_col, int v_col, int docs, int T, int Sk,
                   int n_heads, int causal, const float* bias, __nv_bfloat16* out, int ld_out,
                   cudaStream_t st, const int32_t* klen = nullptr,   // klen [docs]: keys >= klen[doc] are padding
                   PnDrop drop = PnDrop());                        // dropout of the probabilities (training)

// Backward of pn_t5_dec_attn (pn_t5_dec_attn_bwd.cu).  dq rows are scaled by rs_q, dk / dv rows by rs_kv (both
// nullable); grad_rel (nullable) accumulates the decoder relative_attention_bias gradient [buckets, H].
int pn_t5_dec_attn_bwd(const __nv_bfloat16* q, int ld_q, int q_col, const __nv_bfloat16* k,
                       const __nv_bfloat16* v, int ld_kv, int k_col, int v_col, const __nv_bfloat16* dout,
                       int ld_do, int docs, int T, int Sk, int n_heads, int causal, const float* bias,
                       const float* rs_q, const float* rs_kv, __nv_bfloat16* dq, int ld_dq, int dq_col,
                       __nv_bfloat16* dkv, int ld_dkv, int dk_col, int dv_col, float* grad_rel,
                       cudaStream_t st, PnDrop drop = PnDrop());
