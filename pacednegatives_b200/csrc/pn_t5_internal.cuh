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
};
int pn_gemm_bf16_ex(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M, int N, int K,
                    int mode, int relu, const PnGemmExtra* ex, int cta_group, cudaStream_t st);
int pn_gemm_sumsq_parts(int N);
// weight gradient on the tcgen05 path: C[M,N] (fp32, ldc) += A[K,M]^T . B[K,N], A and B bf16 row-major with K rows
// (dW = dY^T X); the K range is split across work items and accumulated with fp32 vector atomics
int pn_gemm_wgrad(const void* A, int lda, const void* B, int ldb, float* out, int ldc, int M, int N, int K,
                  int cta_group, cudaStream_t st);

int pn_make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols);

#include <cuda_bf16.h>
// T5 encoder self-attention on tcgen05 (pn_t5_attn.cu).  qkv rows are doc*S + position.
int pn_t5_self_attn(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                    int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out,
                    cudaStream_t st);
// decoder self / cross attention for T <= 2 query tokens per document
int pn_t5_dec_attn(const __nv_bfloat16* q, int ld_q, int q_col, const __nv_bfloat16* k,
                   const __nv_bfloat16* v, int ld_kv, int k_col, int v_col, int docs, int T, int Sk,
                   int n_heads, int causal, const float* bias, __nv_bfloat16* out, int ld_out,
                   cudaStream_t st);
