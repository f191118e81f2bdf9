// Internal (non-ABI) declarations shared by the T5 scorer translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

enum { PN_EPI_BF16 = 0, PN_EPI_RESID_F32 = 1 };

// C = A[M,K] . B[N,K]^T on the tcgen05 path (pn_t5_gemm.cu).  mode PN_EPI_BF16: bf16 out (optional
// ReLU, optional per-row scale); PN_EPI_RESID_F32: fp32 out += acc.
int pn_gemm_bf16_internal(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M,
                          int N, int K, int mode, int relu, const float* row_scale, int cta_group,
                          cudaStream_t st);

int pn_make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols);
