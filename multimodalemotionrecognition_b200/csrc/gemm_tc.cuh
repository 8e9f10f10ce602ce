// tcgen05/TMA projection GEMM (gemm_tc.cu).
#pragma once
#include "common.cuh"

namespace xf {
// Y[M,N] = X[M,K]·W[N,K]^T + bias.  X, W: bf16, K contiguous, 16-byte aligned.  out: bf16 or fp32 with row pitch
// `out_pitch` (elements).  N % 128 == 0, K % 64 == 0.
int gemm_tc_bf16(cudaStream_t st, const void* X, const void* W, const float* bias, void* out, bool out_bf16, int64_t out_pitch,
                 int M, int N, int K);
int make_tmap_bf16_2d(void* tmap_out, const void* base, int64_t rows, int64_t cols, int box_rows, int box_cols);
}  // namespace xf
