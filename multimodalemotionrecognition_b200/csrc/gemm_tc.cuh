// tcgen05/TMA projection GEMM (gemm_tc.cu).
#pragma once
#include "common.cuh"

namespace xf {
// Y[M,N] = X[M,K]·W[N,K]^T + bias.  X, W: bf16, K contiguous, 16-byte aligned.  out: bf16 or fp32 with row pitch
// `out_pitch` (elements).  N % 128 == 0, K % 64 == 0.  relu: Y = max(Y, 0) (plain epilogues only).
// Optional extras (sc_w != nullptr; bf16 output with N == 128 only), computed in the epilogue from the tile it holds:
//   tok[i][clip][j] (two planes, kTokPitch floats per clip) = sc_w[i,:]·y_m for row m = clip*seg_len + j, from the fp32 row
//   colsum[clip][slot][128] = sum of the rows of `clip` (rows m in [clip*seg_len, +seg_len)) inside the 32-row group
//                             (clip*seg_len)/32 + slot of the output; a clip overlaps at most kColsumSlots groups.
constexpr int kColsumSlots = 6;
constexpr int kTokPitch = 160;   // floats per clip in each tok plane
int gemm_tc_bf16(cudaStream_t st, const void* X, const void* W, const float* bias, void* out, bool out_bf16, int64_t out_pitch,
                 int M, int N, int K, const float* sc_w = nullptr, float* tok = nullptr, float* colsum = nullptr, int seg_len = 0,
                 bool relu = false);
// Video-side update as two GEMMs over weights folded at pack time (bf16_pack.cu), rows = (clip, token slot), 16 slots per clip:
//   video_mix: v1 = LayerNorm(v0 + U[16·clips, 512]·G[128, 512]^T + bias) as bf16 [16·clips, 128]  (fusion.py:394-395);
//              pooled[clip][0..128) = mean over the T real slots (:401); cv[clip][16h + i] = v1_i·wc[h] + wc[512 + h]
//   video_kv : [Kt | Vt][clip][16h + i][:] = v1[16·clip + i]·F[128·(4·which + h) .., 128]^T + bias   (F: [1024, 128])
int gemm_tc_video_mix(cudaStream_t st, const void* U, const void* G, const float* bias, const float* v0, const float* gamma,
                      const float* beta, const float* wc, void* v1, float* cv, float* pooled, int clips, int T);
int gemm_tc_video_kv(cudaStream_t st, const void* v1, const void* F, const float* bias, void* kt, void* vt, int clips);
// 2D bf16 tensor map with SWIZZLE_128B boxes: `cols` contiguous elements per row, rows `pitch` elements apart (0 = cols).
int make_tmap_bf16_2d(void* tmap_out, const void* base, int64_t rows, int64_t cols, int box_rows, int box_cols, int64_t pitch = 0);
int make_tmap_bf16_3d(void* tmap_out, const void* base, int64_t d0, int64_t d1, int64_t d2, int b0, int b1, int b2);

// ---- tensor-core GEMMs of the training path (gemm_tf32.cu), precision = XATTN_PREC_TF32 ----
// C[M,N] (+)= A[M,K]·W^T (+bias)(ReLU) with W given as [N,K] (w_is_kn = false) or as [K,N] (true: dX = dY·W, no transposed copy);
// fp32 tensors, tf32 products, fp32 accumulation.  N % 128 == 0, K % 32 == 0.
bool gemm_tf32_rows_ok(int M, int N, int K, int64_t lda, int64_t ldc);
int gemm_tf32_rows(cudaStream_t st, const float* A, int64_t lda, const float* W, bool w_is_kn, const float* bias, float* C, int64_t ldc,
                   int M, int N, int K, bool relu, bool accumulate);
// C[N1,N2] += A[M,N1]^T·B[M,N2] (dW = dY^T·X), A and B both fp32 (tf32 products) or both bf16; N1, N2 % 128 == 0.
bool gemm_cols_ok(int M, int N1, int N2, int64_t lda, int64_t ldb, bool bf16);
int gemm_cols(cudaStream_t st, const void* A, int64_t lda, const void* B, int64_t ldb, bool bf16, float* C, int64_t ldc, int M, int N1,
              int N2);
}  // namespace xf
