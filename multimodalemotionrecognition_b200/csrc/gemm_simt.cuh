// Generic strided fp32 SIMT GEMM used by the fp32 parity/training path:
//   C[m,n] (op)= sum_k A(m,k) * B(k,n) (+ bias[n]) (ReLU)
// A(m,k) = A[m*am + k*ak], B(k,n) = B[k*bk + n*bn].  One kernel covers X·W^T (forward), dY·W (input
// gradients) and dY^T·X (weight gradients, split over the long reduction with atomics).
// fp32 accumulate, FFMA only — this path exists for the 1e-4 fp32 parity gate; the bf16 tcgen05 kernels
// in fwd_bf16.cu are the performance path.
#pragma once
#include <type_traits>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace xf {

enum GemmFlags : int { GEMM_RELU = 1, GEMM_ACCUM = 2, GEMM_ATOMIC = 4 };

constexpr int GT_M = 64, GT_N = 64, GT_K = 16, GT_PAD = 4;

template <typename TA, typename TB>
__global__ void __launch_bounds__(256) gemm_generic_kernel(const TA* __restrict__ A, int64_t am, int64_t ak,
                                                           const TB* __restrict__ B, int64_t bk, int64_t bn,
                                                           float* __restrict__ C, int64_t ldc, int M, int N, int K,
                                                           const float* __restrict__ bias, int flags, int k_chunk) {
  __shared__ __align__(16) float As[GT_K][GT_M + GT_PAD];
  __shared__ __align__(16) float Bs[GT_K][GT_N + GT_PAD];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * GT_M, n0 = blockIdx.x * GT_N;
  const int kbeg = blockIdx.z * k_chunk;
  const int kend = min(K, kbeg + k_chunk);
  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const bool a_mfast = (am == 1);
  const bool b_nfast = (bn == 1);
  for (int kt = kbeg; kt < kend; kt += GT_K) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int m, k;
      if (a_mfast) { m = tid & 63; k = (tid >> 6) + 4 * i; } else { k = tid & 15; m = (tid >> 4) + 16 * i; }
      float v = 0.f;
      if (m0 + m < M && kt + k < kend) v = to_f32(A[(int64_t)(m0 + m) * am + (int64_t)(kt + k) * ak]);
      As[k][m] = v;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int n, k;
      if (b_nfast) { n = tid & 63; k = (tid >> 6) + 4 * i; } else { k = tid & 15; n = (tid >> 4) + 16 * i; }
      float v = 0.f;
      if (n0 + n < N && kt + k < kend) v = to_f32(B[(int64_t)(kt + k) * bk + (int64_t)(n0 + n) * bn]);
      Bs[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GT_K; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      float v = acc[i][j];
      if (bias != nullptr && blockIdx.z == 0) v += bias[n];
      float* dst = C + (int64_t)m * ldc + n;
      if (flags & GEMM_ATOMIC) {
        atomicAdd(dst, v);
      } else {
        if (flags & GEMM_ACCUM) v += *dst;
        if (flags & GEMM_RELU) v = fmaxf(v, 0.f);
        *dst = v;
      }
    }
  }
}

// Host launcher.  splits > 1 => GEMM_ATOMIC into a pre-zeroed (or to-be-accumulated) C.
template <typename TA, typename TB>
inline void launch_gemm(cudaStream_t st, const TA* A, int64_t am, int64_t ak, const TB* B, int64_t bk, int64_t bn, float* C,
                        int64_t ldc, int M, int N, int K, const float* bias, int flags, int splits = 1) {
  if (M <= 0 || N <= 0 || K <= 0) return;
  int k_chunk = K;
  if (splits > 1) {
    k_chunk = ((K + splits - 1) / splits + GT_K - 1) / GT_K * GT_K;
    splits = (K + k_chunk - 1) / k_chunk;
    flags |= GEMM_ATOMIC;
  }
  dim3 grid((N + GT_N - 1) / GT_N, (M + GT_M - 1) / GT_M, splits);
  gemm_generic_kernel<TA, TB><<<grid, 256, 0, st>>>(A, am, ak, B, bk, bn, C, ldc, M, N, K, bias, flags, k_chunk);
  count_launch();
}

// Y[M,N] = X[M,K] · W[N,K]^T + b      (nn.Linear forward)
template <typename TX>
inline void linear_fwd(cudaStream_t st, const TX* X, const float* W, const float* b, float* Y, int M, int N, int K,
                       bool relu = false) {
  if (std::is_same<TX, float>::value && g_tc_gemm && gemm_tf32_rows_ok(M, N, K, K, N)) {
    const int rc = gemm_tf32_rows(st, reinterpret_cast<const float*>(X), K, W, false, b, Y, N, M, N, K, relu, false);
    if (rc != XATTN_OK) g_tc_status = rc;
    return;
  }
  launch_gemm<TX, float>(st, X, K, 1, W, 1, K, Y, N, M, N, K, b, relu ? GEMM_RELU : 0);
}
// dX[M,K] (+)= dY[M,N] · W[N,K]
inline void linear_bwd_dx(cudaStream_t st, const float* dY, int64_t ldy, const float* W, float* dX, int M, int N, int K,
                          bool accumulate) {
  if (g_tc_gemm && gemm_tf32_rows_ok(M, K, N, ldy, K)) {   // reduction over N, K output columns, W[N,K] read as [reduction, out]
    const int rc = gemm_tf32_rows(st, dY, ldy, W, true, nullptr, dX, K, M, K, N, false, accumulate);
    if (rc != XATTN_OK) g_tc_status = rc;
    return;
  }
  launch_gemm<float, float>(st, dY, ldy, 1, W, K, 1, dX, K, M, K, N, nullptr, accumulate ? GEMM_ACCUM : 0);
}
// dW[N,K] += dY[M,N]^T · X[M,K]   (dW must be zeroed or hold a partial sum; long reduction over M is split)
template <typename TX>
inline void linear_bwd_dw(cudaStream_t st, const float* dY, int64_t ldy, const TX* X, int64_t ldx, float* dW, int M, int N,
                          int K) {
  if (std::is_same<TX, float>::value && g_tc_gemm && gemm_cols_ok(M, N, K, ldy, ldx, false)) {
    const int rc = gemm_cols(st, dY, ldy, X, ldx, false, dW, K, M, N, K);
    if (rc != XATTN_OK) g_tc_status = rc;
    return;
  }
  int splits = M / 128;   // (short reductions too: one CTA walking 1024 rows alone is a 180 us kernel)
  if (M <= 4096) splits = M / 32;   // batch-sized reductions (head / prior weights): each 16-row step is a dependent global round trip
  if (splits < 1) splits = 1;
  if (splits > 512) splits = 512;
  if (splits == 1) {
    launch_gemm<float, TX>(st, dY, 1, ldy, X, ldx, 1, dW, K, N, K, M, nullptr, GEMM_ACCUM, 1);
  } else {
    launch_gemm<float, TX>(st, dY, 1, ldy, X, ldx, 1, dW, K, N, K, M, nullptr, 0, splits);
  }
}

// db[N] += column sums of dY[M,N]
__global__ void colsum_kernel(const float* __restrict__ dY, int64_t ldy, int M, int N, float* __restrict__ db, int rows_per_block);
void launch_colsum(cudaStream_t st, const float* dY, int64_t ldy, int M, int N, float* db);

}  // namespace xf
