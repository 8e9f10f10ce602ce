// Test-only entry points (include/xattn_fusion_debug.h): single kernels of the product in isolation and the tcgen05 hardware
// probe.  Built into libxattn_fusion_debug.so only — the product library libxattn_fusion.so does not carry them.
#include "../../include/xattn_fusion_debug.h"
#include "gemm_tc.cuh"
#include "kernels_bf16.cuh"
#include "kernels_f32.cuh"

namespace xf {
int umma_probe(cudaStream_t st, const void* A, const void* B, float* out32, float* out16, int N, int K, int mode);
}
using namespace xf;

static inline bool misaligned(const void* p) { return ((uintptr_t)p & 15) != 0; }

extern "C" XATTN_API int xattn_debug_gemm_tc(const void* X, const void* W, const float* bias, void* Y, int32_t out_bf16, int64_t out_pitch,
                                   int32_t M, int32_t N, int32_t K, void* stream) {
  if (!X || !W || !bias || !Y) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(X) || misaligned(W) || misaligned(Y)) XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  return gemm_tc_bf16((cudaStream_t)stream, X, W, bias, Y, out_bf16 != 0, out_pitch, M, N, K);
}

extern "C" XATTN_API int xattn_debug_gemm_tf32(int32_t form, const void* A, const void* B, const float* bias, float* C, int32_t M, int32_t N,
                                               int32_t K, int32_t relu, int32_t accumulate, void* stream) {
  if (!A || !B || !C) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  XF_TRY(check_device_sm100());
  cudaStream_t st = (cudaStream_t)stream;
  if (form == 0 || form == 1)
    return gemm_tf32_rows(st, (const float*)A, K, (const float*)B, form == 1, bias, C, N, M, N, K, relu != 0, accumulate != 0);
  if (form == 2 || form == 3) return gemm_cols(st, A, N, B, K, form == 3, C, K, M, N, K);
  XF_FAIL(XATTN_ERR_SHAPE, "xattn_debug_gemm_tf32: unknown form %d", form);
}

extern "C" XATTN_API int xattn_debug_umma_probe(const void* A, const void* B, float* out32, float* out16, int32_t N, int32_t K,
                                                int32_t mode, void* stream) {
  XF_TRY(check_device_sm100());
  if (!A || !B || !out32 || !out16) XF_FAIL(XATTN_ERR_NULL, "xattn_debug_umma_probe: null pointer");
  return umma_probe(reinterpret_cast<cudaStream_t>(stream), A, B, out32, out16, N, K, mode);
}

// The dropout factors (0 or 1/keep) every kernel of the training path draws: drop_factor(p, seed, offset, stream, idx) for idx in
// [idx0, idx0 + n).  tests/test_gpu_rng.py reconstructs each dropout site of the reference from them (fusion.py:18-26, :161,
// :277/:280, :315; temporal.py:19) and checks keep-rate and 1/keep scaling.
namespace {
__global__ void dropout_factors_kernel(float p, uint64_t seed, uint64_t offset, uint32_t rng_stream, uint64_t idx0, int64_t n,
                                       float* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = xf::drop_factor(p, seed, offset, rng_stream, idx0 + (uint64_t)i);
}
}  // namespace
extern "C" XATTN_API int xattn_debug_dropout_factors(float p, uint64_t seed, uint64_t offset, uint32_t rng_stream, uint64_t idx0,
                                                     int64_t n, float* out, void* stream) {
  if (!out) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  XF_TRY(check_device_sm100());
  if (n <= 0) return XATTN_OK;
  dropout_factors_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p, seed, offset, rng_stream, idx0, n, out);
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
