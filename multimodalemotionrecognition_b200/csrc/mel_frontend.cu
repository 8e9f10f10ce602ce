// Mel front end of the xattn path (reference src/models/fusion.py:379-384): encoders without encode_sequence() hand the
// fusion head a mel spectrogram [B, 1, n_mels, Ta]; the reference runs Conv1d(n_mels -> a_dim, kernel 3, padding 1) over time
// and permutes to [B, Ta, a_dim].
//
// A width-3 convolution IS the audio_seq_proj Linear applied to unfolded rows:
//     a_seq[b,t,o] = bias[o] + sum_{c,k} w[o,c,k] * mel[b,c,t+k-1]  =  (U[b,t,:] . W[o,:]) + bias[o]
// with U[b,t,c*3+k] = mel[b,c,t+k-1] (zero outside [0,Ta)) and W = audio_time_conv.weight viewed [a_dim, n_mels*3] — its
// memory order already is (c major, k minor).  So this file only unfolds: mel -> U [B, Ta, 3*n_mels] in the same dtype; the
// head then runs unchanged with Da = 3*n_mels and the XP_SEQ_W / XP_SEQ_B slots bound to audio_time_conv.{weight,bias}
// (forward, backward, fp32 / tf32 / bf16 alike; the weight gradient lands in the conv weight's own layout).
#include "common.cuh"

namespace xf {

constexpr int kMelTT = 32;   // time steps per CTA

// One CTA: clip b, time steps [t0, t0+32).  mel rows (contiguous in t) are staged in shared memory with a one-step halo
// so both the reads (along t) and the writes (along c*3+k) are coalesced.
template <typename T>
__global__ void __launch_bounds__(256) mel_unfold_kernel(const T* __restrict__ mel, T* __restrict__ out, int n_mels, int Ta) {
  extern __shared__ unsigned char smem_raw[];
  T* tile = reinterpret_cast<T*>(smem_raw);            // [n_mels][kMelTT + 2]
  constexpr int W = kMelTT + 2;
  const int b = blockIdx.y, t0 = blockIdx.x * kMelTT;
  const T* src = mel + (size_t)b * n_mels * Ta;
  for (int i = threadIdx.x; i < n_mels * W; i += blockDim.x) {
    int c = i / W, j = i - c * W, t = t0 + j - 1;
    tile[i] = (t >= 0 && t < Ta) ? src[(size_t)c * Ta + t] : T(0.f);
  }
  __syncthreads();
  const int cols = 3 * n_mels;
  const int rows = min(kMelTT, Ta - t0);
  T* dst = out + ((size_t)b * Ta + t0) * cols;
  for (int i = threadIdx.x; i < rows * cols; i += blockDim.x) {
    int tt = i / cols, col = i - tt * cols, c = col / 3, k = col - c * 3;
    dst[i] = tile[c * W + tt + k];
  }
}

int mel_unfold(cudaStream_t st, const void* mel, void* out, int io_dtype, int B, int n_mels, int Ta) {
  if (B == 0) return XATTN_OK;
  const size_t esz = io_dtype == XATTN_DTYPE_BF16 ? 2 : 4;
  const size_t smem = (size_t)n_mels * (kMelTT + 2) * esz;
  if (smem > 48 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "xattn_mel_unfold: n_mels=%d exceeds the staged tile (max %d)", n_mels, (int)(48 * 1024 / ((kMelTT + 2) * esz)));
  if (B > 65535) XF_FAIL(XATTN_ERR_SHAPE, "xattn_mel_unfold: B=%d > 65535 clips per call", B);
  dim3 grid((Ta + kMelTT - 1) / kMelTT, B);
  if (io_dtype == XATTN_DTYPE_BF16)
    mel_unfold_kernel<__nv_bfloat16><<<grid, 256, smem, st>>>((const __nv_bfloat16*)mel, (__nv_bfloat16*)out, n_mels, Ta);
  else
    mel_unfold_kernel<float><<<grid, 256, smem, st>>>((const float*)mel, (float*)out, n_mels, Ta);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf

// include/xattn_fusion.h: xattn_mel_unfold
extern "C" XATTN_API int xattn_mel_unfold(const void* mel, void* unfolded, int32_t io_dtype, int32_t B, int32_t n_mels, int32_t Ta,
                                          void* stream) {
  using namespace xf;
  if (B < 0 || n_mels <= 0 || Ta <= 0) XF_FAIL(XATTN_ERR_SHAPE, "xattn_mel_unfold: B=%d n_mels=%d Ta=%d", B, n_mels, Ta);
  if (io_dtype != XATTN_DTYPE_F32 && io_dtype != XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "xattn_mel_unfold: io_dtype %d", io_dtype);
  if (B == 0) return XATTN_OK;
  if (!mel || !unfolded) XF_FAIL(XATTN_ERR_NULL, "xattn_mel_unfold: NULL pointer argument");
  XF_TRY(check_device_sm100());
  return mel_unfold((cudaStream_t)stream, mel, unfolded, io_dtype, B, n_mels, Ta);
}
