// Backbone tail -> head hand-off (reference src/models/fusion.py:368-370, :377).  The frozen backbones end in
// an fp32 tensor — ResNet18's last feature map [B·T, 512, h, w] in front of its AdaptiveAvgPool2d (src/models/video.py:23), WavLM's
// last_hidden_state [B, Ta, 768] (src/models/wavlm_audio.py:183) — that the reference hands to the fusion head as fp32.  This
// kernel is that tail: the spatial mean (R = h·w values per channel; R = 1 for the audio states) fused with the bf16 rounding the
// tensor-core kernels read, written straight into the rows of a resident embedding table (FrozenEmbeddingCache) chosen by a
// per-clip index — no fp32 embedding tensor, no separate cast, no scatter pass.
#include "common.cuh"

namespace xf {

namespace {

// x [n_clips · rows_per_clip][D][R] fp32 -> out[(dst[clip] · rows_per_clip + j)][D] bf16.  One warp per (row, channel) when R > 1
// (lanes stride over the R contiguous values), one thread per 8 channels when R == 1 (plain vectorised cast).
__global__ void __launch_bounds__(256) emit_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                        const int64_t* __restrict__ dst, int64_t rows, int rows_per_clip, int D, int R) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (R == 1) {
    const int64_t n8 = rows * (D / 8);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
      const int64_t row = i / (D / 8);
      const int c8 = (int)(i - row * (D / 8)) * 8;
      const int64_t clip = row / rows_per_clip, orow = (dst ? dst[clip] : clip) * rows_per_clip + (row - clip * rows_per_clip);
      const float4 a = __ldg(reinterpret_cast<const float4*>(x + row * D + c8)), b = __ldg(reinterpret_cast<const float4*>(x + row * D + c8 + 4));
      __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w), p2 = __floats2bfloat162_rn(b.x, b.y),
                     p3 = __floats2bfloat162_rn(b.z, b.w);
      *reinterpret_cast<uint4*>(out + orow * D + c8) = make_uint4(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1),
                                                                  *reinterpret_cast<uint32_t*>(&p2), *reinterpret_cast<uint32_t*>(&p3));
    }
    return;
  }
  const float inv = 1.f / (float)R;
  for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
    const int64_t clip = row / rows_per_clip, orow = (dst ? dst[clip] : clip) * rows_per_clip + (row - clip * rows_per_clip);
    const float* xr = x + row * (int64_t)D * R;
    for (int c = warp; c < D; c += nw) {
      float s = 0.f;
      for (int r = lane; r < R; r += 32) s += __ldg(xr + (int64_t)c * R + r);
      s = warp_sum(s);
      if (lane == 0) out[orow * D + c] = __float2bfloat16_rn(s * inv);
    }
  }
}

}  // namespace

int emit_embeddings_bf16(cudaStream_t st, const float* x, void* out, const int64_t* dst, int n_clips, int rows_per_clip, int D, int R) {
  if (n_clips <= 0) return XATTN_OK;
  if (rows_per_clip <= 0 || D <= 0 || R <= 0) XF_FAIL(XATTN_ERR_SHAPE, "emit_embeddings: bad shape rows_per_clip=%d D=%d R=%d", rows_per_clip, D, R);
  if (R == 1 && D % 8 != 0) XF_FAIL(XATTN_ERR_SHAPE, "emit_embeddings: D %% 8 != 0 (D=%d)", D);
  const int64_t rows = (int64_t)n_clips * rows_per_clip;
  int64_t blocks = R == 1 ? (rows * (D / 8) + 255) / 256 : rows;
  if (blocks > 148 * 32) blocks = 148 * 32;
  emit_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, reinterpret_cast<__nv_bfloat16*>(out), dst, rows, rows_per_clip, D, R);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
