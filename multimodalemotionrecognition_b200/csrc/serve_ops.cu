// Epilogue of the serving path: what TorchModelRunner.predict_probs (optimized_runtime.py:99-108) and the
// worker loop (inference_worker.py:136-143) do between the logits and the JSON result — softmax(dim=1), argmax of the row and
// the probability at it — as one launch over the batch instead of a softmax kernel plus a per-row .argmax().item() sync.
#include "common.cuh"

namespace xf {

// One thread per clip (C is the class count, 8 here).  probs may alias logits.  top1: first maximal class (torch.argmax).
__global__ void __launch_bounds__(256) softmax_top1_kernel(const float* __restrict__ logits, int B, int C, float* probs,
                                                           int32_t* __restrict__ top1, float* __restrict__ top1_prob) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float* x = logits + (int64_t)b * C;
  float mx = x[0];
  int arg = 0;
  for (int c = 1; c < C; ++c) {
    const float v = x[c];
    if (v > mx) { mx = v; arg = c; }
  }
  float sum = 0.f;
  for (int c = 0; c < C; ++c) sum += expf(x[c] - mx);
  const float inv = 1.f / sum;
  if (probs)
    for (int c = 0; c < C; ++c) probs[(int64_t)b * C + c] = expf(x[c] - mx) * inv;
  if (top1) top1[b] = arg;
  if (top1_prob) top1_prob[b] = inv;   // exp(0) / sum
}

}  // namespace xf

// include/xattn_fusion.h: xattn_softmax_top1
extern "C" XATTN_API int xattn_softmax_top1(const float* logits, int32_t B, int32_t C, float* probs, int32_t* top1, float* top1_prob,
                                            void* stream) {
  using namespace xf;
  if (B < 0 || C <= 0) XF_FAIL(XATTN_ERR_SHAPE, "xattn_softmax_top1: B=%d C=%d", B, C);
  if (B == 0) return XATTN_OK;
  if (!logits) XF_FAIL(XATTN_ERR_NULL, "xattn_softmax_top1: NULL logits");
  XF_TRY(check_device_sm100());
  softmax_top1_kernel<<<(B + 255) / 256, 256, 0, (cudaStream_t)stream>>>(logits, B, C, probs, top1, top1_prob);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
