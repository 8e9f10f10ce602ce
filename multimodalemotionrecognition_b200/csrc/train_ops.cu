// Loss and optimizer kernels of the Stage-1 step.
//  - cross entropy with label smoothing, fused forward + gradient (reference: nn.CrossEntropyLoss at
//    src/train.py:1033, applied at :218)
//  - fused multi-tensor Adam over the flat parameter buffer (reference: torch.optim.Adam(weight_decay) at
//    src/train.py:872 — coupled L2; decoupled AdamW selectable because BASELINE.json names AdamW)
#include <math.h>

#include "common.cuh"

namespace xf {

__global__ void __launch_bounds__(256) ce_fwd_bwd_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int B,
                                                         int C, float eps, float grad_scale, float* __restrict__ loss_out,
                                                         float* __restrict__ dlogits) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  float loss = 0.f;
  if (row < B) {
    const float* z = logits + (int64_t)row * C;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, z[c]);
    mx = warp_max(mx);
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += expf(z[c] - mx);
    se = warp_sum(se);
    const float lse = mx + logf(se);
    const int y = (int)labels[row];
    const float invB = 1.f / (float)B;
    for (int c = lane; c < C; c += 32) {
      const float logp = z[c] - lse;
      const float t = eps / (float)C + (c == y ? 1.f - eps : 0.f);
      loss -= t * logp;
      if (dlogits) dlogits[(int64_t)row * C + c] = (expf(logp) - t) * invB * grad_scale;
    }
    loss = warp_sum(loss) * invB;
  }
  __shared__ float red[8];
  if (lane == 0) red[threadIdx.x >> 5] = loss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
    atomicAdd(loss_out, s);
  }
}

int ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int B, int C, float label_smoothing, float grad_scale,
                    float* loss_out, float* dlogits, cudaStream_t st) {
  XF_CUDA(cudaMemsetAsync(loss_out, 0, sizeof(float), st));
  if (B <= 0) return XATTN_OK;
  ce_fwd_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(logits, labels, B, C, label_smoothing, grad_scale, loss_out, dlogits);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

struct AdamHyper {
  float lr, beta1, beta2, eps, wd, step_size, inv_sqrt_bc2;
  int decoupled;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamHyper& h) {
  if (h.decoupled) {
    p *= 1.f - h.lr * h.wd;
  } else {
    g = fmaf(h.wd, p, g);
  }
  m = fmaf(h.beta1, m, (1.f - h.beta1) * g);
  v = fmaf(h.beta2, v, (1.f - h.beta2) * g * g);
  const float denom = sqrtf(v) * h.inv_sqrt_bc2 + h.eps;
  p -= h.step_size * (m / denom);
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t n, AdamHyper h) {
  const int64_t n4 = n >> 2;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 p4 = reinterpret_cast<float4*>(p)[i];
    const float4 g4 = reinterpret_cast<const float4*>(g)[i];
    float4 m4 = reinterpret_cast<float4*>(m)[i];
    float4 v4 = reinterpret_cast<float4*>(v)[i];
    adam_one(p4.x, g4.x, m4.x, v4.x, h);
    adam_one(p4.y, g4.y, m4.y, v4.y, h);
    adam_one(p4.z, g4.z, m4.z, v4.z, h);
    adam_one(p4.w, g4.w, m4.w, v4.w, h);
    reinterpret_cast<float4*>(p)[i] = p4;
    reinterpret_cast<float4*>(m)[i] = m4;
    reinterpret_cast<float4*>(v)[i] = v4;
  }
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) adam_one(p[i], g[i], m[i], v[i], h);
}

int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, float wd,
              int step, int decoupled, cudaStream_t st) {
  if (n <= 0) return XATTN_OK;
  if (step < 1) XF_FAIL(XATTN_ERR_SHAPE, "adam step must be >= 1 (got %d)", step);
  AdamHyper h;
  h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.wd = wd; h.decoupled = decoupled;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  h.step_size = (float)((double)lr / bc1);
  h.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  int64_t blocks = ((n >> 2) + 255) / 256;
  if (blocks < 1) blocks = 1;
  const int64_t cap = (int64_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  adam_kernel<<<(int)blocks, 256, 0, st>>>(p, g, m, v, n, h);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
