// bf16 tensor-core forward of the xattn fusion head (precision == XATTN_DTYPE_BF16): launch sequence + classifier head.
// Stage map against reference src/models/fusion.py is in bf16_layout.cuh.
#include "bf16_layout.cuh"
#include "gemm_tc.cuh"
#include "kernels_bf16.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kHeadClips = 16;   // clips per CTA of the head kernel: each weight element is read once per 16 clips

struct HeadPooledArgs {
  const float* pooled;   // [B,2E] = [pooled_v ; pooled_a]
  const __nv_bfloat16* w0b;   // [HW,2E] layer 0 weight, bf16
  const float *b0, *w1, *b1, *wc, *bc;
  float* logits;
  int B, HW, C, gated;
};

// concat head: Linear(2E->HW), ReLU, (Dropout eval), Linear(HW->C)                       (fusion.py:405-407)
// gated head : g = sigmoid(Linear(E->1)(ReLU(Linear(2E->E)))), fused = g*v + (1-g)*a, Linear(E->C)   (:408-411)
__global__ void __launch_bounds__(256) head_pooled_kernel(HeadPooledArgs a) {
  __shared__ __align__(16) float pooled[kHeadClips][2 * BE];
  __shared__ __align__(16) float hid[kHeadClips][256];
  __shared__ float gate[kHeadClips];
  const int tid = threadIdx.x, c0 = blockIdx.x * kHeadClips;
  const int nc = min(kHeadClips, a.B - c0);
  for (int i = tid; i < kHeadClips * 2 * BE; i += blockDim.x) {
    const int cc = i / (2 * BE);
    pooled[cc][i % (2 * BE)] = cc < nc ? a.pooled[(int64_t)(c0 + cc) * 2 * BE + (i % (2 * BE))] : 0.f;
  }
  __syncthreads();
  {
    // hid = relu(pooled · W0^T + b0) with bf16 mma.sync: M = the 16 clips, N = HW, K = 2E.  The fp32 activations are fed as
    // a hi + lo bf16 pair (two MMAs) so only the weights are rounded, as everywhere else on this path.
    const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int ntw = a.HW / 64;   // n-tiles per warp (8 warps)
    for (int n = 0; n < ntw; ++n) {
      const int nt = warp * ntw + n;
      const __nv_bfloat16* wrow = a.w0b + (int64_t)(nt * 8 + g) * (2 * BE);
      float d[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
      for (int ks = 0; ks < 2 * BE / 16; ++ks) {
        const int col = ks * 16 + 2 * t;
        const uint32_t b0 = __ldg(reinterpret_cast<const uint32_t*>(wrow + col));
        const uint32_t b1 = __ldg(reinterpret_cast<const uint32_t*>(wrow + col + 8));
        const float2 x[4] = {*reinterpret_cast<const float2*>(&pooled[g][col]), *reinterpret_cast<const float2*>(&pooled[g + 8][col]),
                             *reinterpret_cast<const float2*>(&pooled[g][col + 8]),
                             *reinterpret_cast<const float2*>(&pooled[g + 8][col + 8])};
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const __nv_bfloat16 hx = __float2bfloat16(x[i].x), hy = __float2bfloat16(x[i].y);
          hi[i] = tc::pack_bf16(x[i].x, x[i].y);
          lo[i] = tc::pack_bf16(x[i].x - __bfloat162float(hx), x[i].y - __bfloat162float(hy));
        }
        tc::mma_bf16(d, hi, b0, b1);
        tc::mma_bf16(d, lo, b0, b1);
      }
      const int col = nt * 8 + 2 * t;
      const float c0 = a.b0[col], c1 = a.b0[col + 1];
      hid[g][col] = fmaxf(d[0] + c0, 0.f);
      hid[g][col + 1] = fmaxf(d[1] + c1, 0.f);
      hid[g + 8][col] = fmaxf(d[2] + c0, 0.f);
      hid[g + 8][col + 1] = fmaxf(d[3] + c1, 0.f);
    }
  }
  __syncthreads();
  if (a.gated) {
    if (tid < kHeadClips) {
      float s = a.b1[0];
      for (int k = 0; k < a.HW; ++k) s = fmaf(a.w1[k], hid[tid][k], s);
      gate[tid] = 1.f / (1.f + expf(-s));
    }
    __syncthreads();
    for (int i = tid; i < kHeadClips * BE; i += blockDim.x) {   // fused overwrites hid[:, :E]
      const int cc = i / BE, e = i % BE;
      hid[cc][e] = gate[cc] * pooled[cc][e] + (1.f - gate[cc]) * pooled[cc][BE + e];
    }
    __syncthreads();
  }
  const float* w = a.gated ? a.wc : a.w1;
  const float* b = a.gated ? a.bc : a.b1;
  const int K = a.gated ? BE : a.HW;
  for (int o = tid; o < nc * a.C; o += blockDim.x) {
    const int cc = o / a.C, cls = o % a.C;
    float s0 = b[cls], s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int k = 0; k < K; k += 4) {
      const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (int64_t)cls * K + k));
      const float4 hv = *reinterpret_cast<const float4*>(&hid[cc][k]);
      s0 = fmaf(wv.x, hv.x, s0); s1 = fmaf(wv.y, hv.y, s1); s2 = fmaf(wv.z, hv.z, s2); s3 = fmaf(wv.w, hv.w, s3);
    }
    const float s = (s0 + s1) + (s2 + s3);
    a.logits[(int64_t)(c0 + cc) * a.C + cls] = s;
  }
}

}  // namespace

int launch_head_pooled(const Bf16Ctx& x, float* logits) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  HeadPooledArgs a;
  a.pooled = reinterpret_cast<const float*>(x.ws + x.w.pooled);
  a.w0b = reinterpret_cast<const __nv_bfloat16*>(x.packed + x.pk.W0b);
  a.b0 = P(XP_HEAD0_B); a.w1 = P(XP_HEAD1_W); a.b1 = P(XP_HEAD1_B); a.wc = P(XP_CLS_W); a.bc = P(XP_CLS_B);
  a.logits = logits;
  a.B = c.B; a.C = c.C; a.gated = c.head_kind == XATTN_HEAD_GATED;
  a.HW = a.gated ? c.E : c.common_dim;
  head_pooled_kernel<<<(c.B + kHeadClips - 1) / kHeadClips, 256, 0, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int forward_bf16(const xattn_config& c, const float* params, const void* packed, const void* xv, const void* xa, float* logits,
                 void* ws, cudaStream_t st) {
  XF_TRY(bf16_supported(c));
  Bf16Ctx x;
  x.cfg = &c;
  x.params = params;
  compute_param_layout(c, x.pl);
  x.packed = reinterpret_cast<const uint8_t*>(packed);
  compute_pack_bf16(c, x.pk);
  x.ws = reinterpret_cast<uint8_t*>(ws);
  compute_ws_bf16(c, x.w);
  x.st = st;
  // v_in_proj (fusion.py:372) and a_in_proj∘audio_seq_proj (:378, :386) on tcgen05
  XF_TRY(gemm_tc_bf16(st, xv, x.packed + x.pk.Wv, params + x.pl.off[XP_V_IN_B], x.ws + x.w.v0, false, BE, c.B * c.T, BE, c.Dv));
  XF_TRY(gemm_tc_bf16(st, xa, x.packed + x.pk.Wfold, reinterpret_cast<const float*>(x.packed + x.pk.bfold), x.ws + x.w.a0, true,
                      BEP, c.B * c.Ta, BE, c.Da));
  XF_TRY(launch_video_pre(x));
  XF_TRY(launch_v2a_attn(x));
  XF_TRY(launch_video_update(x));
  XF_TRY(launch_a2v_attn(x));
  XF_TRY(launch_head_pooled(x, logits));
  return XATTN_OK;
}

}  // namespace xf
