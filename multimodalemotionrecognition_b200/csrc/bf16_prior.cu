// Emotion-prior network of the bf16 path, batched over clips (reference src/models/fusion.py:158-181).
//   pooled = [mean_t v0 ; mean_t a0]                      :180   (mean_t a0 from the a0 GEMM's per-tile column sums)
//   prior  = Linear(Hd->P)(ReLU(Linear(2E->Hd)(pooled)))  :158-163, :181   (Dropout: eval identity)
//   pc[i]  = w_i[E:]·prior + b_i  for the four Linear(E+P -> 1) bias heads v_query, a_key, a_query, v_key   :171-174
// The token parts w_i[:E]·x_t of those heads are produced where the tokens are (video_pre, a0 GEMM epilogue), so the
// attention kernels only add pc[i] — concat([x_t, prior]) is never materialised.
// One CTA handles 16 clips: layer 0 is a 16 x Hd x 256 product on bf16 mma.sync with the fp32 activations split into a
// hi + lo bf16 pair (only the weights are rounded).
#include "bf16_layout.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kPriorClips = 16;
constexpr int kXP = 2 * BE + 8;   // bf16 pitch of the split activations (row g -> bank 4g: conflict-free fragment loads)

struct PriorArgs {
  const float *meanv, *acol;            // [B,E], [B,kColsumSlots,E]
  const __nv_bfloat16* pw0b;            // prior_net.0.weight as bf16 [Hd,2E]
  const float *pb0, *pw3, *pb3;         // prior_net.0 bias, prior_net.3 [P,Hd], [P]
  const float* wh[4];                   // v_query, a_key, a_query, v_key bias heads: weight [E+P]
  const float* bh[4];
  float* pc;                            // [B,4]
  int B, Ta, P, Hd;
};

constexpr int kPriorThreads = 256;
__global__ void __launch_bounds__(kPriorThreads) prior_kernel(PriorArgs a) {
  __shared__ __align__(16) __nv_bfloat16 xhi[kPriorClips][kXP];
  __shared__ __align__(16) __nv_bfloat16 xlo[kPriorClips][kXP];
  __shared__ float hid[kPriorClips][128 + 1];
  __shared__ float prior[kPriorClips][32 + 1];
  const int tid = threadIdx.x, c0 = blockIdx.x * kPriorClips;
  const int nc = min(kPriorClips, a.B - c0);
  const float invTa = 1.f / (float)a.Ta;
  for (int i = tid; i < kPriorClips * 2 * BE / 4; i += kPriorThreads) {   // one float4 of the pooled input per step
    const int cc = i / (2 * BE / 4), k = (i % (2 * BE / 4)) * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (cc < nc) {
      const int64_t c = c0 + cc;
      if (k < BE) {
        x = *reinterpret_cast<const float4*>(a.meanv + c * BE + k);
      } else {   // add the partial sums of the 32-row output groups the clip overlaps, first group first
        const int64_t r0 = c * a.Ta;
        const int nslot = (int)((r0 + a.Ta - 1) / 32 - r0 / 32) + 1;
        float4 v[kColsumSlots];
#pragma unroll
        for (int s = 0; s < kColsumSlots; ++s)
          v[s] = s < nslot ? *reinterpret_cast<const float4*>(a.acol + (c * kColsumSlots + s) * BE + (k - BE)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int s = 0; s < kColsumSlots; ++s) {
          x.x += v[s].x; x.y += v[s].y; x.z += v[s].z; x.w += v[s].w;
        }
        x.x *= invTa; x.y *= invTa; x.z *= invTa; x.w *= invTa;
      }
    }
    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat16 h = __float2bfloat16(xs[e]);
      xhi[cc][k + e] = h;
      xlo[cc][k + e] = __float2bfloat16(xs[e] - __bfloat162float(h));
    }
  }
  __syncthreads();
  {
    const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int ntiles = (a.Hd + 7) / 8;
    for (int nt = warp; nt < ntiles; nt += kPriorThreads / 32) {
      const int n = nt * 8 + g;
      const __nv_bfloat16* wrow = a.pw0b + (int64_t)min(n, a.Hd - 1) * (2 * BE) + 2 * t;
      uint32_t b[2 * BE / 16][2];
#pragma unroll
      for (int ks = 0; ks < 2 * BE / 16; ++ks) {
        asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(b[ks][0]) : "l"(wrow + ks * 16));
        asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(b[ks][1]) : "l"(wrow + ks * 16 + 8));
      }
      float d[4] = {0.f, 0.f, 0.f, 0.f}, e[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int ks = 0; ks < 2 * BE / 16; ++ks) {
        const int col = ks * 16 + 2 * t;
        uint32_t hi[4], lo[4];
        hi[0] = *reinterpret_cast<const uint32_t*>(&xhi[g][col]);      hi[1] = *reinterpret_cast<const uint32_t*>(&xhi[g + 8][col]);
        hi[2] = *reinterpret_cast<const uint32_t*>(&xhi[g][col + 8]);  hi[3] = *reinterpret_cast<const uint32_t*>(&xhi[g + 8][col + 8]);
        lo[0] = *reinterpret_cast<const uint32_t*>(&xlo[g][col]);      lo[1] = *reinterpret_cast<const uint32_t*>(&xlo[g + 8][col]);
        lo[2] = *reinterpret_cast<const uint32_t*>(&xlo[g][col + 8]);  lo[3] = *reinterpret_cast<const uint32_t*>(&xlo[g + 8][col + 8]);
        tc::mma_bf16(d, hi, b[ks][0], b[ks][1]);
        tc::mma_bf16(e, lo, b[ks][0], b[ks][1]);
      }
      // C fragment: rows g, g+8 (clips), columns nt*8 + 2t, +1 (hidden units)
      const int u0 = nt * 8 + 2 * t;
      if (u0 < a.Hd) {
        const float bb = a.pb0[u0];
        hid[g][u0] = fmaxf(d[0] + e[0] + bb, 0.f);
        hid[g + 8][u0] = fmaxf(d[2] + e[2] + bb, 0.f);
      }
      if (u0 + 1 < a.Hd) {
        const float bb = a.pb0[u0 + 1];
        hid[g][u0 + 1] = fmaxf(d[1] + e[1] + bb, 0.f);
        hid[g + 8][u0 + 1] = fmaxf(d[3] + e[3] + bb, 0.f);
      }
    }
  }
  __syncthreads();
  for (int o = tid; o < kPriorClips * a.P; o += kPriorThreads) {
    const int cc = o / a.P, p = o % a.P;
    const float* w = a.pw3 + (int64_t)p * a.Hd;
    float s0 = a.pb3[p], s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int k = 0;
    for (; k + 3 < a.Hd; k += 4) {
      s0 = fmaf(__ldg(w + k), hid[cc][k], s0);
      s1 = fmaf(__ldg(w + k + 1), hid[cc][k + 1], s1);
      s2 = fmaf(__ldg(w + k + 2), hid[cc][k + 2], s2);
      s3 = fmaf(__ldg(w + k + 3), hid[cc][k + 3], s3);
    }
    for (; k < a.Hd; ++k) s0 = fmaf(__ldg(w + k), hid[cc][k], s0);
    prior[cc][p] = (s0 + s1) + (s2 + s3);
  }
  __syncthreads();
  if (tid < kPriorClips * 4) {
    const int cc = tid >> 2, i = tid & 3;
    if (cc < nc) {
      const float* w = a.wh[i] + BE;
      float s = a.bh[i][0];
      for (int p = 0; p < a.P; ++p) s = fmaf(w[p], prior[cc][p], s);
      a.pc[(int64_t)(c0 + cc) * 4 + i] = s;
    }
  }
}

}  // namespace

int launch_prior(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  PriorArgs a;
  a.meanv = reinterpret_cast<const float*>(x.ws + x.w.meanv);
  a.acol = reinterpret_cast<const float*>(x.ws + x.w.acol);
  a.pw0b = reinterpret_cast<const __nv_bfloat16*>(x.packed + x.pk.pW0);
  a.pb0 = P(XP_PRIOR0_B); a.pw3 = P(XP_PRIOR3_W); a.pb3 = P(XP_PRIOR3_B);
  a.wh[0] = P(XP_VQ_W); a.bh[0] = P(XP_VQ_B); a.wh[1] = P(XP_AK_W); a.bh[1] = P(XP_AK_B);
  a.wh[2] = P(XP_AQ_W); a.bh[2] = P(XP_AQ_B); a.wh[3] = P(XP_VK_W); a.bh[3] = P(XP_VK_B);
  a.pc = reinterpret_cast<float*>(x.ws + x.w.pc);
  a.B = c.B; a.Ta = c.Ta; a.P = c.prior_dim; a.Hd = c.prior_hidden;
  prior_kernel<<<(c.B + kPriorClips - 1) / kPriorClips, kPriorThreads, 0, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
