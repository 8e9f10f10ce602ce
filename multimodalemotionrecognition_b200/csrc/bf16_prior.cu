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

constexpr int kPriorClips = 16;   // (32 clips per CTA, as in the head kernel, measured slower here: 12.1 vs 10.9 us at B = 4096)
constexpr int kXP = 2 * BE + 8;   // bf16 pitch of the split activations (row g -> bank 4g: conflict-free fragment loads)

struct PriorArgs {
  const float *meanv, *acol;            // [B,E], [B,kColsumSlots,E]
  const __nv_bfloat16* pw0b;            // prior_net.0.weight as bf16, mma B-fragment order (frag_pack_kernel)
  const float *pb0, *pw3, *pb3;         // prior_net.0 bias, prior_net.3 [P,Hd], [P]
  const float* wh[4];                   // v_query, a_key, a_query, v_key bias heads: weight [E+P]
  const float* bh[4];
  float* pc;                            // [B,4]
  int B, Ta, P, Hd;
};

constexpr int kPriorThreads = 256;
constexpr int kPriorStage = kPriorClips * kColsumSlots * BE * 4;
constexpr int kPriorSmem = kPriorStage + 2 * kPriorClips * kXP * 2 + kPriorClips * 129 * 4 + kPriorClips * 33 * 4;
__global__ void __launch_bounds__(kPriorThreads) prior_kernel(PriorArgs a) {
  // dynamic smem: [astage: the clip block's a0 column sums][xhi][xlo][hid][prior]
  extern __shared__ __align__(16) unsigned char psm[];
  float* astage = reinterpret_cast<float*>(psm);   // [kPriorClips][kColsumSlots][BE]
  auto xhi = reinterpret_cast<__nv_bfloat16(*)[kXP]>(psm + kPriorStage);
  auto xlo = reinterpret_cast<__nv_bfloat16(*)[kXP]>(psm + kPriorStage + kPriorClips * kXP * 2);
  auto hid = reinterpret_cast<float(*)[128 + 1]>(psm + kPriorStage + 2 * kPriorClips * kXP * 2);
  auto prior = reinterpret_cast<float(*)[32 + 1]>(psm + kPriorStage + 2 * kPriorClips * kXP * 2 + kPriorClips * 129 * 4);
  const int tid = threadIdx.x, c0 = blockIdx.x * kPriorClips;
  const int nc = min(kPriorClips, a.B - c0);
  const float invTa = 1.f / (float)a.Ta;
  // weight fragments of a layer-0 tile: 8 x 512 contiguous bytes per warp (fragment order); the warp's first tile is fetched now,
  // its L2 round trip hides behind the activation loads
  const int wid = tid >> 5;
  auto fetchw = [&](uint4 (&bw)[2 * BE / 32], int nt) {
    const uint4* wf = reinterpret_cast<const uint4*>(a.pw0b) + ((int64_t)nt * (2 * BE / 32)) * 32 + (tid & 31);
#pragma unroll
    for (int kp = 0; kp < 2 * BE / 32; ++kp) bw[kp] = __ldg(wf + kp * 32);
  };
  uint4 bq[2 * BE / 32];
  if (wid < (a.Hd + 7) / 8) fetchw(bq, wid);
  tc::pdl_wait();     // (weights above are static; everything below reads what the previous kernels wrote)
  tc::pdl_launch();
  // the clip block's column sums (16 clips x 6 slots x 128 floats, contiguous) arrive as one bulk async copy
  __shared__ uint64_t abar;
  if (tid == 0) {
    tc::mbar_init(&abar, 1);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(&abar, (uint32_t)(nc * kColsumSlots * BE * 4));
    tc::bulk_load_1d(astage, a.acol + (int64_t)c0 * kColsumSlots * BE, (uint32_t)(nc * kColsumSlots * BE * 4), &abar);
  }
  auto put = [&](int cc, int k, float4 x) {
    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat16 h = __float2bfloat16(xs[e]);
      xhi[cc][k + e] = h;
      xlo[cc][k + e] = __float2bfloat16(xs[e] - __bfloat162float(h));
    }
  };
  {
    float4 xv[kPriorClips * BE / 4 / kPriorThreads];   // mean_t v0
#pragma unroll
    for (int r = 0; r < kPriorClips * BE / 4 / kPriorThreads; ++r) {
      const int i = tid + r * kPriorThreads, cc = i / (BE / 4), k = (i % (BE / 4)) * 4;
      xv[r] = cc < nc ? *reinterpret_cast<const float4*>(a.meanv + (int64_t)(c0 + cc) * BE + k) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int r = 0; r < kPriorClips * BE / 4 / kPriorThreads; ++r) {
      const int i = tid + r * kPriorThreads;
      put(i / (BE / 4), (i % (BE / 4)) * 4, xv[r]);
    }
  }
  __syncthreads();   // the barrier is initialised
  tc::mbar_wait(&abar, 0);
  for (int i = tid; i < kPriorClips * BE / 4; i += kPriorThreads) {
    const int cc = i / (BE / 4), k = (i % (BE / 4)) * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (cc < nc) {   // add the partial sums of the 32-row output groups the clip overlaps, first group first
      const int64_t r0 = (int64_t)(c0 + cc) * a.Ta;
      const int nslot = (int)((r0 + a.Ta - 1) / 32 - r0 / 32) + 1;
#pragma unroll
      for (int s = 0; s < kColsumSlots; ++s) {
        if (s < nslot) {
          const float4 v = *reinterpret_cast<const float4*>(astage + (cc * kColsumSlots + s) * BE + k);
          x.x += v.x; x.y += v.y; x.z += v.z; x.w += v.w;
        }
      }
      x.x *= invTa; x.y *= invTa; x.z *= invTa; x.w *= invTa;
    }
    put(cc, BE + k, x);
  }
  __syncthreads();
  {
    const int warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int ntiles = (a.Hd + 7) / 8;
    for (int nt = warp; nt < ntiles; nt += kPriorThreads / 32) {
      if (nt != wid) fetchw(bq, nt);   // (the warp's first tile was fetched at kernel entry)
      float d[kPriorClips / 16][4], e[kPriorClips / 16][4];
#pragma unroll
      for (int mt = 0; mt < kPriorClips / 16; ++mt)
#pragma unroll
        for (int r = 0; r < 4; ++r) d[mt][r] = e[mt][r] = 0.f;
      const uint32_t ahi = tc::smem_u32(&xhi[(lane & 7) + 8 * ((lane >> 3) & 1)][8 * (lane >> 4)]);
      const uint32_t alo = tc::smem_u32(&xlo[(lane & 7) + 8 * ((lane >> 3) & 1)][8 * (lane >> 4)]);
#pragma unroll
      for (int ks = 0; ks < 2 * BE / 16; ++ks) {
        const uint32_t b0 = (ks & 1) ? bq[ks >> 1].z : bq[ks >> 1].x, b1 = (ks & 1) ? bq[ks >> 1].w : bq[ks >> 1].y;
#pragma unroll
        for (int mt = 0; mt < kPriorClips / 16; ++mt) {
          uint32_t hi[4], lo[4];
          tc::ldsm_x4(ahi + mt * 16 * kXP * 2 + ks * 32, hi[0], hi[1], hi[2], hi[3]);
          tc::ldsm_x4(alo + mt * 16 * kXP * 2 + ks * 32, lo[0], lo[1], lo[2], lo[3]);
          tc::mma_bf16(d[mt], hi, b0, b1);
          tc::mma_bf16(e[mt], lo, b0, b1);
        }
      }
      // C fragment: rows g, g+8 (clips), columns nt*8 + 2t, +1 (hidden units)
      const int u0 = nt * 8 + 2 * t;
#pragma unroll
      for (int mt = 0; mt < kPriorClips / 16; ++mt) {
        if (u0 < a.Hd) {
          const float bb = a.pb0[u0];
          hid[16 * mt + g][u0] = fmaxf(d[mt][0] + e[mt][0] + bb, 0.f);
          hid[16 * mt + g + 8][u0] = fmaxf(d[mt][2] + e[mt][2] + bb, 0.f);
        }
        if (u0 + 1 < a.Hd) {
          const float bb = a.pb0[u0 + 1];
          hid[16 * mt + g][u0 + 1] = fmaxf(d[mt][1] + e[mt][1] + bb, 0.f);
          hid[16 * mt + g + 8][u0 + 1] = fmaxf(d[mt][3] + e[mt][3] + bb, 0.f);
        }
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
  constexpr int dyn = kPriorSmem;
  XF_TRY(ensure_dyn_smem(prior_kernel, dyn));
  XF_CUDA(launch_chain(prior_kernel, dim3((c.B + kPriorClips - 1) / kPriorClips), dim3(kPriorThreads), dyn, x.st, a));
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
