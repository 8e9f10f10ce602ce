// Video-token side of the bf16 path: one WARP per clip (the 16 video tokens are exactly one m16 MMA tile), all weight
// matrices resident in shared memory for the lifetime of a persistent CTA (one per SM), bf16 mma.sync with fp32
// accumulation.  Two kernels, on either side of the video-queries-audio attention:
//   video_pre_kernel     v0 -> Qt (the per-clip 64x128 operand that replaces Q·K^T's audio-side K projection) + prior inputs
//   video_update_kernel  U -> attention output -> out_proj -> v1 = LN(v0 + .) -> pooled_v, Kt, c, Vt
// Reference: src/models/fusion.py:394-395 (v2a_attn, v_norm) and the K/V side of :398 (a2v_attn).
#include <cstdlib>

#include "bf16_layout.cuh"
#include "frag.cuh"

namespace xf {

namespace {

constexpr int kWBytes = BE * BEP * 2;   // one 128x128 bf16 matrix with the 136-element pitch: 34,816 B

struct VideoPreArgs {
  const float* v0;
  const __nv_bfloat16 *Wq1, *Wk1;
  const float *bq1, *wvq, *wvk;   // wvq / wvk: first E entries of v_query_bias / v_key_bias weights (or nullptr)
  __nv_bfloat16* qt;
  float *meanv, *tqv, *tkv;
  int B, T;
  float scale;
};

__global__ void __launch_bounds__(256, 2) video_pre_kernel(VideoPreArgs a) {
  extern __shared__ __align__(16) uint8_t sm[];
  __nv_bfloat16* Wq1s = reinterpret_cast<__nv_bfloat16*>(sm);
  __nv_bfloat16* Wk1s = reinterpret_cast<__nv_bfloat16*>(sm + kWBytes);
  float* bq1s = reinterpret_cast<float*>(sm + 2 * kWBytes);
  float* wvqs = bq1s + BE;
  float* wvks = wvqs + BE;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  fr::copy_g2s(Wq1s, a.Wq1, kWBytes, tid, blockDim.x);
  fr::copy_g2s(Wk1s, a.Wk1, kWBytes, tid, blockDim.x);
  for (int i = tid; i < BE; i += blockDim.x) {
    bq1s[i] = a.bq1[i];
    wvqs[i] = a.wvq ? a.wvq[i] : 0.f;
    wvks[i] = a.wvk ? a.wvk[i] : 0.f;
  }
  __syncthreads();
  const uint32_t wq = tc::smem_u32(Wq1s), wk = tc::smem_u32(Wk1s);
  const int T = a.T;
  const float invT = 1.f / (float)T;
  for (int c = blockIdx.x * 8 + warp; c < a.B; c += gridDim.x * 8) {
    const float* v = a.v0 + (int64_t)c * T * BE;
    uint32_t af[8][4];
    float sq0 = 0.f, sq1 = 0.f, sk0 = 0.f, sk1 = 0.f;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int col = ks * 16 + 2 * t;
      float2 x00 = make_float2(0.f, 0.f), x01 = x00, x10 = x00, x11 = x00;
      if (g < T) {
        x00 = *reinterpret_cast<const float2*>(v + g * BE + col);
        x01 = *reinterpret_cast<const float2*>(v + g * BE + col + 8);
      }
      if (g + 8 < T) {
        x10 = *reinterpret_cast<const float2*>(v + (g + 8) * BE + col);
        x11 = *reinterpret_cast<const float2*>(v + (g + 8) * BE + col + 8);
      }
      af[ks][0] = tc::pack_bf16(x00.x, x00.y);
      af[ks][1] = tc::pack_bf16(x10.x, x10.y);
      af[ks][2] = tc::pack_bf16(x01.x, x01.y);
      af[ks][3] = tc::pack_bf16(x11.x, x11.y);
      sq0 += wvqs[col] * x00.x + wvqs[col + 1] * x00.y + wvqs[col + 8] * x01.x + wvqs[col + 9] * x01.y;
      sq1 += wvqs[col] * x10.x + wvqs[col + 1] * x10.y + wvqs[col + 8] * x11.x + wvqs[col + 9] * x11.y;
      sk0 += wvks[col] * x00.x + wvks[col + 1] * x00.y + wvks[col + 8] * x01.x + wvks[col + 9] * x01.y;
      sk1 += wvks[col] * x10.x + wvks[col + 1] * x10.y + wvks[col + 8] * x11.x + wvks[col + 9] * x11.y;
      float c0 = x00.x + x10.x, c1 = x00.y + x10.y, c2 = x01.x + x11.x, c3 = x01.y + x11.y;   // column sums (mean_t v0)
#pragma unroll
      for (int o = 4; o <= 16; o <<= 1) {
        c0 += __shfl_xor_sync(0xffffffffu, c0, o);
        c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        c2 += __shfl_xor_sync(0xffffffffu, c2, o);
        c3 += __shfl_xor_sync(0xffffffffu, c3, o);
      }
      if (g == 0) {
        float* mv = a.meanv + (int64_t)c * BE + col;
        mv[0] = c0 * invT; mv[1] = c1 * invT; mv[8] = c2 * invT; mv[9] = c3 * invT;
      }
    }
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
      sq0 += __shfl_xor_sync(0xffffffffu, sq0, o);
      sq1 += __shfl_xor_sync(0xffffffffu, sq1, o);
      sk0 += __shfl_xor_sync(0xffffffffu, sk0, o);
      sk1 += __shfl_xor_sync(0xffffffffu, sk1, o);
    }
    if (t == 0) {
      a.tqv[(int64_t)c * 16 + g] = sq0; a.tqv[(int64_t)c * 16 + g + 8] = sq1;
      a.tkv[(int64_t)c * 16 + g] = sk0; a.tkv[(int64_t)c * 16 + g + 8] = sk1;
    }
    // per head: Q~_h = (v0 · Wq1_h^T + bq1_h) / sqrt(dh)  (torch scales q before the product), then
    // Qt_h[i, e] = sum_d Q~_h[i, d] * Wk1[32h+d, e]
#pragma unroll 1
    for (int h = 0; h < BH; ++h) {
      float acc[4][4];
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        const int nt = 4 * h + n;
        acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ks += 2) {
          uint32_t b[4];
          fr::ldB_nk(wq, nt, ks, lane, b);
          tc::mma_bf16(acc[n], af[ks], b[0], b[1]);
          tc::mma_bf16(acc[n], af[ks + 1], b[2], b[3]);
        }
        const int col = nt * 8 + 2 * t;
        acc[n][0] = (acc[n][0] + bq1s[col]) * a.scale;
        acc[n][1] = (acc[n][1] + bq1s[col + 1]) * a.scale;
        acc[n][2] = (acc[n][2] + bq1s[col]) * a.scale;
        acc[n][3] = (acc[n][3] + bq1s[col + 1]) * a.scale;
      }
      uint32_t aq[2][4];
      fr::c2a(acc[0], acc[1], aq[0]);
      fr::c2a(acc[2], acc[3], aq[1]);
      __nv_bfloat16* q0 = a.qt + ((int64_t)c * 64 + h * 16 + g) * BOP + 2 * t;
#pragma unroll
      for (int nt2 = 0; nt2 < 8; ++nt2) {
        float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          uint32_t b[4];
          fr::ldB_kn(wk, 32 * h + 16 * kk, nt2, lane, b);
          tc::mma_bf16(d0, aq[kk], b[0], b[1]);
          tc::mma_bf16(d1, aq[kk], b[2], b[3]);
        }
        *reinterpret_cast<uint32_t*>(q0 + nt2 * 16) = tc::pack_bf16(d0[0], d0[1]);
        *reinterpret_cast<uint32_t*>(q0 + nt2 * 16 + 8 * BOP) = tc::pack_bf16(d0[2], d0[3]);
        *reinterpret_cast<uint32_t*>(q0 + nt2 * 16 + 8) = tc::pack_bf16(d1[0], d1[1]);
        *reinterpret_cast<uint32_t*>(q0 + nt2 * 16 + 8 + 8 * BOP) = tc::pack_bf16(d1[2], d1[3]);
      }
    }
  }
}

struct VideoUpdArgs {
  const __nv_bfloat16* u;    // [B,4,32,8,4] bf16x2: U_h = softmax(S_h) · a0 in mma A-fragment order (see v2a_attn_kernel)
  const float* v0;           // [B,T,E]
  const __nv_bfloat16* W[6]; // Wv1, Wo1, Wk2, Wv2, Wq2, Wo2
  const float* vec[8];       // bv1, bo1, bk2, bv2, bq2, v_norm.weight, v_norm.bias, bo2
  __nv_bfloat16 *kt, *vt;    // [B,64,BOP]
  float *cv, *pooled;        // [B,64], [B,2E] (first E = pooled_v)
  int B, T;
  float scale;
};

// acc[16][4] (16 x 128, fp32) = A(16 x 128 as 8 k-step fragments) · W^T with W stored [n][k] in smem, + bias
__device__ __forceinline__ void proj128(float (&acc)[16][4], const uint32_t (&af)[8][4], uint32_t w, const float* bias, int lane) {
  const int t = lane & 3;
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) {
    acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 8; ks += 2) {
      uint32_t b[4];
      fr::ldB_nk(w, nt, ks, lane, b);
      tc::mma_bf16(acc[nt], af[ks], b[0], b[1]);
      tc::mma_bf16(acc[nt], af[ks + 1], b[2], b[3]);
    }
    const float b0 = bias[nt * 8 + 2 * t], b1 = bias[nt * 8 + 2 * t + 1];
    acc[nt][0] += b0; acc[nt][1] += b1; acc[nt][2] += b0; acc[nt][3] += b1;
  }
}

constexpr int kUpdWarps = 16;

// 4 n-tiles (32 output columns [32h, 32h+32)) of A(16 x 128) · W^T, + bias
__device__ __forceinline__ void proj32(float (&acc)[4][4], const uint32_t (&af)[8][4], uint32_t w, const float* bias, int h, int lane) {
  const int t = lane & 3;
#pragma unroll
  for (int n = 0; n < 4; ++n) {
    const int nt = 4 * h + n;
    acc[n][0] = acc[n][1] = acc[n][2] = acc[n][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 8; ks += 2) {
      uint32_t b[4];
      fr::ldB_nk(w, nt, ks, lane, b);
      tc::mma_bf16(acc[n], af[ks], b[0], b[1]);
      tc::mma_bf16(acc[n], af[ks + 1], b[2], b[3]);
    }
    const float b0 = bias[nt * 8 + 2 * t], b1 = bias[nt * 8 + 2 * t + 1];
    acc[n][0] += b0; acc[n][1] += b1; acc[n][2] += b0; acc[n][3] += b1;
  }
}

__global__ void __launch_bounds__(kUpdWarps * 32, 1) video_update_kernel(VideoUpdArgs a) {
  extern __shared__ __align__(16) uint8_t sm[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  // the six weight matrices arrive as bulk async copies (one thread issues, the copy engine streams 204 KB from L2)
  __shared__ uint64_t wbar;
  if (tid == 0) {
    tc::mbar_init(&wbar, 1);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(&wbar, 6 * kWBytes);
#pragma unroll
    for (int i = 0; i < 6; ++i) tc::bulk_load_1d(sm + i * kWBytes, a.W[i], kWBytes, &wbar);
  }
  float* vecs = reinterpret_cast<float*>(sm + 6 * kWBytes);
  for (int i = tid; i < 8 * BE; i += blockDim.x) vecs[i] = a.vec[i / BE][i % BE];
  __syncthreads();
  tc::mbar_wait(&wbar, 0);
  const uint32_t wv1 = tc::smem_u32(sm), wo1 = wv1 + kWBytes, wk2 = wv1 + 2 * kWBytes, wv2 = wv1 + 3 * kWBytes,
                 wq2 = wv1 + 4 * kWBytes, wo2 = wv1 + 5 * kWBytes;
  const float *bv1 = vecs, *bo1 = vecs + BE, *bk2 = vecs + 2 * BE, *bv2 = vecs + 3 * BE, *bq2 = vecs + 4 * BE,
              *gam = vecs + 5 * BE, *bet = vecs + 6 * BE, *bo2 = vecs + 7 * BE;
  const int T = a.T;
  const float invT = 1.f / (float)T;
  for (int c = blockIdx.x * kUpdWarps + warp; c < a.B; c += gridDim.x * kUpdWarps) {
    // ---- attention output of v2a, head by head: O[i, 32h+d] = sum_e U_h[i,e] Wv1[32h+d, e] + bv1 (softmax rows sum to 1) ----
    uint32_t af[8][4];   // first the A fragments of O (k-steps 2h, 2h+1 <- head h), later those of v1
#pragma unroll
    for (int h = 0; h < BH; ++h) {
      const uint4* up = reinterpret_cast<const uint4*>(a.u) + ((int64_t)c * BH + h) * 256 + lane;   // fragment order, [ks][lane]
      uint32_t au[8][4];
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint4 v = __ldg(up + ks * 32);   // a warp reads 512 contiguous bytes
        au[ks][0] = v.x; au[ks][1] = v.y; au[ks][2] = v.z; au[ks][3] = v.w;
      }
      float o4[4][4];
      proj32(o4, au, wv1, bv1, h, lane);
      fr::c2a(o4[0], o4[1], af[2 * h]);
      fr::c2a(o4[2], o4[3], af[2 * h + 1]);
    }
    // ---- out_proj, residual, LayerNorm (fusion.py:395) ----
    float acc[16][4];
    proj128(acc, af, wo1, bo1, lane);
    const float* v = a.v0 + (int64_t)c * T * BE;
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      const int col = nt * 8 + 2 * t;
      if (g < T) {
        const float2 x = *reinterpret_cast<const float2*>(v + g * BE + col);
        acc[nt][0] += x.x; acc[nt][1] += x.y;
      }
      if (g + 8 < T) {
        const float2 x = *reinterpret_cast<const float2*>(v + (g + 8) * BE + col);
        acc[nt][2] += x.x; acc[nt][3] += x.y;
      }
      s0 += acc[nt][0] + acc[nt][1];
      s1 += acc[nt][2] + acc[nt][3];
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
    const float mu0 = s0 * (1.f / BE), mu1 = s1 * (1.f / BE);
    float q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      float d;
      d = acc[nt][0] - mu0; q0 = fmaf(d, d, q0);
      d = acc[nt][1] - mu0; q0 = fmaf(d, d, q0);
      d = acc[nt][2] - mu1; q1 = fmaf(d, d, q1);
      d = acc[nt][3] - mu1; q1 = fmaf(d, d, q1);
    }
    q0 += __shfl_xor_sync(0xffffffffu, q0, 1); q0 += __shfl_xor_sync(0xffffffffu, q0, 2);
    q1 += __shfl_xor_sync(0xffffffffu, q1, 1); q1 += __shfl_xor_sync(0xffffffffu, q1, 2);
    const float r0 = rsqrtf(q0 * (1.f / BE) + 1e-5f), r1 = rsqrtf(q1 * (1.f / BE) + 1e-5f);
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      const int col = nt * 8 + 2 * t;
      const float g0 = gam[col], g1 = gam[col + 1], e0 = bet[col], e1 = bet[col + 1];
      acc[nt][0] = (acc[nt][0] - mu0) * r0 * g0 + e0;
      acc[nt][1] = (acc[nt][1] - mu0) * r0 * g1 + e1;
      acc[nt][2] = (acc[nt][2] - mu1) * r1 * g0 + e0;
      acc[nt][3] = (acc[nt][3] - mu1) * r1 * g1 + e1;
      // temporal mean of the updated video tokens (fusion.py:401, temporal.py:108-109)
      float p0 = (g < T ? acc[nt][0] : 0.f) + (g + 8 < T ? acc[nt][2] : 0.f);
      float p1 = (g < T ? acc[nt][1] : 0.f) + (g + 8 < T ? acc[nt][3] : 0.f);
#pragma unroll
      for (int o = 4; o <= 16; o <<= 1) {
        p0 += __shfl_xor_sync(0xffffffffu, p0, o);
        p1 += __shfl_xor_sync(0xffffffffu, p1, o);
      }
      if (g == 0) {
        a.pooled[(int64_t)c * 2 * BE + col] = p0 * invT;
        a.pooled[(int64_t)c * 2 * BE + col + 1] = p1 * invT;
      }
    }
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) fr::c2a(acc[2 * ks], acc[2 * ks + 1], af[ks]);   // v1 as the A operand
    // ---- keys of a2v, head by head: K2_h = v1·Wk2_h^T + bk2_h ; Kt_h = scale·K2_h·Wq2_h ; c_hi = scale·K2_h[i]·bq2_h ----
#pragma unroll 1
    for (int h = 0; h < BH; ++h) {
      float k4[4][4];
      proj32(k4, af, wk2, bk2, h, lane);
      uint32_t aq[2][4];
      fr::c2a(k4[0], k4[1], aq[0]);
      fr::c2a(k4[2], k4[3], aq[1]);
      float c0 = 0.f, c1 = 0.f;
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        const float w0 = bq2[(4 * h + n) * 8 + 2 * t], w1 = bq2[(4 * h + n) * 8 + 2 * t + 1];
        c0 += k4[n][0] * w0 + k4[n][1] * w1;
        c1 += k4[n][2] * w0 + k4[n][3] * w1;
      }
      c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
      c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
      if (t == 0) {
        a.cv[(int64_t)c * 64 + h * 16 + g] = c0 * a.scale;
        a.cv[(int64_t)c * 64 + h * 16 + g + 8] = c1 * a.scale;
      }
      __nv_bfloat16* k0 = a.kt + ((int64_t)c * 64 + h * 16 + g) * BOP + 2 * t;
#pragma unroll
      for (int nt2 = 0; nt2 < 8; ++nt2) {
        float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          uint32_t b[4];
          fr::ldB_kn(wq2, 32 * h + 16 * kk, nt2, lane, b);
          tc::mma_bf16(d0, aq[kk], b[0], b[1]);
          tc::mma_bf16(d1, aq[kk], b[2], b[3]);
        }
        *reinterpret_cast<uint32_t*>(k0 + nt2 * 16) = tc::pack_bf16(d0[0] * a.scale, d0[1] * a.scale);
        *reinterpret_cast<uint32_t*>(k0 + nt2 * 16 + 8 * BOP) = tc::pack_bf16(d0[2] * a.scale, d0[3] * a.scale);
        *reinterpret_cast<uint32_t*>(k0 + nt2 * 16 + 8) = tc::pack_bf16(d1[0] * a.scale, d1[1] * a.scale);
        *reinterpret_cast<uint32_t*>(k0 + nt2 * 16 + 8 + 8 * BOP) = tc::pack_bf16(d1[2] * a.scale, d1[3] * a.scale);
      }
    }
    // ---- values of a2v with out_proj folded in: V2_h = v1·Wv2_h^T + bv2_h ; Vt_h[i, e'] = sum_d V2_h[i,d] Wo2[e', 32h+d] ----
#pragma unroll 1
    for (int h = 0; h < BH; ++h) {
      float v4[4][4];
      proj32(v4, af, wv2, bv2, h, lane);
      uint32_t aq[2][4];
      fr::c2a(v4[0], v4[1], aq[0]);
      fr::c2a(v4[2], v4[3], aq[1]);
      __nv_bfloat16* v0p = a.vt + ((int64_t)c * 64 + h * 16 + g) * BOP + 2 * t;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        // + bo2/4 on every head's rows: the 4 heads' probabilities each sum to one, so P·Vt arrives with out_proj's bias
        const float bq0 = 0.25f * bo2[nt * 8 + 2 * t], bq1 = 0.25f * bo2[nt * 8 + 2 * t + 1];
        float d[4] = {bq0, bq1, bq0, bq1};
        uint32_t b[4];
        fr::ldB_nk(wo2, nt, 0, lane, b, 32 * h);
        tc::mma_bf16(d, aq[0], b[0], b[1]);
        tc::mma_bf16(d, aq[1], b[2], b[3]);
        *reinterpret_cast<uint32_t*>(v0p + nt * 8) = tc::pack_bf16(d[0], d[1]);
        *reinterpret_cast<uint32_t*>(v0p + nt * 8 + 8 * BOP) = tc::pack_bf16(d[2], d[3]);
      }
    }
  }
}

}  // namespace

static int video_grid(int B, int ctas_per_sm, int warps = 8) {
  const int need = (B + warps - 1) / warps;
  const int cap = num_sms() * ctas_per_sm;
  return need < cap ? need : cap;
}

int launch_video_pre_tc(const Bf16Ctx& x);

int launch_video_pre(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  static const bool legacy = getenv("XATTN_VIDEO_PRE_LEGACY") != nullptr;
  if (!legacy) return launch_video_pre_tc(x);
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  VideoPreArgs a;
  a.v0 = reinterpret_cast<const float*>(x.ws + x.w.v0);
  a.Wq1 = reinterpret_cast<const __nv_bfloat16*>(x.packed + x.pk.Wq1);
  a.Wk1 = reinterpret_cast<const __nv_bfloat16*>(x.packed + x.pk.Wk1);
  a.bq1 = P(XP_V2A_IN_B);
  a.wvq = c.use_prior ? P(XP_VQ_W) : nullptr;
  a.wvk = c.use_prior ? P(XP_VK_W) : nullptr;
  a.qt = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.qt);
  a.meanv = reinterpret_cast<float*>(x.ws + x.w.meanv);
  a.tqv = reinterpret_cast<float*>(x.ws + x.w.tqv);
  a.tkv = reinterpret_cast<float*>(x.ws + x.w.tkv);
  a.B = c.B; a.T = c.T;
  a.scale = 1.0f / sqrtf((float)BDH);
  const int smem = 2 * kWBytes + 3 * BE * 4;
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(video_pre_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  video_pre_kernel<<<video_grid(c.B, 2), 256, smem, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int launch_video_update(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  auto W = [&](int64_t off) { return reinterpret_cast<const __nv_bfloat16*>(x.packed + off); };
  VideoUpdArgs a;
  a.u = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.u);
  a.v0 = reinterpret_cast<const float*>(x.ws + x.w.v0);
  a.W[0] = W(x.pk.Wv1); a.W[1] = W(x.pk.Wo1); a.W[2] = W(x.pk.Wk2); a.W[3] = W(x.pk.Wv2); a.W[4] = W(x.pk.Wq2); a.W[5] = W(x.pk.Wo2);
  a.vec[0] = P(XP_V2A_IN_B) + 2 * BE;   // bv1
  a.vec[1] = P(XP_V2A_OUT_B);           // bo1
  a.vec[2] = P(XP_A2V_IN_B) + BE;       // bk2
  a.vec[3] = P(XP_A2V_IN_B) + 2 * BE;   // bv2
  a.vec[4] = P(XP_A2V_IN_B);            // bq2
  a.vec[5] = P(XP_VN_W);
  a.vec[6] = P(XP_VN_B);
  a.vec[7] = P(XP_A2V_OUT_B);          // bo2
  a.kt = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.kt);
  a.vt = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.vt);
  a.cv = reinterpret_cast<float*>(x.ws + x.w.cv);
  a.pooled = reinterpret_cast<float*>(x.ws + x.w.pooled);
  a.B = c.B; a.T = c.T;
  a.scale = 1.0f / sqrtf((float)BDH);
  const int smem = 6 * kWBytes + 8 * BE * 4;
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(video_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  video_update_kernel<<<video_grid(c.B, 1, kUpdWarps), kUpdWarps * 32, smem, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
