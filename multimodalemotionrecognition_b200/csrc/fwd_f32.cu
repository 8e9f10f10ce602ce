#include <stdlib.h>
// fp32 forward of the xattn fusion head (parity + training path).
// Stage map (reference src/models/fusion.py):
//   v_in_proj :372 | audio_seq_proj :378 | a_in_proj :386 | EmotionPriorBiasAdapter :170-184 |
//   v2a_attn :394 | v_norm(v + drop_path(v2)) :395 | a2v_attn :398 | a_norm :399 | mean pool :401-402
//   (temporal.py:108-109) | head :405-411.
// Every intermediate the backward needs is kept in the caller's workspace (layout: compute_ws_f32).
#include "gemm_simt.cuh"
#include "smem_tile.cuh"
#include "kernels_f32.cuh"

namespace xf {

// ---------------------------------------------------------------------------------------------------------
// Emotion prior (fusion.py:178-184): pooled means -> prior_net -> prior[P]; then the four token scores
//   qv[i] = w_vq[:E]·v_i + w_vq[E:]·prior + c_vq     ka[j] = w_ak[:E]·a_j + w_ak[E:]·prior + c_ak
//   qa[j] = w_aq[:E]·a_j + ...                        kvs[i] = w_vk[:E]·v_i + ...
// (the reference builds these by expand+cat+Linear(E+P -> 1), :171-174; the sum decomposes).
// One CTA per clip.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) prior_fwd_kernel(PriorArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int E = a.E, P = a.P, Hd = a.Hd, T = a.T, Ta = a.Ta;
  float* pooled = sm;            // [2E]
  float* hid = pooled + 2 * E;   // [Hd]
  float* prior = hid + Hd;       // [P]
  float* pc = prior + P;         // [4] prior contribution + bias of the four score heads
  float* scratch = pc + 4;       // [nw][E]
  const float* v0 = a.v0 + (int64_t)b * T * E;
  const float* a0 = a.a0 + (int64_t)b * Ta * E;
  clip_mean(v0, T, E, pooled, scratch, tid, blockDim.x);
  clip_mean(a0, Ta, E, pooled + E, scratch, tid, blockDim.x);
  for (int e = tid; e < 2 * E; e += blockDim.x) a.pooled0[(int64_t)b * 2 * E + e] = pooled[e];
  for (int h = warp; h < Hd; h += nw) {
    const float* w = a.w0 + (int64_t)h * 2 * E;
    float s = 0.f;
    for (int k = lane; k < 2 * E; k += 32) s = fmaf(w[k], pooled[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      float r = fmaxf(s + a.b0[h], 0.f);
      r *= drop_factor(a.p_drop, a.seed, a.offset, RS_PRIOR, (uint64_t)b * Hd + h);
      hid[h] = r;
      a.phid[(int64_t)b * Hd + h] = r;
    }
  }
  __syncthreads();
  for (int p = warp; p < P; p += nw) {
    const float* w = a.w3 + (int64_t)p * Hd;
    float s = 0.f;
    for (int k = lane; k < Hd; k += 32) s = fmaf(w[k], hid[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      prior[p] = s + a.b3[p];
      a.prior[(int64_t)b * P + p] = s + a.b3[p];
    }
  }
  __syncthreads();
  if (tid < 4) {
    const float* w = a.wh[tid] + E;
    float s = a.bh[tid][0];
    for (int p = 0; p < P; ++p) s = fmaf(w[p], prior[p], s);
    pc[tid] = s;
  }
  __syncthreads();
  // video tokens: heads 0 (v_query -> qv) and 3 (v_key -> kvs); audio tokens: heads 1 (a_key -> ka), 2 (a_query -> qa)
  if ((E & 31) == 0 && ((Hd + P) & 3) == 0) {   // (the second condition keeps `scratch` 16-byte aligned)
    // eight lanes per token (four tokens per warp and step): float4 loads, the four weight vectors from shared memory (the mean
    // scratch is free by now), three shuffle steps per dot product instead of five for a token per warp
    for (int i = tid; i < 4 * E; i += blockDim.x) scratch[i] = a.wh[i / E][i % E];
    __syncthreads();
    const int grp = lane >> 3, sl = lane & 7, n4 = E >> 2;
    for (int i0 = warp * 4; i0 < T + Ta; i0 += nw * 4) {
      const int i = min(i0 + grp, T + Ta - 1);
      const bool is_v = i < T;
      const float4* tok = reinterpret_cast<const float4*>(is_v ? v0 + (int64_t)i * E : a0 + (int64_t)(i - T) * E);
      const float4* wA = reinterpret_cast<const float4*>(scratch + (is_v ? 0 : 1) * E);
      const float4* wB = reinterpret_cast<const float4*>(scratch + (is_v ? 3 : 2) * E);
      float sA = 0.f, sB = 0.f;
#pragma unroll 4
      for (int j = sl; j < n4; j += 8) {
        const float4 x = __ldg(tok + j), u = wA[j], v = wB[j];
        sA = fmaf(u.x, x.x, fmaf(u.y, x.y, fmaf(u.z, x.z, fmaf(u.w, x.w, sA))));
        sB = fmaf(v.x, x.x, fmaf(v.y, x.y, fmaf(v.z, x.z, fmaf(v.w, x.w, sB))));
      }
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        sA += __shfl_xor_sync(0xffffffffu, sA, o);
        sB += __shfl_xor_sync(0xffffffffu, sB, o);
      }
      if (sl == 0 && i0 + grp < T + Ta) {
        if (is_v) {
          a.qv[(int64_t)b * T + i] = sA + pc[0];
          a.kvs[(int64_t)b * T + i] = sB + pc[3];
        } else {
          a.ka[(int64_t)b * Ta + (i - T)] = sA + pc[1];
          a.qa[(int64_t)b * Ta + (i - T)] = sB + pc[2];
        }
      }
    }
    return;
  }
  for (int i = warp; i < T + Ta; i += nw) {
    const bool is_v = i < T;
    const float* tok = is_v ? v0 + (int64_t)i * E : a0 + (int64_t)(i - T) * E;
    const float* wA = is_v ? a.wh[0] : a.wh[1];
    const float* wB = is_v ? a.wh[3] : a.wh[2];
    float sA = 0.f, sB = 0.f;
    for (int e = lane; e < E; e += 32) {
      const float x = tok[e];
      sA = fmaf(wA[e], x, sA);
      sB = fmaf(wB[e], x, sB);
    }
    sA = warp_sum(sA);
    sB = warp_sum(sB);
    if (lane == 0) {
      if (is_v) {
        a.qv[(int64_t)b * T + i] = sA + pc[0];
        a.kvs[(int64_t)b * T + i] = sB + pc[3];
      } else {
        a.ka[(int64_t)b * Ta + (i - T)] = sA + pc[1];
        a.qa[(int64_t)b * Ta + (i - T)] = sB + pc[2];
      }
    }
  }
}

// The same for d_model = 128 (a warp covers a token row with one float4 per lane) as ONE pass over the clip's tokens and a short
// dependency chain: the column sums for the two means and the raw token dot products with the four score vectors come from the
// same loads (four rows per warp in flight), the 2E -> Hd layer runs on all threads at once (four threads per hidden unit, 16
// float4 weight loads each: one L2 round trip instead of Hd / 8 dependent ones), and the per-clip constants are added to the raw
// dots at the end.  (The two-pass kernel above spent ~30 us per CTA on that chain; 1024 clips are two waves of CTAs.)
__device__ __forceinline__ float dot4(const float4& a, const float4& b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, a.w * b.w))); }
__global__ void __launch_bounds__(256, 4) prior_fwd128_kernel(PriorArgs a) {
  extern __shared__ __align__(16) float sm[];
  constexpr int E = 128;
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int P = a.P, Hd = a.Hd, T = a.T, Ta = a.Ta;
  float* pooled = sm;                  // [2E]
  float* scr = pooled + 2 * E;         // [2][nw][E] per-warp column sums (video | audio)
  float* rawA = scr + 2 * nw * E;      // [T + Ta] token · wA  (v_query | a_key)
  float* rawB = rawA + T + Ta;         // [T + Ta] token · wB  (v_key | a_query)
  float* hid = rawB + T + Ta;          // [Hd]
  float* prior = hid + Hd;             // [P]
  float* pc = prior + P;               // [4]
  const float* v0 = a.v0 + (int64_t)b * T * E;
  const float* a0 = a.a0 + (int64_t)b * Ta * E;
  auto rows = [&](const float* x, int L, const float4 wA, const float4 wB, float* outA, float* outB, float* colsum) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i0 = warp; i0 < L; i0 += 4 * nw) {
      float4 r[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * nw;
        r[u] = i < L ? __ldg(reinterpret_cast<const float4*>(x + (int64_t)i * E) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * nw;
        s.x += r[u].x; s.y += r[u].y; s.z += r[u].z; s.w += r[u].w;
        const float dA = warp_sum(dot4(r[u], wA)), dB = warp_sum(dot4(r[u], wB));
        if (lane == 0 && i < L) { outA[i] = dA; outB[i] = dB; }
      }
    }
    *reinterpret_cast<float4*>(colsum + warp * E + 4 * lane) = s;
  };
  rows(v0, T, __ldg(reinterpret_cast<const float4*>(a.wh[0]) + lane), __ldg(reinterpret_cast<const float4*>(a.wh[3]) + lane), rawA, rawB, scr);
  rows(a0, Ta, __ldg(reinterpret_cast<const float4*>(a.wh[1]) + lane), __ldg(reinterpret_cast<const float4*>(a.wh[2]) + lane), rawA + T,
       rawB + T, scr + nw * E);
  __syncthreads();
  {   // thread e < E: video mean of column e; e >= E: audio mean (warps in order, as clip_mean sums them)
    const float* src = scr + (tid >= E ? nw * E + (tid - E) : tid);
    float s = 0.f;
    for (int w = 0; w < nw; ++w) s += src[w * E];
    const float m = s / (float)(tid >= E ? Ta : T);
    pooled[tid] = m;
    a.pooled0[(int64_t)b * 2 * E + tid] = m;
  }
  __syncthreads();
  for (int h0 = 0; h0 < Hd; h0 += 64) {   // prior_net.0 + ReLU + Dropout: four threads per hidden unit
    const int h = h0 + (tid >> 2), part = tid & 3;
    float s = 0.f;
    if (h < Hd) {
      const float4* w = reinterpret_cast<const float4*>(a.w0 + (int64_t)h * 2 * E + part * 64);
      const float4* x = reinterpret_cast<const float4*>(pooled + part * 64);
      float4 wr[16];
#pragma unroll
      for (int q = 0; q < 16; ++q) wr[q] = __ldg(w + q);
#pragma unroll
      for (int q = 0; q < 16; ++q) s += dot4(wr[q], x[q]);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (h < Hd && part == 0) {
      float r = fmaxf(s + a.b0[h], 0.f);
      r *= drop_factor(a.p_drop, a.seed, a.offset, RS_PRIOR, (uint64_t)b * Hd + h);
      hid[h] = r;
      a.phid[(int64_t)b * Hd + h] = r;
    }
  }
  __syncthreads();
  for (int p = warp; p < P; p += nw) {
    const float* w = a.w3 + (int64_t)p * Hd;
    float s = 0.f;
    for (int k = lane; k < Hd; k += 32) s = fmaf(w[k], hid[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      prior[p] = s + a.b3[p];
      a.prior[(int64_t)b * P + p] = s + a.b3[p];
    }
  }
  __syncthreads();
  if (tid < 4) {
    const float* w = a.wh[tid] + E;
    float s = a.bh[tid][0];
    for (int p = 0; p < P; ++p) s = fmaf(w[p], prior[p], s);
    pc[tid] = s;
  }
  __syncthreads();
  for (int i = tid; i < T; i += blockDim.x) {
    a.qv[(int64_t)b * T + i] = rawA[i] + pc[0];
    a.kvs[(int64_t)b * T + i] = rawB[i] + pc[3];
  }
  for (int j = tid; j < Ta; j += blockDim.x) {
    a.ka[(int64_t)b * Ta + j] = rawA[T + j] + pc[1];
    a.qa[(int64_t)b * Ta + j] = rawB[T + j] + pc[2];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Attention core of nn.MultiheadAttention (need_weights branch of F.multi_head_attention_forward):
//   S_h = (Q_h / sqrt(dh)) K_h^T + bias ;  P_h = softmax(S_h) ;  O_h = dropout(P_h) V_h
// with bias[l,s] = tanh(qs[l] + ks[s]) * bias_scale computed on the fly (fusion.py:175-176; identical for
// every head, :351-354 — the reference materialises it H times).  One CTA per clip, heads in sequence.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) attn_fwd_kernel(AttnArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, S = a.S, E = a.E, H = a.H, dh = E / H, dp = dh + 1;
  float* Qs = sm;                 // [L][dp]
  float* Ks = Qs + L * dp;        // [S][dp]
  float* Vs = Ks + S * dp;        // [S][dp]
  float* prow = Vs + S * dp;      // [nw][S]
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float bscale = a.bias_scale ? a.bias_scale[0] : 0.f;
  const float* qs = a.qs ? a.qs + (int64_t)b * L : nullptr;
  const float* ks = a.ks ? a.ks + (int64_t)b * S : nullptr;
  for (int h = 0; h < H; ++h) {
    for (int i = tid; i < L * dh; i += blockDim.x) {
      const int l = i / dh, d = i - l * dh;
      Qs[l * dp + d] = q[(int64_t)l * E + h * dh + d] * a.scale;   // q scaled before the product, as torch does
    }
    for (int i = tid; i < S * dh; i += blockDim.x) {
      const int s = i / dh, d = i - s * dh;
      Ks[s * dp + d] = kv[(int64_t)s * 2 * E + h * dh + d];
      Vs[s * dp + d] = kv[(int64_t)s * 2 * E + E + h * dh + d];
    }
    __syncthreads();
    float* pr = prow + warp * S;
    for (int l = warp; l < L; l += nw) {
      float mx = -INFINITY;
      for (int s = lane; s < S; s += 32) {
        float acc = 0.f;
        for (int d = 0; d < dh; ++d) acc = fmaf(Qs[l * dp + d], Ks[s * dp + d], acc);
        if (qs) acc += tanhf(qs[l] + ks[s]) * bscale;
        pr[s] = acc;
        mx = fmaxf(mx, acc);
      }
      mx = warp_max(mx);
      float sum = 0.f;
      for (int s = lane; s < S; s += 32) {
        const float e = expf(pr[s] - mx);
        pr[s] = e;
        sum += e;
      }
      sum = warp_sum(sum);
      const float inv = 1.f / sum;
      const int64_t prow_idx = (((int64_t)b * H + h) * L + l) * S;
      for (int s = lane; s < S; s += 32) {
        float p = pr[s] * inv;
        if (a.P) a.P[prow_idx + s] = p;
        if (a.p_drop > 0.f) p *= drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)(prow_idx + s));
        pr[s] = p;
      }
      __syncwarp();
      for (int d = lane; d < dh; d += 32) {
        float acc = 0.f;
        for (int s = 0; s < S; ++s) acc = fmaf(pr[s], Vs[s * dp + d], acc);
        a.O[((int64_t)b * L + l) * E + h * dh + d] = acc;
      }
      __syncwarp();
    }
    __syncthreads();
  }
}


// The same for dh == 32 (one lane per head column) with register-tiled products (smem_tile.cuh): scores of the whole head go to
// shared memory as [L][SP] (SP = S rounded up to 4, padding zeroed), one warp normalises a row, then O = Pd·V.
__global__ void __launch_bounds__(256, 4) attn_fwd32_kernel(AttnArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, S = a.S, E = a.E, H = a.H, SP = (S + 3) & ~3;
  constexpr int dh = 32, dp = 33;
  float* Sc = sm;                 // [L][SP]
  float* Qs = Sc + L * SP;        // [L][dp]
  float* Ks = Qs + L * dp;        // [S][dp]
  float* Vs = Ks + S * dp;        // [S][dp]
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float bscale = a.bias_scale ? a.bias_scale[0] : 0.f;
  const float* qs = a.qs ? a.qs + (int64_t)b * L : nullptr;
  const float* ks = a.ks ? a.ks + (int64_t)b * S : nullptr;
  const bool lanes_on_s = S >= L;   // the longer side runs along the lanes
  const int AR = lanes_on_s ? L : S, BR = lanes_on_s ? S : L;
  const float* Am = lanes_on_s ? Qs : Ks;
  const float* Bm = lanes_on_s ? Ks : Qs;
  const int tiles_a = (AR + 1) / 2, ntile = tiles_a * ((BR + 159) / 160);
  const bool short_rows = (S & 3) == 0 && S <= 32;
  const bool p_vec = (reinterpret_cast<uintptr_t>(a.P) & 15) == 0;
  for (int h = 0; h < H; ++h) {
    stage_rows32(Qs, q + h * dh, L, E, a.scale, tid, blockDim.x);   // q scaled before the product, as torch does
    stage_rows32x2(Ks, Vs, kv + h * dh, E, S, 2 * E, tid, blockDim.x);
    __syncthreads();
    for (int t = warp; t < ntile; t += nw) {
      const int a0 = (t % tiles_a) * 2, b0 = (t / tiles_a) * 160;
      float acc[2][5];
      tile_nt<2, 5>(Am, a0, AR, dp, Bm, b0, BR, dp, dh, lane, acc);
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 5; ++j) {
          const int ar = a0 + i, br = b0 + lane + 32 * j;
          if (ar < AR && br < BR) {
            const int l = lanes_on_s ? ar : br, s2 = lanes_on_s ? br : ar;
            float v = acc[i][j];
            if (qs) v += tanhf(qs[l] + ks[s2]) * bscale;
            Sc[l * SP + s2] = v;
          }
        }
    }
    __syncthreads();
    if (short_rows) {
      // rows of <= 32 scores (a2v: 16): lpr lanes per row with four scores each, 32/lpr rows per warp, one Philox evaluation
      // per lane (a row's first element index is a multiple of 4) — a warp per row left half the lanes idle and spent one
      // Philox evaluation and ten shuffles on every 16 scores.
      const int lpr = S <= 16 ? 4 : 8, rpw = 32 / lpr, sub = lane / lpr, sl = lane - sub * lpr;
      const bool act = 4 * sl < S;
      for (int l0 = warp * rpw; l0 < L; l0 += nw * rpw) {
        const int l = min(l0 + sub, L - 1);
        const bool ok = act && l0 + sub < L;
        float4* pr4 = reinterpret_cast<float4*>(Sc + l * SP) + sl;
        float4 x = act ? *pr4 : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        float mx = fmaxf(fmaxf(x.x, x.y), fmaxf(x.z, x.w));
        for (int o = lpr >> 1; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        x.x = expf(x.x - mx); x.y = expf(x.y - mx); x.z = expf(x.z - mx); x.w = expf(x.w - mx);
        float sum = (x.x + x.y) + (x.z + x.w);
        for (int o = lpr >> 1; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const float inv = 1.f / sum;
        x.x *= inv; x.y *= inv; x.z *= inv; x.w *= inv;
        const int64_t idx0 = (((int64_t)b * H + h) * L + l) * S + 4 * sl;
        if (ok && a.P) {
          if (p_vec) *reinterpret_cast<float4*>(a.P + idx0) = x;
          else { a.P[idx0] = x.x; a.P[idx0 + 1] = x.y; a.P[idx0 + 2] = x.z; a.P[idx0 + 3] = x.w; }
        }
        if (a.p_drop > 0.f) {
          float f[4];
          drop_factor4(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)idx0, f);
          x.x *= f[0]; x.y *= f[1]; x.z *= f[2]; x.w *= f[3];
        }
        if (ok) *pr4 = x;
      }
    } else
    for (int l = warp; l < L; l += nw) {
      float* pr = Sc + l * SP;
      float mx = -INFINITY;
      for (int s2 = lane; s2 < S; s2 += 32) mx = fmaxf(mx, pr[s2]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int s2 = lane; s2 < S; s2 += 32) {
        const float e = expf(pr[s2] - mx);
        pr[s2] = e;
        sum += e;
      }
      sum = warp_sum(sum);
      const float inv = 1.f / sum;
      const int64_t prow_idx = (((int64_t)b * H + h) * L + l) * S;
      for (int s2 = lane; s2 < SP; s2 += 32) {
        float p = 0.f;
        if (s2 < S) {
          p = pr[s2] * inv;
          if (a.P) a.P[prow_idx + s2] = p;
          if (a.p_drop > 0.f) p *= drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)(prow_idx + s2));
        }
        pr[s2] = p;   // (padding columns: zero, the product below reads them)
      }
    }
    __syncthreads();
    if (L > 32) {
      for (int t = warp; t < (L + 3) / 4; t += nw) {
        float o[4];
        tile_ak<4>(Sc, 4 * t, L, SP, Vs, dp, S, lane, o);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (4 * t + i < L) a.O[((int64_t)b * L + 4 * t + i) * E + h * dh + lane] = o[i];
      }
    } else {
      for (int t = warp; t < (L + 1) / 2; t += nw) {
        float o[2];
        tile_ak<2>(Sc, 2 * t, L, SP, Vs, dp, S, lane, o);
#pragma unroll
        for (int i = 0; i < 2; ++i)
          if (2 * t + i < L) a.O[((int64_t)b * L + 2 * t + i) * E + h * dh + lane] = o[i];
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------------------------------------
// s = x + drop_path(y) ; out = LayerNorm(s) (fusion.py:395 / :399).  One warp per token row.  `y` is
// overwritten with s (the backward needs it) and mean / rstd are saved.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) residual_ln_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              float* __restrict__ out, float* __restrict__ mean_out,
                                                              float* __restrict__ rstd_out, float* __restrict__ dpf_out,
                                                              int rows, int L, int E, float p_path, uint64_t seed,
                                                              uint64_t offset, uint32_t stream) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int b = row / L;
  const float f = drop_factor(p_path, seed, offset, stream, (uint64_t)b);
  if (dpf_out && lane == 0 && row % L == 0) dpf_out[b] = f;
  const float* xr = x + (int64_t)row * E;
  float* yr = y + (int64_t)row * E;
  float sum = 0.f;
  for (int e = lane; e < E; e += 32) {
    const float s = xr[e] + f * yr[e];
    yr[e] = s;
    sum += s;
  }
  const float mu = warp_sum(sum) / (float)E;
  float var = 0.f;
  for (int e = lane; e < E; e += 32) {
    const float d = yr[e] - mu;
    var = fmaf(d, d, var);
  }
  const float rstd = rsqrtf(warp_sum(var) / (float)E + 1e-5f);
  for (int e = lane; e < E; e += 32) out[(int64_t)row * E + e] = (yr[e] - mu) * rstd * gamma[e] + beta[e];
  if (lane == 0) {
    mean_out[row] = mu;
    rstd_out[row] = rstd;
  }
}

// E == 128: one float4 per lane covers the row and stays in registers across the three passes (the generic kernel re-reads the row
// it has just written); two rows per warp so that four 16-byte loads are in flight before the first reduction.
__global__ void __launch_bounds__(256) residual_ln_fwd128_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 float* __restrict__ out, float* __restrict__ mean_out,
                                                                 float* __restrict__ rstd_out, float* __restrict__ dpf_out,
                                                                 int rows, int L, float p_path, uint64_t seed, uint64_t offset,
                                                                 uint32_t stream) {
  constexpr int E = 128;
  const int row0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 2, lane = threadIdx.x & 31;
  if (row0 >= rows) return;
  const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma) + lane), b4 = __ldg(reinterpret_cast<const float4*>(beta) + lane);
  float4 s[2];
  bool ok[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int row = row0 + i;
    ok[i] = row < rows;
    const int rr = ok[i] ? row : rows - 1, b = rr / L;
    const float f = drop_factor(p_path, seed, offset, stream, (uint64_t)b);
    if (dpf_out && ok[i] && lane == 0 && rr % L == 0) dpf_out[b] = f;
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x + (int64_t)rr * E) + lane);
    const float4 yv = reinterpret_cast<const float4*>(y + (int64_t)rr * E)[lane];
    s[i] = make_float4(fmaf(f, yv.x, xv.x), fmaf(f, yv.y, xv.y), fmaf(f, yv.z, xv.z), fmaf(f, yv.w, xv.w));
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float mu = warp_sum((s[i].x + s[i].y) + (s[i].z + s[i].w)) * (1.f / (float)E);
    const float dx = s[i].x - mu, dy = s[i].y - mu, dz = s[i].z - mu, dw = s[i].w - mu;
    const float rstd = rsqrtf(warp_sum(fmaf(dx, dx, fmaf(dy, dy, fmaf(dz, dz, dw * dw)))) * (1.f / (float)E) + 1e-5f);
    if (ok[i]) {
      const int64_t o = (int64_t)(row0 + i) * E;
      reinterpret_cast<float4*>(y + o)[lane] = s[i];
      reinterpret_cast<float4*>(out + o)[lane] = make_float4(fmaf(dx * rstd, g4.x, b4.x), fmaf(dy * rstd, g4.y, b4.y),
                                                             fmaf(dz * rstd, g4.z, b4.z), fmaf(dw * rstd, g4.w, b4.w));
      if (lane == 0) {
        mean_out[row0 + i] = mu;
        rstd_out[row0 + i] = rstd;
      }
    }
  }
}
static void launch_residual_ln(cudaStream_t st, const float* x, float* y, const float* gamma, const float* beta, float* out, float* mean_out,
                               float* rstd_out, float* dpf_out, int rows, int L, int E, float p_path, uint64_t seed, uint64_t offset,
                               uint32_t stream) {
  if (E == 128)
    residual_ln_fwd128_kernel<<<(rows + 15) / 16, 256, 0, st>>>(x, y, gamma, beta, out, mean_out, rstd_out, dpf_out, rows, L, p_path, seed,
                                                                offset, stream);
  else
    residual_ln_fwd_kernel<<<(rows + 7) / 8, 256, 0, st>>>(x, y, gamma, beta, out, mean_out, rstd_out, dpf_out, rows, L, E, p_path, seed,
                                                           offset, stream);
  count_launch();
}

// ---------------------------------------------------------------------------------------------------------
// Temporal mean pooling (temporal.py:108-109) + classifier head (fusion.py:405-411).  One CTA per clip.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pool_head_fwd_kernel(HeadArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int E = a.E, T = a.T, Ta = a.Ta, HW = a.HW, C = a.C;
  float* pooled = sm;            // [2E]
  float* hid = pooled + 2 * E;   // [HW]
  float* fused = hid + HW;       // [E] (gated)
  float* scratch = fused + E;    // [nw][E]
  __shared__ float s_gate;
  const float* v1 = a.v1 + (int64_t)b * T * E;
  const float* a1 = a.a1 + (int64_t)b * Ta * E;
  // stage 0: everything in this launch.  Large batches run it as three launches instead — stage 1: the means only (-> a.pooled);
  // the hidden Linear + ReLU as a GEMM over the batch (tensor cores under PREC_TF32; per clip it re-read W0 from L2 and spent
  // five shuffles per output); stage 2: dropout on a.hhid (holding relu(pre-activation)) and everything after it.
  if (a.stage == 2) {
    for (int e = tid; e < 2 * E; e += blockDim.x) pooled[e] = a.pooled[(int64_t)b * 2 * E + e];
    for (int h = tid; h < HW; h += blockDim.x) {
      float r = a.hhid[(int64_t)b * HW + h];
      r *= drop_factor(a.p_drop, a.seed, a.offset, RS_HEAD, (uint64_t)b * HW + h);
      hid[h] = r;
      a.hhid[(int64_t)b * HW + h] = r;
    }
  } else {
    clip_mean(v1, T, E, pooled, scratch, tid, blockDim.x);
    clip_mean(a1, Ta, E, pooled + E, scratch, tid, blockDim.x);
    if (a.pooled)
      for (int e = tid; e < 2 * E; e += blockDim.x) a.pooled[(int64_t)b * 2 * E + e] = pooled[e];
    if (a.stage == 1) return;
  }
  for (int h = warp; a.stage == 0 && h < HW; h += nw) {
    const float* w = a.w0 + (int64_t)h * 2 * E;
    float s = 0.f;
    for (int k = lane; k < 2 * E; k += 32) s = fmaf(w[k], pooled[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      float r = fmaxf(s + a.b0[h], 0.f);
      r *= drop_factor(a.p_drop, a.seed, a.offset, RS_HEAD, (uint64_t)b * HW + h);
      hid[h] = r;
      if (a.hhid) a.hhid[(int64_t)b * HW + h] = r;
    }
  }
  __syncthreads();
  if (!a.gated) {
    for (int c = warp; c < C; c += nw) {
      const float* w = a.w1 + (int64_t)c * HW;
      float s = 0.f;
      for (int k = lane; k < HW; k += 32) s = fmaf(w[k], hid[k], s);
      s = warp_sum(s);
      if (lane == 0) a.logits[(int64_t)b * C + c] = s + a.b1[c];
    }
    return;
  }
  if (warp == 0) {
    float s = 0.f;
    for (int k = lane; k < HW; k += 32) s = fmaf(a.w1[k], hid[k], s);
    s = warp_sum(s);
    if (lane == 0) {
      const float g = 1.f / (1.f + expf(-(s + a.b1[0])));
      s_gate = g;
      if (a.gate) a.gate[b] = g;
    }
  }
  __syncthreads();
  const float g = s_gate;
  for (int e = tid; e < E; e += blockDim.x) {
    const float f = g * pooled[e] + (1.f - g) * pooled[E + e];
    fused[e] = f;
    if (a.fusedg) a.fusedg[(int64_t)b * E + e] = f;
  }
  __syncthreads();
  for (int c = warp; c < C; c += nw) {
    const float* w = a.wc + (int64_t)c * E;
    float s = 0.f;
    for (int k = lane; k < E; k += 32) s = fmaf(w[k], fused[k], s);
    s = warp_sum(s);
    if (lane == 0) a.logits[(int64_t)b * C + c] = s + a.bc[c];
  }
}


// ---------------------------------------------------------------------------------------------------------
// TemporalAttentionPooling (temporal.py:9-26): score_l = w2·GELU(W1·LN(x_l) + b1) + b2, attn = softmax_l(score),
// out = sum_l attn_l · x_l (the UN-normalised tokens).  One CTA per clip; W1 is staged transposed in shared memory
// ([e][j], so lane j walks a column without bank conflicts), one warp per token row.  The Dropout of the score MLP
// (temporal.py:19) multiplies GELU's output by a Philox mask indexed (clip, token, hidden unit); attn_pool_bwd_kernel regenerates it.
// ---------------------------------------------------------------------------------------------------------
// TIN = float: the fp32 family's token tensors.  TIN = bf16 (the tensor-core inference path): `x` holds the tokens as that path
// leaves them — clips `clip_stride` elements apart, optionally still without the affine of the LayerNorm that produced them
// (pre_g / pre_b: token = pre_g ⊙ x + pre_b; nullptr = identity) — and `out` is a slice of the [B, 2E] pooled buffer (ld_out).
template <typename TIN>
__global__ void __launch_bounds__(256) attn_pool_kernel(const TIN* __restrict__ x, int64_t clip_stride, const float* __restrict__ pre_g,
                                                        const float* __restrict__ pre_b, const float* __restrict__ ln_w,
                                                        const float* __restrict__ ln_b, const float* __restrict__ w1,
                                                        const float* __restrict__ b1, const float* __restrict__ w2,
                                                        const float* __restrict__ b2, float* __restrict__ out, int64_t ld_out, int L,
                                                        int E, float p_drop, uint64_t seed, uint64_t offset, uint32_t rng_stream) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int H2 = E / 2 > 0 ? E / 2 : 1;
  float* w1t = sm;                 // [E][H2]
  float* yrow = w1t + E * H2;      // [nw][E]
  float* score = yrow + nw * E;    // [L]
  __shared__ float s_stat[2];
  for (int i = tid; i < E * H2; i += blockDim.x) {
    const int j = i / E, e = i - j * E;
    w1t[e * H2 + j] = w1[i];
  }
  __syncthreads();
  const TIN* xb = x + (int64_t)b * clip_stride;
  auto tokv = [&](const TIN* row, int e) -> float {
    const float v = to_f32(row[e]);
    return pre_g != nullptr ? fmaf(pre_g[e], v, pre_b[e]) : v;
  };
  float* y = yrow + warp * E;
  for (int l = warp; l < L; l += nw) {
    const TIN* xr = xb + (int64_t)l * E;
    float sum = 0.f;
    for (int e = lane; e < E; e += 32) sum += tokv(xr, e);
    const float mu = warp_sum(sum) / (float)E;
    float var = 0.f;
    for (int e = lane; e < E; e += 32) {
      const float d = tokv(xr, e) - mu;
      var = fmaf(d, d, var);
    }
    const float rstd = rsqrtf(warp_sum(var) / (float)E + 1e-5f);
    for (int e = lane; e < E; e += 32) y[e] = (tokv(xr, e) - mu) * rstd * ln_w[e] + ln_b[e];
    __syncwarp();
    float sc = 0.f;
    for (int j = lane; j < H2; j += 32) {
      float h = b1[j];
      for (int e = 0; e < E; ++e) h = fmaf(w1t[e * H2 + j], y[e], h);
      float g = 0.5f * h * (1.f + erff(h * 0.70710678118654752f));   // nn.GELU() (exact)
      if (p_drop > 0.f) g *= drop_factor(p_drop, seed, offset, rng_stream, ((uint64_t)b * L + l) * H2 + j);
      sc = fmaf(w2[j], g, sc);
    }
    sc = warp_sum(sc);
    if (lane == 0) score[l] = sc + b2[0];
    __syncwarp();
  }
  __syncthreads();
  if (warp == 0) {   // softmax over time
    float mx = -INFINITY;
    for (int l = lane; l < L; l += 32) mx = fmaxf(mx, score[l]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int l = lane; l < L; l += 32) {
      const float e = expf(score[l] - mx);
      score[l] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) s_stat[0] = 1.f / sum;
  }
  __syncthreads();
  const float inv = s_stat[0];
  for (int e = tid; e < E; e += blockDim.x) {
    float acc = 0.f;
    for (int l = 0; l < L; ++l) acc = fmaf(score[l], tokv(xb + (int64_t)l * E, e), acc);
    out[(int64_t)b * ld_out + e] = acc * inv;
  }
}

template <typename TIN>
static int launch_attn_pool_t(cudaStream_t st, const TIN* x, int64_t clip_stride, const float* pre_g, const float* pre_b,
                              const float* const* p, float* out, int64_t ld_out, int B, int L, int E, float p_drop, uint64_t seed,
                              uint64_t offset, uint32_t rng_stream) {
  const int H2 = E / 2 > 0 ? E / 2 : 1;
  const size_t smem = sizeof(float) * ((size_t)E * H2 + 8 * (size_t)E + (size_t)L);
  if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention pooling needs %zu B of shared memory (> 200 KiB)", smem);
  XF_TRY(ensure_dyn_smem(attn_pool_kernel<TIN>, smem));
  attn_pool_kernel<TIN><<<B, 256, smem, st>>>(x, clip_stride, pre_g, pre_b, p[0], p[1], p[2], p[3], p[4], p[5], out, ld_out, L, E, p_drop,
                                              seed, offset, rng_stream);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
static int launch_attn_pool(cudaStream_t st, const float* x, const float* const* p, float* out, int B, int L, int E, float p_drop,
                            uint64_t seed, uint64_t offset, uint32_t rng_stream) {
  return launch_attn_pool_t<float>(st, x, (int64_t)L * E, nullptr, nullptr, p, out, E, B, L, E, p_drop, seed, offset, rng_stream);
}
// The learned pooling over bf16 token rows of the tensor-core inference path (fwd_bf16.cu), eval mode
int launch_attn_pool_bf16(cudaStream_t st, const void* x_bf16, int64_t clip_stride, const float* pre_g, const float* pre_b,
                          const float* const* p, float* out, int64_t ld_out, int B, int L, int E) {
  return launch_attn_pool_t<__nv_bfloat16>(st, reinterpret_cast<const __nv_bfloat16*>(x_bf16), clip_stride, pre_g, pre_b, p, out, ld_out,
                                           B, L, E, 0.f, 0, 0, 0);
}

// ---------------------------------------------------------------------------------------------------------
// TemporalTransformerPooling (temporal.py:46-75): x + sinusoidal positions -> pool_layers pre-norm nn.TransformerEncoderLayers
// (x += drop(SA(LN1(x))); x += drop(W2·drop(GELU(W1·LN2(x))))) -> TemporalAttentionPooling.  The Linears go through linear_fwd
// (tensor cores under PREC_TF32), the self-attention through the attention kernels of the fusion block (no bias; attention
// dropout p_pool when training); what is left are the row-wise pieces below.  With training / save_for_backward every layer
// keeps its activations in the workspace (TxSaved) for transformer_pool_bwd (bwd_f32.cu).
// ---------------------------------------------------------------------------------------------------------
// out[row, e] = x[row, e] + pe[row % L, e],  pe[l, 2i] = sin(l·w_i), pe[l, 2i+1] = cos(l·w_i), w_i = exp(2i·(-ln 10000 / E))  (temporal.py:33-38)
__global__ void posenc_add_kernel(const float* __restrict__ x, float* __restrict__ out, int64_t n, int L, int E) {
  const float c = -9.210340371976184f / (float)(E > 1 ? E : 1);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int e = (int)(i % E), l = (int)((i / E) % L);
    const float w = expf((float)(e & ~1) * c), arg = (float)l * w;
    out[i] = x[i] + ((e & 1) ? cosf(arg) : sinf(arg));
  }
}
// y = LayerNorm(x)·gamma + beta, one warp per row
__global__ void __launch_bounds__(256) ln_rows_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                      const float* __restrict__ beta, float* __restrict__ y, int rows, int E) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* xr = x + (int64_t)row * E;
  float sum = 0.f;
  for (int e = lane; e < E; e += 32) sum += xr[e];
  const float mu = warp_sum(sum) / (float)E;
  float var = 0.f;
  for (int e = lane; e < E; e += 32) {
    const float d = xr[e] - mu;
    var = fmaf(d, d, var);
  }
  const float rstd = rsqrtf(warp_sum(var) / (float)E + 1e-5f);
  for (int e = lane; e < E; e += 32) y[(int64_t)row * E + e] = (xr[e] - mu) * rstd * gamma[e] + beta[e];
}
int launch_ln_rows(cudaStream_t st, const float* x, const float* gamma, const float* beta, float* y, int rows, int E) {
  ln_rows_kernel<<<(rows + 7) / 8, 256, 0, st>>>(x, gamma, beta, y, rows, E);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
// act = Dropout(GELU(pre)): nn.GELU() (exact) and the FFN's inner dropout (TransformerEncoderLayer._ff_block); act may be pre
__global__ void gelu_drop_kernel(const float* pre, float* act, int64_t n, float p, uint64_t seed, uint64_t offset, uint32_t stream) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float h = pre[i];
    act[i] = 0.5f * h * (1.f + erff(h * 0.70710678118654752f)) * drop_factor(p, seed, offset, stream, (uint64_t)i);
  }
}
// out = x + Dropout(y): the residual around a branch (dropout1 / dropout2 of the encoder layer); out may be x
__global__ void add_drop_kernel(float* out, const float* x, const float* __restrict__ y, int64_t n, float p, uint64_t seed,
                                uint64_t offset, uint32_t stream) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = fmaf(y[i], drop_factor(p, seed, offset, stream, (uint64_t)i), x[i]);
}
static inline int ew_grid(int64_t n) {
  const int64_t g = (n + 255) / 256, cap = (int64_t)num_sms() * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}

// ---------------------------------------------------------------------------------------------------------
// launch helpers shared with bwd_f32.cu
// ---------------------------------------------------------------------------------------------------------
size_t attn_smem_bytes(int L, int S, int dh, int nwarps) { return sizeof(float) * ((size_t)(L + 2 * S) * (dh + 1) + (size_t)nwarps * S); }

int launch_attn_fwd(cudaStream_t st, const AttnArgs& a, int B) {
  static const bool simt_env = getenv("XATTN_ATTN_SIMT") != nullptr;   // (comparison runs: keep the fp32 SIMT attention under tf32)
  if (g_tc_gemm && !simt_env && attn_mma_ok(a.L, a.S, a.E, a.H)) return launch_attn_fwd_mma(st, a, B);
  if (a.E / a.H == 32) {   // one lane per head column: the register-tiled kernel
    const size_t smem = sizeof(float) * ((size_t)a.L * ((a.S + 3) & ~3) + (size_t)(a.L + 2 * a.S) * 33);
    if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention tile needs %zu B of shared memory (> 200 KiB)", smem);
    XF_TRY(ensure_dyn_smem(attn_fwd32_kernel, smem));
    attn_fwd32_kernel<<<B, 256, smem, st>>>(a);
    count_launch();
    XF_LAUNCH_CHECK();
    return XATTN_OK;
  }
  const size_t smem = attn_smem_bytes(a.L, a.S, a.E / a.H, 8);
  if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention tile needs %zu B of shared memory (> 200 KiB)", smem);
  XF_TRY(ensure_dyn_smem(attn_fwd_kernel, smem));
  attn_fwd_kernel<<<B, 256, smem, st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

template <typename TIO>
static int forward_f32_impl(const xattn_config& c, const float* params, const TIO* xv, const TIO* xa, float* logits, float* ws,
                            cudaStream_t st) {
  ParamLayout pl;
  compute_param_layout(c, pl);
  WsF32 w;
  compute_ws_f32(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  const int B = c.B, T = c.T, Ta = c.Ta, E = c.E, H = c.H;
  const bool train = c.training != 0;
  const bool save = train || c.save_for_backward;
  const float p_attn = train ? c.p_attn : 0.f, p_path = train ? c.p_path : 0.f, p_head = train ? c.p_head : 0.f,
              p_prior = train ? c.p_prior : 0.f;

  NvtxRange nvtx_fwd(g_tc_gemm ? "xattn_forward[tf32]" : "xattn_forward[fp32]");
  // projections (fusion.py:372, :378, :386)
  if (g_tc_gemm && std::is_same<TIO, __nv_bfloat16>::value && E % 128 == 0 && c.Dv % 64 == 0 && c.Da % 64 == 0) {
    // bf16 embeddings: kind::f16 products against bf16-rounded copies of the two weights (gemm_tc.cu), fp32 out
    __nv_bfloat16* wv = reinterpret_cast<__nv_bfloat16*>(ws + w.wbf);
    __nv_bfloat16* wa = wv + (int64_t)E * c.Dv;
    cvt_f32_to_bf16(st, P(XP_V_IN_W), wv, (int64_t)E * c.Dv);
    cvt_f32_to_bf16(st, P(XP_SEQ_W), wa, (int64_t)E * c.Da);
    XF_TRY(gemm_tc_bf16(st, xv, wv, P(XP_V_IN_B), ws + w.v0, false, E, B * T, E, c.Dv));
    XF_TRY(gemm_tc_bf16(st, xa, wa, P(XP_SEQ_B), ws + w.aseq, false, E, B * Ta, E, c.Da));
  } else {
    linear_fwd<TIO>(st, xv, P(XP_V_IN_W), P(XP_V_IN_B), ws + w.v0, B * T, E, c.Dv);
    linear_fwd<TIO>(st, xa, P(XP_SEQ_W), P(XP_SEQ_B), ws + w.aseq, B * Ta, E, c.Da);
  }
  linear_fwd<float>(st, ws + w.aseq, P(XP_A_IN_W), P(XP_A_IN_B), ws + w.a0, B * Ta, E, E);
  XF_LAUNCH_CHECK();

  const float scale = 1.0f / sqrtf((float)(E / H));
  // ---- video queries audio (fusion.py:394) ----
  // (Running the prior kernel beside these two GEMMs on a side stream, as the backward does with its counterpart, was measured:
  // no gain here — the prior CTAs reach the SMs first and hold every register until their first wave retires.)
  linear_fwd<float>(st, ws + w.v0, P(XP_V2A_IN_W), P(XP_V2A_IN_B), ws + w.q1, B * T, E, E);
  linear_fwd<float>(st, ws + w.a0, P(XP_V2A_IN_W) + (int64_t)E * E, P(XP_V2A_IN_B) + E, ws + w.kv1, B * Ta, 2 * E, E);
  if (c.use_prior) {
    PriorArgs pa;
    pa.v0 = ws + w.v0; pa.a0 = ws + w.a0;
    pa.w0 = P(XP_PRIOR0_W); pa.b0 = P(XP_PRIOR0_B); pa.w3 = P(XP_PRIOR3_W); pa.b3 = P(XP_PRIOR3_B);
    pa.wh[0] = P(XP_VQ_W); pa.bh[0] = P(XP_VQ_B); pa.wh[1] = P(XP_AK_W); pa.bh[1] = P(XP_AK_B);
    pa.wh[2] = P(XP_AQ_W); pa.bh[2] = P(XP_AQ_B); pa.wh[3] = P(XP_VK_W); pa.bh[3] = P(XP_VK_B);
    pa.pooled0 = ws + w.pooled0; pa.phid = ws + w.phid; pa.prior = ws + w.prior;
    pa.qv = ws + w.qv; pa.ka = ws + w.ka; pa.qa = ws + w.qa; pa.kvs = ws + w.kvs;
    pa.T = T; pa.Ta = Ta; pa.E = E; pa.P = c.prior_dim; pa.Hd = c.prior_hidden;
    pa.p_drop = p_prior; pa.seed = c.seed; pa.offset = c.offset;
    if (E == 128) {
      const size_t smem128 = sizeof(float) * (2 * E + 2 * 8 * E + 2 * (T + Ta) + c.prior_hidden + c.prior_dim + 4);
      if (smem128 > 48 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "prior kernel needs %zu B of shared memory (T + Ta too large)", smem128);
      prior_fwd128_kernel<<<B, 256, smem128, st>>>(pa);
    } else {
      const size_t smem = sizeof(float) * (2 * E + c.prior_hidden + c.prior_dim + 4 + 8 * E);
      prior_fwd_kernel<<<B, 256, smem, st>>>(pa);
    }
    count_launch();
    XF_LAUNCH_CHECK();
  }

  {
    AttnArgs aa;
    aa.q = ws + w.q1; aa.kv = ws + w.kv1;
    aa.qs = c.use_prior ? ws + w.qv : nullptr; aa.ks = c.use_prior ? ws + w.ka : nullptr;
    aa.bias_scale = c.use_prior ? P(XP_BIAS_SCALE) : nullptr;
    aa.P = save ? ws + w.p1 : nullptr; aa.O = ws + w.o1;
    aa.L = T; aa.S = Ta; aa.E = E; aa.H = H; aa.scale = scale;
    aa.p_drop = p_attn; aa.seed = c.seed; aa.offset = c.offset; aa.rng_stream = RS_ATTN1; aa.ldp = p_row_stride(c, Ta);
    XF_TRY(launch_attn_fwd(st, aa, B));
  }
  linear_fwd<float>(st, ws + w.o1, P(XP_V2A_OUT_W), P(XP_V2A_OUT_B), ws + w.s1, B * T, E, E);
  {
    const int rows = B * T;
    launch_residual_ln(st, ws + w.v0, ws + w.s1, P(XP_VN_W), P(XP_VN_B), ws + w.v1, ws + w.mean1, ws + w.rstd1, ws + w.dp1, rows, T, E,
                       p_path, c.seed, c.offset, RS_PATH1);
    XF_LAUNCH_CHECK();
  }
  // ---- audio queries the UPDATED video (fusion.py:398) ----
  linear_fwd<float>(st, ws + w.a0, P(XP_A2V_IN_W), P(XP_A2V_IN_B), ws + w.q2, B * Ta, E, E);
  linear_fwd<float>(st, ws + w.v1, P(XP_A2V_IN_W) + (int64_t)E * E, P(XP_A2V_IN_B) + E, ws + w.kv2, B * T, 2 * E, E);
  {
    AttnArgs aa;
    aa.q = ws + w.q2; aa.kv = ws + w.kv2;
    aa.qs = c.use_prior ? ws + w.qa : nullptr; aa.ks = c.use_prior ? ws + w.kvs : nullptr;
    aa.bias_scale = c.use_prior ? P(XP_BIAS_SCALE) : nullptr;
    aa.P = save ? ws + w.p2 : nullptr; aa.O = ws + w.o2;
    aa.L = Ta; aa.S = T; aa.E = E; aa.H = H; aa.scale = scale;
    aa.p_drop = p_attn; aa.seed = c.seed; aa.offset = c.offset; aa.rng_stream = RS_ATTN2; aa.ldp = p_row_stride(c, T);
    XF_TRY(launch_attn_fwd(st, aa, B));
  }
  linear_fwd<float>(st, ws + w.o2, P(XP_A2V_OUT_W), P(XP_A2V_OUT_B), ws + w.s2, B * Ta, E, E);
  {
    const int rows = B * Ta;
    launch_residual_ln(st, ws + w.a0, ws + w.s2, P(XP_AN_W), P(XP_AN_B), ws + w.a1, ws + w.mean2, ws + w.rstd2, ws + w.dp2, rows, Ta, E,
                       p_path, c.seed, c.offset, RS_PATH2);
    XF_LAUNCH_CHECK();
  }
  // ---- pool + head (fusion.py:401-411) ----
  {
    HeadArgs ha;
    ha.v1 = ws + w.v1; ha.a1 = ws + w.a1;
    int hT = T, hTa = Ta;
    if (c.pool_mode == XATTN_POOL_TRANSFORMER) {   // encoder layers over each side's tokens, then the learned pooling
      const int F = 4 * E, PH = c.pool_heads, layers = pool_layers(c);
      const float p_tx = train ? c.p_pool : 0.f;
      const float* pv[6] = {P(XP_VPOOL_LN_W), P(XP_VPOOL_LN_B), P(XP_VPOOL_W1), P(XP_VPOOL_B1), P(XP_VPOOL_W2), P(XP_VPOOL_B2)};
      const float* pa[6] = {P(XP_APOOL_LN_W), P(XP_APOOL_LN_B), P(XP_APOOL_W1), P(XP_APOOL_B1), P(XP_APOOL_W2), P(XP_APOOL_B2)};
      float *th = ws + w.tx_h, *tmp = ws + w.tx_q;
      for (int side = 0; side < 2; ++side) {
        const int L = side == 0 ? T : Ta;
        const int M = B * L;
        const int64_t n = (int64_t)M * E;
        // without a backward everything lives in one scratch set and x is updated in place
        TxSaved sv;
        sv.x_in = sv.x_mid = ws + w.tx_x; sv.q = ws + w.tx_q; sv.kv = ws + w.tx_kv; sv.o = th; sv.f_pre = sv.f_act = ws + w.tx_f;
        sv.P = nullptr;
        float* x_out = ws + w.tx_x;
        if (save) sv = tx_saved(ws, w, c, side, 0);
        posenc_add_kernel<<<ew_grid(n), 256, 0, st>>>(side == 0 ? ws + w.v1 : ws + w.a1, sv.x_in, n, L, E);   // temporal.py:72
        count_launch();
        for (int layer = 0; layer < layers; ++layer) {
          auto TP = [&](int k) { return P(XATTN_TX_PARAM_ID(side, layer, k)); };
          if (save) {
            sv = tx_saved(ws, w, c, side, layer);
            x_out = layer + 1 < layers ? tx_saved(ws, w, c, side, layer + 1).x_in : ws + w.tx_out[side];
          }
          ln_rows_kernel<<<(M + 7) / 8, 256, 0, st>>>(sv.x_in, TP(8), TP(9), th, M, E);                          // norm1 (norm_first)
          count_launch();
          linear_fwd<float>(st, th, TP(0), TP(1), sv.q, M, E, E);
          linear_fwd<float>(st, th, TP(0) + (int64_t)E * E, TP(1) + E, sv.kv, M, 2 * E, E);
          AttnArgs aa;
          aa.q = sv.q; aa.kv = sv.kv; aa.qs = nullptr; aa.ks = nullptr; aa.bias_scale = nullptr; aa.P = sv.P; aa.O = sv.o;
          aa.L = L; aa.S = L; aa.E = E; aa.H = PH; aa.scale = 1.0f / sqrtf((float)(E / PH)); aa.p_drop = p_tx; aa.seed = c.seed;
          aa.offset = c.offset; aa.rng_stream = rs_tx(side, layer, 0);
          XF_TRY(launch_attn_fwd(st, aa, B));
          linear_fwd<float>(st, sv.o, TP(2), TP(3), tmp, M, E, E);                                               // out_proj (q is dead)
          add_drop_kernel<<<ew_grid(n), 256, 0, st>>>(sv.x_mid, sv.x_in, tmp, n, p_tx, c.seed, c.offset, rs_tx(side, layer, 1));
          ln_rows_kernel<<<(M + 7) / 8, 256, 0, st>>>(sv.x_mid, TP(10), TP(11), th, M, E);                       // norm2
          count_launch(2);
          linear_fwd<float>(st, th, TP(4), TP(5), sv.f_pre, M, F, E);
          gelu_drop_kernel<<<ew_grid((int64_t)M * F), 256, 0, st>>>(sv.f_pre, sv.f_act, (int64_t)M * F, p_tx, c.seed, c.offset,
                                                                     rs_tx(side, layer, 2));
          linear_fwd<float>(st, sv.f_act, TP(6), TP(7), tmp, M, E, F);
          add_drop_kernel<<<ew_grid(n), 256, 0, st>>>(x_out, sv.x_mid, tmp, n, p_tx, c.seed, c.offset, rs_tx(side, layer, 3));
          count_launch(2);
          XF_LAUNCH_CHECK();
        }
        XF_TRY(launch_attn_pool(st, x_out, side == 0 ? pv : pa, side == 0 ? ws + w.poolv : ws + w.poola, B, L, E, p_tx, c.seed,
                                c.offset, side == 0 ? RS_POOLV : RS_POOLA));
      }
      ha.v1 = ws + w.poolv; ha.a1 = ws + w.poola;
      hT = hTa = 1;
    }
    if (c.pool_mode == XATTN_POOL_ATTN) {   // learned pooling first; the head kernel then sees one (already pooled) token per clip
      const float* pv[6] = {P(XP_VPOOL_LN_W), P(XP_VPOOL_LN_B), P(XP_VPOOL_W1), P(XP_VPOOL_B1), P(XP_VPOOL_W2), P(XP_VPOOL_B2)};
      const float* pa[6] = {P(XP_APOOL_LN_W), P(XP_APOOL_LN_B), P(XP_APOOL_W1), P(XP_APOOL_B1), P(XP_APOOL_W2), P(XP_APOOL_B2)};
      const float p_pool = train ? c.p_pool : 0.f;
      XF_TRY(launch_attn_pool(st, ws + w.v1, pv, ws + w.poolv, B, T, E, p_pool, c.seed, c.offset, RS_POOLV));
      XF_TRY(launch_attn_pool(st, ws + w.a1, pa, ws + w.poola, B, Ta, E, p_pool, c.seed, c.offset, RS_POOLA));
      ha.v1 = ws + w.poolv; ha.a1 = ws + w.poola;
      hT = hTa = 1;
    }
    ha.w0 = P(XP_HEAD0_W); ha.b0 = P(XP_HEAD0_B); ha.w1 = P(XP_HEAD1_W); ha.b1 = P(XP_HEAD1_B);
    ha.wc = P(XP_CLS_W); ha.bc = P(XP_CLS_B);
    ha.pooled = ws + w.pooled; ha.hhid = ws + w.hhid; ha.gate = ws + w.gate; ha.fusedg = ws + w.fusedg;
    ha.logits = logits;
    ha.T = hT; ha.Ta = hTa; ha.E = E; ha.C = c.C; ha.gated = c.head_kind == XATTN_HEAD_GATED;
    ha.HW = ha.gated ? E : c.common_dim;
    ha.p_drop = p_head; ha.seed = c.seed; ha.offset = c.offset;
    const size_t smem = sizeof(float) * (3 * E + ha.HW + 8 * E);
    // always the three-launch form (a clip's logits must not depend on the batch it arrives in — shards are compared bit for bit)
    ha.stage = 1;
    pool_head_fwd_kernel<<<B, 256, smem, st>>>(ha);
    linear_fwd<float>(st, ws + w.pooled, ha.w0, ha.b0, ws + w.hhid, B, ha.HW, 2 * E, true);
    ha.stage = 2;
    pool_head_fwd_kernel<<<B, 256, smem, st>>>(ha);
    count_launch();
    count_launch();
    XF_LAUNCH_CHECK();
  }
  return XATTN_OK;
}

int forward_f32(const xattn_config& c, const float* params, const void* xv, const void* xa, float* logits, void* ws,
                cudaStream_t st) {
  g_tc_gemm = c.precision == XATTN_PREC_TF32;   // the Linear helpers route to the tensor-core GEMMs for this call
  g_tc_status = XATTN_OK;
  int rc;
  if (c.io_dtype == XATTN_DTYPE_BF16)
    rc = forward_f32_impl<__nv_bfloat16>(c, params, (const __nv_bfloat16*)xv, (const __nv_bfloat16*)xa, logits, (float*)ws, st);
  else
    rc = forward_f32_impl<float>(c, params, (const float*)xv, (const float*)xa, logits, (float*)ws, st);
  g_tc_gemm = false;
  return rc != XATTN_OK ? rc : g_tc_status;
}

}  // namespace xf
