#include <stdlib.h>
// fp32 backward of the xattn fusion head for the Stage-1 step (reference train.py:227: loss.backward()
// with both backbones frozen, train.py:808-812 — so no gradient flows into video_emb / audio_emb).
// Hand-derived reverse of fwd_f32.cu; activations come from the forward's workspace, dropout masks are
// regenerated from the same Philox counters.
#include "gemm_simt.cuh"
#include "smem_tile.cuh"
#include "kernels_f32.cuh"

namespace xf {

// ---- db[N] += column sums ----------------------------------------------------------------------------------
__global__ void colsum_kernel(const float* __restrict__ dY, int64_t ldy, int M, int N, float* __restrict__ db, int rows_per_block) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int r0 = blockIdx.y * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int r = r0;
  for (; r + 8 <= r1; r += 8) {   // eight independent row loads in flight per thread (the plain loop was one round trip per row)
    const float* p = dY + (int64_t)r * ldy + n;
    const float a0 = p[0], a1 = p[ldy], a2 = p[2 * ldy], a3 = p[3 * ldy], a4 = p[4 * ldy], a5 = p[5 * ldy], a6 = p[6 * ldy],
                a7 = p[7 * ldy];
    s0 += a0 + a4; s1 += a1 + a5; s2 += a2 + a6; s3 += a3 + a7;
  }
  for (; r < r1; ++r) s0 += dY[(int64_t)r * ldy + n];
  atomicAdd(db + n, (s0 + s1) + (s2 + s3));
}
// The same with float4 lanes (N % 4 == 0, N <= 1024, rows 16-byte aligned): N/4 threads cover a row, the CTA's 256/(N/4) row
// groups stride over the block's rows with eight 16-byte loads in flight each, partial sums meet in shared memory and one
// thread per column group issues the atomics.  The scalar kernel kept 4 bytes per load in flight and reached a quarter of HBM.
__global__ void __launch_bounds__(256) colsum4_kernel(const float* __restrict__ dY, int64_t ldy, int M, int N, float* __restrict__ db,
                                                      int rows_per_block) {
  __shared__ float4 red[256];
  const int tid = threadIdx.x, nc4 = N >> 2, nry = 256 / nc4, cx = tid % nc4, ry = tid / nc4;
  const int r0 = blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (ry < nry) {
#pragma unroll 8
    for (int r = r0 + ry; r < r1; r += nry) {
      const float4 u = __ldg(reinterpret_cast<const float4*>(dY + (int64_t)r * ldy) + cx);
      s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
    }
  }
  red[tid] = s;
  __syncthreads();
  if (ry == 0) {
    for (int k = 1; k < nry; ++k) {
      const float4 u = red[cx + k * nc4];
      s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
    }
    atomicAdd(db + 4 * cx, s.x);
    atomicAdd(db + 4 * cx + 1, s.y);
    atomicAdd(db + 4 * cx + 2, s.z);
    atomicAdd(db + 4 * cx + 3, s.w);
  }
}
void launch_colsum(cudaStream_t st, const float* dY, int64_t ldy, int M, int N, float* db) {
  if (M <= 0 || N <= 0) return;
  if ((N & 3) == 0 && N <= 1024 && (ldy & 3) == 0 && (reinterpret_cast<uintptr_t>(dY) & 15) == 0) {
    int rpb = (M + 4 * num_sms() - 1) / (4 * num_sms());
    if (rpb < 64) rpb = 64;
    colsum4_kernel<<<(M + rpb - 1) / rpb, 256, 0, st>>>(dY, ldy, M, N, db, rpb);
    count_launch();
    return;
  }
  const int rpb = 256;
  dim3 grid((N + 127) / 128, (M + rpb - 1) / rpb);
  colsum_kernel<<<grid, 128, 0, st>>>(dY, ldy, M, N, db, rpb);
  count_launch();
}

// gb_out[k] += Σ_n gb_in[n]·W[n,k]: the bias gradient of a Linear that feeds another Linear y = x·Wᵀ (+ b) directly, from the
// bias gradient of that one (column sums commute with the product).
// grid (column blocks, 16-row slices of n): coalesced rows of W, partial sums meet in gb_out atomically.
__global__ void bias_chain_kernel(const float* __restrict__ gb_in, const float* __restrict__ W, float* __restrict__ gb_out, int N, int K) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= K) return;
  const int n0 = blockIdx.y * 16, n1 = min(N, n0 + 16);
  float s = 0.f;
#pragma unroll 4
  for (int n = n0; n < n1; ++n) s = fmaf(gb_in[n], W[(int64_t)n * K + k], s);
  atomicAdd(gb_out + k, s);
}

// ---- head: gradient w.r.t. the hidden pre-activation (and the gated blend) ---------------------------------
// concat: d_hh = (dlogits·W1) ⊙ relu'/dropout mask.
// gated : dfused given; dg = <dfused, v̄-ā>; dz = dg·g(1-g); d_pooled = [g·dfused, (1-g)·dfused]; d_hh = dz·w3 ⊙ mask.
__global__ void __launch_bounds__(256) head_bwd_mid_kernel(const float* __restrict__ dlogits, const float* __restrict__ w1,
                                                           const float* __restrict__ hhid, const float* __restrict__ pooled,
                                                           const float* __restrict__ gate, const float* __restrict__ dfused,
                                                           float* __restrict__ d_hh, float* __restrict__ d_pooled,
                                                           float* __restrict__ d_z, int E, int HW, int C, int gated, float inv_keep) {
  const int b = blockIdx.x, tid = threadIdx.x;
  __shared__ float red[32];
  __shared__ float s_dz;
  if (!gated) {
    for (int h = tid; h < HW; h += blockDim.x) {
      float s = 0.f;
      for (int c = 0; c < C; ++c) s = fmaf(dlogits[(int64_t)b * C + c], w1[(int64_t)c * HW + h], s);
      d_hh[(int64_t)b * HW + h] = hhid[(int64_t)b * HW + h] > 0.f ? s * inv_keep : 0.f;
    }
    return;
  }
  const float g = gate[b];
  float part = 0.f;
  for (int e = tid; e < E; e += blockDim.x) {
    const float df = dfused[(int64_t)b * E + e];
    part = fmaf(df, pooled[(int64_t)b * 2 * E + e] - pooled[(int64_t)b * 2 * E + E + e], part);
    d_pooled[(int64_t)b * 2 * E + e] = g * df;
    d_pooled[(int64_t)b * 2 * E + E + e] = (1.f - g) * df;
  }
  part = warp_sum(part);
  if ((tid & 31) == 0) red[tid >> 5] = part;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
    s_dz = s * g * (1.f - g);
    d_z[b] = s_dz;
  }
  __syncthreads();
  const float dz = s_dz;
  for (int h = tid; h < HW; h += blockDim.x)
    d_hh[(int64_t)b * HW + h] = hhid[(int64_t)b * HW + h] > 0.f ? dz * w1[h] * inv_keep : 0.f;
}

// ---- LayerNorm backward -------------------------------------------------------------------------------------
// dy = dy_full (optional) + dy_pool[b]/L (optional; gradient of the temporal mean).  dx -> dx_out,
// dx·dpf[b] -> dbranch (gradient into the drop-path branch).  dgamma/dbeta accumulated atomically.
constexpr int LN_MAXK = 16;  // E <= 512
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ dy_full, const float* __restrict__ dy_pool,
                                                     int64_t pool_ld, float inv_L, const float* __restrict__ s,
                                                     const float* __restrict__ mean, const float* __restrict__ rstd,
                                                     const float* __restrict__ gamma, const float* __restrict__ dpf,
                                                     float* __restrict__ dx_out, float* __restrict__ dbranch,
                                                     float* __restrict__ dgamma, float* __restrict__ dbeta, int rows, int L, int E) {
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  float ag[LN_MAXK], ab[LN_MAXK];
#pragma unroll
  for (int k = 0; k < LN_MAXK; ++k) ag[k] = ab[k] = 0.f;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < rows; row += gridDim.x * wpb) {
    const int b = row / L;
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < LN_MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        float dy = dy_full ? dy_full[(int64_t)row * E + e] : 0.f;
        if (dy_pool) dy += dy_pool[(int64_t)b * pool_ld + e] * inv_L;
        const float xh = (s[(int64_t)row * E + e] - mu) * rs;
        const float dyg = dy * gamma[e];
        s1 += dyg;
        s2 = fmaf(dyg, xh, s2);
        ag[k] = fmaf(dy, xh, ag[k]);
        ab[k] += dy;
      }
    }
    s1 = warp_sum(s1) / (float)E;
    s2 = warp_sum(s2) / (float)E;
    const float f = dpf ? dpf[b] : 1.f;
#pragma unroll
    for (int k = 0; k < LN_MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        float dy = dy_full ? dy_full[(int64_t)row * E + e] : 0.f;
        if (dy_pool) dy += dy_pool[(int64_t)b * pool_ld + e] * inv_L;
        const float xh = (s[(int64_t)row * E + e] - mu) * rs;
        const float dx = rs * (dy * gamma[e] - s1 - xh * s2);
        dx_out[(int64_t)row * E + e] = dx;
        dbranch[(int64_t)row * E + e] = dx * f;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < LN_MAXK; ++k) {
    const int e = lane + 32 * k;
    if (e < E) {
      atomicAdd(dgamma + e, ag[k]);
      atomicAdd(dbeta + e, ab[k]);
    }
  }
}

// E == 128: a lane owns four adjacent columns, so one float4 load per operand covers the row and dy / x stay in registers
// between the two passes; a warp works on two rows at a time (all their loads in flight before the first reduction); the
// dgamma / dbeta partials of the CTA's warps meet in shared memory before the atomics (one per column and CTA).
__global__ void __launch_bounds__(256) ln_bwd128_kernel(const float* __restrict__ dy_full, const float* __restrict__ dy_pool,
                                                        int64_t pool_ld, float inv_L, const float* __restrict__ s,
                                                        const float* __restrict__ mean, const float* __restrict__ rstd,
                                                        const float* __restrict__ gamma, const float* __restrict__ dpf,
                                                        float* __restrict__ dx_out, float* __restrict__ dbranch,
                                                        float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbranch_bias,
                                                        int rows, int L) {
  constexpr int E = 128;
  __shared__ float red[8][3 * E];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float4 g4 = __ldg(reinterpret_cast<const float4*>(gamma) + lane);
  float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = ag, abr = ag;   // dgamma, dbeta, column sums of dbranch
  const int stride = gridDim.x * 8 * 2;
  for (int row0 = (blockIdx.x * 8 + warp) * 2; row0 < rows; row0 += stride) {
    float4 dy[2], xh[2];
    float mu[2], rs[2], f[2];
    bool ok[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int row = row0 + i;
      ok[i] = row < rows;
      const int rr = ok[i] ? row : rows - 1, b = rr / L;
      dy[i] = dy_full ? __ldg(reinterpret_cast<const float4*>(dy_full + (int64_t)rr * E) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (dy_pool) {
        const float4 p = __ldg(reinterpret_cast<const float4*>(dy_pool + (int64_t)b * pool_ld) + lane);
        dy[i].x = fmaf(p.x, inv_L, dy[i].x); dy[i].y = fmaf(p.y, inv_L, dy[i].y);
        dy[i].z = fmaf(p.z, inv_L, dy[i].z); dy[i].w = fmaf(p.w, inv_L, dy[i].w);
      }
      xh[i] = __ldg(reinterpret_cast<const float4*>(s + (int64_t)rr * E) + lane);
      mu[i] = mean[rr]; rs[i] = rstd[rr];
      f[i] = dpf ? dpf[b] : 1.f;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      xh[i].x = (xh[i].x - mu[i]) * rs[i]; xh[i].y = (xh[i].y - mu[i]) * rs[i];
      xh[i].z = (xh[i].z - mu[i]) * rs[i]; xh[i].w = (xh[i].w - mu[i]) * rs[i];
      const float4 dg = make_float4(dy[i].x * g4.x, dy[i].y * g4.y, dy[i].z * g4.z, dy[i].w * g4.w);
      float s1 = (dg.x + dg.y) + (dg.z + dg.w);
      float s2 = fmaf(dg.x, xh[i].x, fmaf(dg.y, xh[i].y, fmaf(dg.z, xh[i].z, dg.w * xh[i].w)));
      s1 = warp_sum(s1) * (1.f / (float)E);
      s2 = warp_sum(s2) * (1.f / (float)E);
      if (ok[i]) {
        float4 dx;
        dx.x = rs[i] * (dg.x - s1 - xh[i].x * s2); dx.y = rs[i] * (dg.y - s1 - xh[i].y * s2);
        dx.z = rs[i] * (dg.z - s1 - xh[i].z * s2); dx.w = rs[i] * (dg.w - s1 - xh[i].w * s2);
        const int64_t o = (int64_t)(row0 + i) * E;
        reinterpret_cast<float4*>(dx_out + o)[lane] = dx;
        reinterpret_cast<float4*>(dbranch + o)[lane] = make_float4(dx.x * f[i], dx.y * f[i], dx.z * f[i], dx.w * f[i]);
        ag.x = fmaf(dy[i].x, xh[i].x, ag.x); ag.y = fmaf(dy[i].y, xh[i].y, ag.y);
        ag.z = fmaf(dy[i].z, xh[i].z, ag.z); ag.w = fmaf(dy[i].w, xh[i].w, ag.w);
        ab.x += dy[i].x; ab.y += dy[i].y; ab.z += dy[i].z; ab.w += dy[i].w;
        abr.x = fmaf(dx.x, f[i], abr.x); abr.y = fmaf(dx.y, f[i], abr.y); abr.z = fmaf(dx.z, f[i], abr.z); abr.w = fmaf(dx.w, f[i], abr.w);
      }
    }
  }
  float* r = red[warp];
  r[4 * lane] = ag.x; r[4 * lane + 1] = ag.y; r[4 * lane + 2] = ag.z; r[4 * lane + 3] = ag.w;
  r[E + 4 * lane] = ab.x; r[E + 4 * lane + 1] = ab.y; r[E + 4 * lane + 2] = ab.z; r[E + 4 * lane + 3] = ab.w;
  r[2 * E + 4 * lane] = abr.x; r[2 * E + 4 * lane + 1] = abr.y; r[2 * E + 4 * lane + 2] = abr.z; r[2 * E + 4 * lane + 3] = abr.w;
  __syncthreads();
  for (int c = threadIdx.x; c < (dbranch_bias ? 3 * E : 2 * E); c += 256) {   // columns of dgamma | dbeta | dbranch_bias
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][c];
    atomicAdd(c < E ? dgamma + c : c < 2 * E ? dbeta + (c - E) : dbranch_bias + (c - 2 * E), t);
  }
}
// dbranch_bias (optional): += column sums of dbranch (the bias gradient of the Linear that produced the branch).  Returns
// whether the kernel accumulated it; if not the caller runs the column-sum pass.
static bool launch_ln_bwd(cudaStream_t st, int grid, const float* dy_full, const float* dy_pool, int64_t pool_ld, float inv_L, const float* s,
                          const float* mean, const float* rstd, const float* gamma, const float* dpf, float* dx_out, float* dbranch,
                          float* dgamma, float* dbeta, float* dbranch_bias, int rows, int L, int E) {
  const bool fast = E == 128 && (pool_ld & 3) == 0;
  if (fast)
    ln_bwd128_kernel<<<grid, 256, 0, st>>>(dy_full, dy_pool, pool_ld, inv_L, s, mean, rstd, gamma, dpf, dx_out, dbranch, dgamma, dbeta,
                                           dbranch_bias, rows, L);
  else
    ln_bwd_kernel<<<grid, 256, 0, st>>>(dy_full, dy_pool, pool_ld, inv_L, s, mean, rstd, gamma, dpf, dx_out, dbranch, dgamma, dbeta, rows, L, E);
  count_launch();
  return fast;
}

// ---- attention backward ---------------------------------------------------------------------------------------
// Per clip, heads in sequence.  With Pd = dropout(P):  dV = Pd^T dO ; dPd = dO V^T ; dP = dPd ⊙ mask ;
// dS = P ⊙ (dP - rowsum(dP ⊙ P)) ; dQ = scale · dS K ; dK = dS^T (scale·Q) ; dBias = Σ_h dS_h and through
// bias = tanh(qs_l + ks_s)·β :  g = dBias·β·(1 - tanh²) ; dqs_l = Σ_s g ; dks_s = Σ_l g ; dβ = Σ dBias·tanh.
__global__ void __launch_bounds__(256) attn_bwd_kernel(AttnBwdArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, S = a.S, E = a.E, H = a.H, dh = E / H, dp = dh + 1;
  float* Qs = sm;                  // [L][dp]  (scaled q)
  float* Ks = Qs + L * dp;         // [S][dp]
  float* Vs = Ks + S * dp;         // [S][dp]
  float* dOs = Vs + S * dp;        // [L][dp]
  float* dSs = dOs + L * dp;       // [L][S]
  float* Pds = dSs + L * S;        // [L][S]
  float* dBs = Pds + L * S;        // [L][S]  Σ_h dS
  __shared__ float red[32];
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float* dO = a.dO + (int64_t)b * L * E;
  for (int i = tid; i < L * S; i += blockDim.x) dBs[i] = 0.f;
  for (int h = 0; h < H; ++h) {
    for (int i = tid; i < L * dh; i += blockDim.x) {
      const int l = i / dh, d = i - l * dh;
      Qs[l * dp + d] = q[(int64_t)l * E + h * dh + d] * a.scale;
      dOs[l * dp + d] = dO[(int64_t)l * E + h * dh + d];
    }
    for (int i = tid; i < S * dh; i += blockDim.x) {
      const int s = i / dh, d = i - s * dh;
      Ks[s * dp + d] = kv[(int64_t)s * 2 * E + h * dh + d];
      Vs[s * dp + d] = kv[(int64_t)s * 2 * E + E + h * dh + d];
    }
    __syncthreads();
    for (int l = warp; l < L; l += nw) {
      const int64_t pbase = (((int64_t)b * H + h) * L + l) * S;
      float rowdot = 0.f;
      for (int s = lane; s < S; s += 32) {
        const float p = a.P[pbase + s];
        const float f = a.p_drop > 0.f ? drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)(pbase + s)) : 1.f;
        float dpd = 0.f;
        for (int d = 0; d < dh; ++d) dpd = fmaf(dOs[l * dp + d], Vs[s * dp + d], dpd);
        const float dP = dpd * f;
        Pds[l * S + s] = p * f;
        dSs[l * S + s] = dP;          // holds dP until the row dot is known
        rowdot = fmaf(dP, p, rowdot);
      }
      rowdot = warp_sum(rowdot);
      for (int s = lane; s < S; s += 32) {
        const float p = a.P[pbase + s];
        dSs[l * S + s] = p * (dSs[l * S + s] - rowdot);
      }
    }
    __syncthreads();
    // dV, dK : [S, dh]
    for (int i = tid; i < S * dh; i += blockDim.x) {
      const int s = i / dh, d = i - s * dh;
      float dv = 0.f, dk = 0.f;
      for (int l = 0; l < L; ++l) {
        dv = fmaf(Pds[l * S + s], dOs[l * dp + d], dv);
        dk = fmaf(dSs[l * S + s], Qs[l * dp + d], dk);
      }
      a.dkv[((int64_t)b * S + s) * 2 * E + h * dh + d] = dk;
      a.dkv[((int64_t)b * S + s) * 2 * E + E + h * dh + d] = dv;
    }
    // dQ : [L, dh]
    for (int i = tid; i < L * dh; i += blockDim.x) {
      const int l = i / dh, d = i - l * dh;
      float dq = 0.f;
      for (int s = 0; s < S; ++s) dq = fmaf(dSs[l * S + s], Ks[s * dp + d], dq);
      a.dq[((int64_t)b * L + l) * E + h * dh + d] = dq * a.scale;
    }
    for (int i = tid; i < L * S; i += blockDim.x) dBs[i] += dSs[i];
    __syncthreads();
  }
  if (a.qs == nullptr) return;
  const float* qs = a.qs + (int64_t)b * L;
  const float* ks = a.ks + (int64_t)b * S;
  const float bscale = a.bias_scale[0];
  float dscale = 0.f;
  for (int i = tid; i < L * S; i += blockDim.x) {
    const int l = i / S, s = i - l * S;
    const float t = tanhf(qs[l] + ks[s]);
    const float db = dBs[i];
    dscale = fmaf(db, t, dscale);
    dBs[i] = db * bscale * (1.f - t * t);
  }
  dscale = warp_sum(dscale);
  if (lane == 0) red[warp] = dscale;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int i = 0; i < nw; ++i) s += red[i];
    atomicAdd(a.dbias_scale, s);
  }
  for (int l = warp; l < L; l += nw) {
    float s = 0.f;
    for (int j = lane; j < S; j += 32) s += dBs[l * S + j];
    s = warp_sum(s);
    if (lane == 0) a.dqs[(int64_t)b * L + l] = s;
  }
  for (int j = tid; j < S; j += blockDim.x) {
    float s = 0.f;
    for (int l = 0; l < L; ++l) s += dBs[l * S + j];
    a.dks[(int64_t)b * S + j] = s;
  }
}


// The same for dh == 32 with register-tiled products (smem_tile.cuh).  [L][SP] matrices (SP = S rounded up to 4) come first in
// shared memory so their rows stay 16-byte aligned.
__global__ void __launch_bounds__(256, 2) attn_bwd32_kernel(AttnBwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  __shared__ float bred[8][96];   // per-warp column sums of dq | dk | dv of the current head
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, S = a.S, E = a.E, H = a.H, SP = (S + 3) & ~3;
  constexpr int dh = 32, dp = 33;
  float* Ps = sm;                  // [L][SP]  P
  float* Pds = Ps + L * SP;        // [L][SP]  dropout(P)
  float* dSs = Pds + L * SP;       // [L][SP]  dP, then dS
  float* dBs = dSs + L * SP;       // [L][SP]  sum over heads of dS
  float* Qs = dBs + L * SP;        // [L][dp]  (scaled q)
  float* Ks = Qs + L * dp;         // [S][dp]
  float* Vs = Ks + S * dp;         // [S][dp]
  float* dOs = Vs + S * dp;        // [L][dp]
  __shared__ float red[32];
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float* dO = a.dO + (int64_t)b * L * E;
  const bool lanes_on_s = S >= L;
  const int AR = lanes_on_s ? L : S, BR = lanes_on_s ? S : L;
  const float* Am = lanes_on_s ? dOs : Vs;
  const float* Bm = lanes_on_s ? Vs : dOs;
  const int tiles_a = (AR + 1) / 2, ntile = tiles_a * ((BR + 159) / 160);
  for (int i = tid; i < L * SP; i += blockDim.x) {
    dBs[i] = 0.f;
    Pds[i] = 0.f;   // (padding columns stay zero: the dQ product reads them)
    dSs[i] = 0.f;
  }
  const bool p_vec = (S & 3) == 0 && (reinterpret_cast<uintptr_t>(a.P) & 15) == 0;   // SP == S: Ps has the layout of P itself
  for (int h = 0; h < H; ++h) {
    stage_rows32(Qs, q + h * dh, L, E, a.scale, tid, blockDim.x);
    stage_rows32(dOs, dO + h * dh, L, E, 1.f, tid, blockDim.x);
    stage_rows32x2(Ks, Vs, kv + h * dh, E, S, 2 * E, tid, blockDim.x);
    const int64_t hbase = ((int64_t)b * H + h) * L * S;
    // P of this head -> Ps with coalesced loads (in the a2v direction the tile epilogue below would read it with a 64-byte stride)
    if (p_vec) {
      const float4* src = reinterpret_cast<const float4*>(a.P + hbase);
      float4* dst = reinterpret_cast<float4*>(Ps);
#pragma unroll 4
      for (int i = tid; i < (L * S) >> 2; i += blockDim.x) dst[i] = __ldg(src + i);
    } else {
      for (int l = warp; l < L; l += nw)
        for (int s2 = lane; s2 < S; s2 += 32) Ps[l * SP + s2] = a.P[hbase + (int64_t)l * S + s2];
    }
    __syncthreads();
    for (int t = warp; t < ntile; t += nw) {   // dPd = dO · V^T
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
            const int64_t pi = hbase + (int64_t)l * S + s2;
            const float p = Ps[l * SP + s2];
            const float f = a.p_drop > 0.f ? drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)pi) : 1.f;
            Pds[l * SP + s2] = p * f;
            dSs[l * SP + s2] = acc[i][j] * f;   // dP until the row dot is known
          }
        }
    }
    __syncthreads();
    if (p_vec && S <= 32) {   // short rows (a2v: 16 scores): lpr lanes per row with four scores each, 32/lpr rows per warp
      const int lpr = S <= 16 ? 4 : 8, rpw = 32 / lpr, sub = lane / lpr, sl = lane - sub * lpr;
      const bool act = 4 * sl < S;
      for (int l0 = warp * rpw; l0 < L; l0 += nw * rpw) {
        const int l = min(l0 + sub, L - 1);
        const bool ok = act && l0 + sub < L;
        const int o = l * SP + 4 * sl;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 pv = act ? *reinterpret_cast<const float4*>(Ps + o) : z;
        float4 dv = act ? *reinterpret_cast<const float4*>(dSs + o) : z;
        float rowdot = fmaf(dv.x, pv.x, fmaf(dv.y, pv.y, fmaf(dv.z, pv.z, dv.w * pv.w)));
        for (int w2 = lpr >> 1; w2 > 0; w2 >>= 1) rowdot += __shfl_xor_sync(0xffffffffu, rowdot, w2);
        if (ok) {
          dv.x = pv.x * (dv.x - rowdot); dv.y = pv.y * (dv.y - rowdot); dv.z = pv.z * (dv.z - rowdot); dv.w = pv.w * (dv.w - rowdot);
          *reinterpret_cast<float4*>(dSs + o) = dv;
          float4 db = *reinterpret_cast<float4*>(dBs + o);
          db.x += dv.x; db.y += dv.y; db.z += dv.z; db.w += dv.w;
          *reinterpret_cast<float4*>(dBs + o) = db;
        }
      }
    } else
    for (int l = warp; l < L; l += nw) {   // dS = P ⊙ (dP - rowsum(dP ⊙ P))
      float rowdot = 0.f;
      for (int s2 = lane; s2 < S; s2 += 32) rowdot = fmaf(dSs[l * SP + s2], Ps[l * SP + s2], rowdot);
      rowdot = warp_sum(rowdot);
      for (int s2 = lane; s2 < S; s2 += 32) {
        const float ds = Ps[l * SP + s2] * (dSs[l * SP + s2] - rowdot);
        dSs[l * SP + s2] = ds;
        dBs[l * SP + s2] += ds;
      }
    }
    __syncthreads();
    // dV = Pd^T dO, dK = dS^T (scale·Q): [S, dh], lane = column.  bq / bk / bv: this warp's column sums of the rows it
    // produced (the in_proj bias gradients, a.dbq / a.dbkv) — the separate column-sum passes over dq / dkv re-read them from HBM.
    float bq = 0.f, bk = 0.f, bv = 0.f;
    if (S > 32) {
      for (int t = warp; t < (S + 3) / 4; t += nw) {
        float dv[4], dk[4];
        tile_am<4>(Pds, 4 * t, SP, dOs, dp, L, lane, dv);
        tile_am<4>(dSs, 4 * t, SP, Qs, dp, L, lane, dk);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (4 * t + i < S) {
            a.dkv[((int64_t)b * S + 4 * t + i) * 2 * E + h * dh + lane] = dk[i];
            a.dkv[((int64_t)b * S + 4 * t + i) * 2 * E + E + h * dh + lane] = dv[i];
            bk += dk[i];
            bv += dv[i];
          }
      }
    } else {
      for (int t = warp; t < (S + 1) / 2; t += nw) {
        float dv[2], dk[2];
        tile_am<2>(Pds, 2 * t, SP, dOs, dp, L, lane, dv);
        tile_am<2>(dSs, 2 * t, SP, Qs, dp, L, lane, dk);
#pragma unroll
        for (int i = 0; i < 2; ++i)
          if (2 * t + i < S) {
            a.dkv[((int64_t)b * S + 2 * t + i) * 2 * E + h * dh + lane] = dk[i];
            a.dkv[((int64_t)b * S + 2 * t + i) * 2 * E + E + h * dh + lane] = dv[i];
            bk += dk[i];
            bv += dv[i];
          }
      }
    }
    // dQ = scale · dS K: [L, dh]
    if (L > 32) {
      for (int t = warp; t < (L + 3) / 4; t += nw) {
        float dq[4];
        tile_ak<4>(dSs, 4 * t, L, SP, Ks, dp, S, lane, dq);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (4 * t + i < L) {
            a.dq[((int64_t)b * L + 4 * t + i) * E + h * dh + lane] = dq[i] * a.scale;
            bq += dq[i] * a.scale;
          }
      }
    } else {
      for (int t = warp; t < (L + 1) / 2; t += nw) {
        float dq[2];
        tile_ak<2>(dSs, 2 * t, L, SP, Ks, dp, S, lane, dq);
#pragma unroll
        for (int i = 0; i < 2; ++i)
          if (2 * t + i < L) {
            a.dq[((int64_t)b * L + 2 * t + i) * E + h * dh + lane] = dq[i] * a.scale;
            bq += dq[i] * a.scale;
          }
      }
    }
    if (a.dbq) {
      bred[warp][lane] = bq;
      bred[warp][32 + lane] = bk;
      bred[warp][64 + lane] = bv;
    }
    __syncthreads();
    if (a.dbq && tid < 96) {   // (bred is rewritten only after the next head's staging barrier)
      float t = 0.f;
      for (int w2 = 0; w2 < nw; ++w2) t += bred[w2][tid];
      const int which = tid >> 5, c = h * dh + (tid & 31);
      atomicAdd(which == 0 ? a.dbq + c : a.dbkv + (which - 1) * E + c, t);
    }
  }
  if (a.qs == nullptr) return;
  const float* qs = a.qs + (int64_t)b * L;
  const float* ks = a.ks + (int64_t)b * S;
  const float bscale = a.bias_scale[0];
  float dscale = 0.f;
  for (int i = tid; i < L * S; i += blockDim.x) {
    const int l = i / S, s2 = i - l * S;
    const float t = tanhf(qs[l] + ks[s2]);
    const float db = dBs[l * SP + s2];
    dscale = fmaf(db, t, dscale);
    dBs[l * SP + s2] = db * bscale * (1.f - t * t);
  }
  dscale = warp_sum(dscale);
  if (lane == 0) red[warp] = dscale;
  __syncthreads();
  if (tid == 0) {
    float s2 = 0.f;
    for (int i = 0; i < nw; ++i) s2 += red[i];
    atomicAdd(a.dbias_scale, s2);
  }
  for (int l = warp; l < L; l += nw) {
    float s2 = 0.f;
    for (int j = lane; j < S; j += 32) s2 += dBs[l * SP + j];
    s2 = warp_sum(s2);
    if (lane == 0) a.dqs[(int64_t)b * L + l] = s2;
  }
  for (int j = tid; j < S; j += blockDim.x) {
    float s2 = 0.f;
    for (int l = 0; l < L; ++l) s2 += dBs[l * SP + j];
    a.dks[(int64_t)b * S + j] = s2;
  }
}

// *bias_fused: the kernel accumulated a.dbq / a.dbkv itself (else the caller runs the column sums over dq / dkv)
static int launch_attn_bwd(cudaStream_t st, const AttnBwdArgs& a, int B, bool* bias_fused) {
  static const bool simt_env = getenv("XATTN_ATTN_SIMT") != nullptr;
  if (g_tc_gemm && !simt_env && attn_mma_ok(a.L, a.S, a.E, a.H)) {
    *bias_fused = a.dbq != nullptr;
    return launch_attn_bwd_mma(st, a, B);
  }
  const int dh = a.E / a.H;
  {   // score matrices too large for the one-CTA-per-clip kernels (self-attention of the transformer pooler's audio side)
    const int SP = (a.S + 3) & ~3;
    const size_t need = dh == 32 ? sizeof(float) * (4 * (size_t)a.L * SP + (size_t)(2 * a.L + 2 * a.S) * 33)
                                 : sizeof(float) * ((size_t)(2 * a.L + 2 * a.S) * (dh + 1) + 3 * (size_t)a.L * a.S);
    if (need > 200 * 1024 && a.qs == nullptr) {
      *bias_fused = false;
      return launch_attn_bwd_rows(st, a, B);
    }
  }
  *bias_fused = dh == 32 && a.dbq != nullptr;
  if (dh == 32) {
    const int SP = (a.S + 3) & ~3;
    const size_t smem32 = sizeof(float) * (4 * (size_t)a.L * SP + (size_t)(2 * a.L + 2 * a.S) * 33);
    if (smem32 > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention backward tile needs %zu B of shared memory (> 200 KiB)", smem32);
    XF_TRY(ensure_dyn_smem(attn_bwd32_kernel, smem32));
    attn_bwd32_kernel<<<B, 256, smem32, st>>>(a);
    count_launch();
    XF_LAUNCH_CHECK();
    return XATTN_OK;
  }
  const size_t smem = sizeof(float) * ((size_t)(2 * a.L + 2 * a.S) * (dh + 1) + 3 * (size_t)a.L * a.S);
  if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention backward tile needs %zu B of shared memory (> 200 KiB)", smem);
  XF_TRY(ensure_dyn_smem(attn_bwd_kernel, smem));
  attn_bwd_kernel<<<B, 256, smem, st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// ---- emotion prior backward -------------------------------------------------------------------------------------
struct PriorBwdArgs {
  const float *v0, *a0, *phid, *prior, *w0, *w3;
  const float* wh[4];
  const float *d_qv, *d_ka, *d_qa, *d_kvs;
  float *d_v0, *d_a0, *d_prior, *d_phid;
  float* g_wh[4];
  float* g_bh[4];
  int T, Ta, E, P, Hd;
  float inv_keep;
  int vec;   // float4 token loop (E % 4 == 0, dpool 16-byte aligned in shared memory is not needed: scalar smem accesses)
  float *g_bv, *g_ba;   // vec path: += column sums of the final d_v0 / d_a0 (the v_in_proj / a_in_proj bias gradients)
};
__global__ void __launch_bounds__(256) prior_bwd_kernel(PriorBwdArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int E = a.E, P = a.P, Hd = a.Hd, T = a.T, Ta = a.Ta;
  float* sq = sm;               // [4]
  float* dprior = sq + 4;       // [P]
  float* dphid = dprior + P;    // [Hd]
  float* dpool = dphid + Hd;    // [2E]
  const float* dsc[4] = {a.d_qv + (int64_t)b * T, a.d_ka + (int64_t)b * Ta, a.d_qa + (int64_t)b * Ta, a.d_kvs + (int64_t)b * T};
  const int len[4] = {T, Ta, Ta, T};
  if (warp < 4) {
    float s = 0.f;
    for (int i = lane; i < len[warp]; i += 32) s += dsc[warp][i];
    s = warp_sum(s);
    if (lane == 0) sq[warp] = s;
  }
  __syncthreads();
  for (int p = tid; p < P; p += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < 4; ++k) s = fmaf(sq[k], a.wh[k][E + p], s);
    dprior[p] = s;
    a.d_prior[(int64_t)b * P + p] = s;
    const float pr = a.prior[(int64_t)b * P + p];
    for (int k = 0; k < 4; ++k) atomicAdd(a.g_wh[k] + E + p, sq[k] * pr);
  }
  if (tid < 4) atomicAdd(a.g_bh[tid], sq[tid]);
  __syncthreads();
  for (int h = tid; h < Hd; h += blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < P; ++p) s = fmaf(dprior[p], a.w3[(int64_t)p * Hd + h], s);
    s = a.phid[(int64_t)b * Hd + h] > 0.f ? s * a.inv_keep : 0.f;
    dphid[h] = s;
    a.d_phid[(int64_t)b * Hd + h] = s;
  }
  __syncthreads();
  for (int k = tid; k < 2 * E; k += blockDim.x) {
    float s = 0.f;
    for (int h = 0; h < Hd; ++h) s = fmaf(dphid[h], a.w0[(int64_t)h * 2 * E + k], s);
    dpool[k] = s;
  }
  __syncthreads();
  const float* v0 = a.v0 + (int64_t)b * T * E;
  const float* a0 = a.a0 + (int64_t)b * Ta * E;
  float* dv0 = a.d_v0 + (int64_t)b * T * E;
  float* da0 = a.d_a0 + (int64_t)b * Ta * E;
  if (a.vec) {
    // token gradients + the token part of the score-head weights: warps stride over the tokens, a lane owns four columns
    // (float4 loads / read-modify-writes, independent across tokens); the four weight-gradient rows of the CTA's warps
    // meet in shared memory, then one atomic per column.  (The scalar loop below walks the tokens with E threads.)
    float* sd = dpool + 2 * E;          // [2T + 2Ta] score gradients of this clip
    float* gred = sd + 2 * (T + Ta);    // [nw][6][E]: the four score-head rows, then the column sums of d_v0 and d_a0
    for (int i = tid; i < T; i += blockDim.x) { sd[i] = dsc[0][i]; sd[T + i] = dsc[3][i]; }
    for (int j = tid; j < Ta; j += blockDim.x) { sd[2 * T + j] = dsc[1][j]; sd[2 * T + Ta + j] = dsc[2][j]; }
    __syncthreads();
    const float inv_T = 1.f / (float)T, inv_Ta = 1.f / (float)Ta;
    for (int c4 = lane; c4 < (E >> 2); c4 += 32) {
      const int e = 4 * c4;
      float wq[4], wk[4], g0[4] = {0.f, 0.f, 0.f, 0.f}, g1[4] = {0.f, 0.f, 0.f, 0.f}, g2[4] = {0.f, 0.f, 0.f, 0.f}, g3[4] = {0.f, 0.f, 0.f, 0.f}, pm[4];
      float4 sv = make_float4(0.f, 0.f, 0.f, 0.f), sa = sv;   // column sums of the finished token gradients
#pragma unroll
      for (int k = 0; k < 4; ++k) { wq[k] = a.wh[0][e + k]; wk[k] = a.wh[3][e + k]; pm[k] = dpool[e + k] * inv_T; }
      // (four tokens per step with all eight loads issued before the first store: the gradient pointers may alias the token
      // pointers as far as the compiler knows, so it will not move a later load above an earlier store by itself)
      for (int i0 = warp; i0 < T; i0 += 4 * nw) {
        float4 x[4], d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * nw;
          if (i < T) {
            x[u] = __ldg(reinterpret_cast<const float4*>(v0 + (int64_t)i * E) + c4);
            d[u] = *(reinterpret_cast<const float4*>(dv0 + (int64_t)i * E) + c4);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i = i0 + u * nw;
          if (i < T) {
            const float dq = sd[i], dk = sd[T + i];
            d[u].x += dq * wq[0] + dk * wk[0] + pm[0]; d[u].y += dq * wq[1] + dk * wk[1] + pm[1];
            d[u].z += dq * wq[2] + dk * wk[2] + pm[2]; d[u].w += dq * wq[3] + dk * wk[3] + pm[3];
            *(reinterpret_cast<float4*>(dv0 + (int64_t)i * E) + c4) = d[u];
            sv.x += d[u].x; sv.y += d[u].y; sv.z += d[u].z; sv.w += d[u].w;
            g0[0] = fmaf(dq, x[u].x, g0[0]); g0[1] = fmaf(dq, x[u].y, g0[1]); g0[2] = fmaf(dq, x[u].z, g0[2]); g0[3] = fmaf(dq, x[u].w, g0[3]);
            g3[0] = fmaf(dk, x[u].x, g3[0]); g3[1] = fmaf(dk, x[u].y, g3[1]); g3[2] = fmaf(dk, x[u].z, g3[2]); g3[3] = fmaf(dk, x[u].w, g3[3]);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) { wk[k] = a.wh[1][e + k]; wq[k] = a.wh[2][e + k]; pm[k] = dpool[E + e + k] * inv_Ta; }
      for (int j0 = warp; j0 < Ta; j0 += 4 * nw) {
        float4 x[4], d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + u * nw;
          if (j < Ta) {
            x[u] = __ldg(reinterpret_cast<const float4*>(a0 + (int64_t)j * E) + c4);
            d[u] = *(reinterpret_cast<const float4*>(da0 + (int64_t)j * E) + c4);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + u * nw;
          if (j < Ta) {
            const float dk = sd[2 * T + j], dq = sd[2 * T + Ta + j];
            d[u].x += dk * wk[0] + dq * wq[0] + pm[0]; d[u].y += dk * wk[1] + dq * wq[1] + pm[1];
            d[u].z += dk * wk[2] + dq * wq[2] + pm[2]; d[u].w += dk * wk[3] + dq * wq[3] + pm[3];
            *(reinterpret_cast<float4*>(da0 + (int64_t)j * E) + c4) = d[u];
            sa.x += d[u].x; sa.y += d[u].y; sa.z += d[u].z; sa.w += d[u].w;
            g1[0] = fmaf(dk, x[u].x, g1[0]); g1[1] = fmaf(dk, x[u].y, g1[1]); g1[2] = fmaf(dk, x[u].z, g1[2]); g1[3] = fmaf(dk, x[u].w, g1[3]);
            g2[0] = fmaf(dq, x[u].x, g2[0]); g2[1] = fmaf(dq, x[u].y, g2[1]); g2[2] = fmaf(dq, x[u].z, g2[2]); g2[3] = fmaf(dq, x[u].w, g2[3]);
          }
        }
      }
      float* gr = gred + (int64_t)warp * 6 * E + e;
#pragma unroll
      for (int k = 0; k < 4; ++k) { gr[k] = g0[k]; gr[E + k] = g1[k]; gr[2 * E + k] = g2[k]; gr[3 * E + k] = g3[k]; }
      gr[4 * E] = sv.x; gr[4 * E + 1] = sv.y; gr[4 * E + 2] = sv.z; gr[4 * E + 3] = sv.w;
      gr[5 * E] = sa.x; gr[5 * E + 1] = sa.y; gr[5 * E + 2] = sa.z; gr[5 * E + 3] = sa.w;
    }
    __syncthreads();
    for (int idx = tid; idx < 6 * E; idx += blockDim.x) {
      float t = 0.f;
      for (int w2 = 0; w2 < nw; ++w2) t += gred[(int64_t)w2 * 6 * E + idx];
      const int which = idx / E, e = idx - which * E;
      atomicAdd(which < 4 ? a.g_wh[which] + e : which == 4 ? a.g_bv + e : a.g_ba + e, t);
    }
    return;
  }
  // token gradients + the token part of the score-head weights (block-local reduction, then one atomic per e)
  for (int e = tid; e < E; e += blockDim.x) {
    const float w_vq = a.wh[0][e], w_ak = a.wh[1][e], w_aq = a.wh[2][e], w_vk = a.wh[3][e];
    const float pv = dpool[e] / (float)T, pa = dpool[E + e] / (float)Ta;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
    for (int i = 0; i < T; ++i) {
      const float dq = dsc[0][i], dk = dsc[3][i], x = v0[(int64_t)i * E + e];
      dv0[(int64_t)i * E + e] += dq * w_vq + dk * w_vk + pv;
      g0 = fmaf(dq, x, g0);
      g3 = fmaf(dk, x, g3);
    }
    for (int j = 0; j < Ta; ++j) {
      const float dk = dsc[1][j], dq = dsc[2][j], x = a0[(int64_t)j * E + e];
      da0[(int64_t)j * E + e] += dk * w_ak + dq * w_aq + pa;
      g1 = fmaf(dk, x, g1);
      g2 = fmaf(dq, x, g2);
    }
    atomicAdd(a.g_wh[0] + e, g0);
    atomicAdd(a.g_wh[1] + e, g1);
    atomicAdd(a.g_wh[2] + e, g2);
    atomicAdd(a.g_wh[3] + e, g3);
  }
}

// The same for d_model = 128 with a short dependency chain in front of the token loop (the kernel above walks W0 with 64 dependent
// rounds of loads per thread and runs at two CTAs per SM): the clip's score gradients are staged in shared memory once, the
// Hd x 2E product for d_pooled runs as 4 x 64 threads with 16 float4 weight loads in flight each, and a warp covers a token row
// with one float4 per lane.  Same arithmetic and outputs as the `vec` path of prior_bwd_kernel.
__global__ void __launch_bounds__(256, 3) prior_bwd128_kernel(PriorBwdArgs a) {
  extern __shared__ __align__(16) float sm[];
  constexpr int E = 128;
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int P = a.P, Hd = a.Hd, T = a.T, Ta = a.Ta;
  float* dpool = sm;                    // [2E]
  float* part = dpool + 2 * E;          // [4][2E] partial d_pooled; later gred [6][E] (starts here): the CTA's sums of the four
  float* gred = part;                   //   score-head rows and of the d_v0 / d_a0 columns, met by shared-memory atomics
  float* sd = part + 4 * 2 * E;         // [2T + 2Ta]: d_qv | d_kvs | d_ka | d_qa
  float* sq = sd + 2 * (T + Ta);        // [4]
  float* dprior = sq + 4;               // [P]
  float* dphid = dprior + P;            // [Hd]
  for (int i = tid; i < T; i += blockDim.x) {
    sd[i] = a.d_qv[(int64_t)b * T + i];
    sd[T + i] = a.d_kvs[(int64_t)b * T + i];
  }
  for (int j = tid; j < Ta; j += blockDim.x) {
    sd[2 * T + j] = a.d_ka[(int64_t)b * Ta + j];
    sd[2 * T + Ta + j] = a.d_qa[(int64_t)b * Ta + j];
  }
  __syncthreads();
  if (warp < 4) {   // sums in the order of the score heads: v_query, a_key, a_query, v_key
    const float* src = warp == 0 ? sd : warp == 1 ? sd + 2 * T : warp == 2 ? sd + 2 * T + Ta : sd + T;
    const int n = (warp == 0 || warp == 3) ? T : Ta;
    float s = 0.f;
    for (int i = lane; i < n; i += 32) s += src[i];
    s = warp_sum(s);
    if (lane == 0) sq[warp] = s;
  }
  __syncthreads();
  for (int p = tid; p < P; p += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k < 4; ++k) s = fmaf(sq[k], a.wh[k][E + p], s);
    dprior[p] = s;
    a.d_prior[(int64_t)b * P + p] = s;
    const float pr = a.prior[(int64_t)b * P + p];
    for (int k = 0; k < 4; ++k) atomicAdd(a.g_wh[k] + E + p, sq[k] * pr);
  }
  if (tid < 4) atomicAdd(a.g_bh[tid], sq[tid]);
  __syncthreads();
  for (int h = tid; h < Hd; h += blockDim.x) {
    float s = 0.f;
    for (int p = 0; p < P; ++p) s = fmaf(dprior[p], a.w3[(int64_t)p * Hd + h], s);
    s = a.phid[(int64_t)b * Hd + h] > 0.f ? s * a.inv_keep : 0.f;
    dphid[h] = s;
    a.d_phid[(int64_t)b * Hd + h] = s;
  }
  __syncthreads();
  {   // d_pooled = d_phid · W0: thread (k4, hp) takes four columns and every fourth-of-the-range hidden unit
    const int k4 = tid & 63, hp = tid >> 6;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int h0 = hp * 16; h0 < Hd; h0 += 64) {
      float4 w[16];
#pragma unroll
      for (int q = 0; q < 16; ++q)
        w[q] = h0 + q < Hd ? __ldg(reinterpret_cast<const float4*>(a.w0 + (int64_t)(h0 + q) * 2 * E) + k4) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < 16; ++q) {
        const float d = h0 + q < Hd ? dphid[h0 + q] : 0.f;
        acc.x = fmaf(d, w[q].x, acc.x); acc.y = fmaf(d, w[q].y, acc.y); acc.z = fmaf(d, w[q].z, acc.z); acc.w = fmaf(d, w[q].w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(part + hp * 2 * E + 4 * k4) = acc;
  }
  __syncthreads();
  dpool[tid] = part[tid] + part[2 * E + tid] + part[4 * E + tid] + part[6 * E + tid];   // (blockDim.x == 2E)
  __syncthreads();
  for (int i = tid; i < 6 * E; i += blockDim.x) gred[i] = 0.f;   // (part is dead)
  __syncthreads();
  const float* v0 = a.v0 + (int64_t)b * T * E;
  const float* a0 = a.a0 + (int64_t)b * Ta * E;
  float* dv0 = a.d_v0 + (int64_t)b * T * E;
  float* da0 = a.d_a0 + (int64_t)b * Ta * E;
  float4 gA, gB, cs;   // a side's two score-head rows and the column sums of its finished token gradients
  // x += dA·wA + dB·wB + pm over the side's tokens (four rows per warp in flight, all loads before the first store: the gradient
  // pointers may alias the token pointers as far as the compiler knows)
  auto side = [&](const float* x, float* dx, int L, const float* sA, const float* sB, const float* wAp, const float* wBp, const float* pool,
                  float inv_L) {
    const float4 wA = __ldg(reinterpret_cast<const float4*>(wAp) + lane), wB = __ldg(reinterpret_cast<const float4*>(wBp) + lane);
    const float4 pl = *reinterpret_cast<const float4*>(pool + 4 * lane);
    const float4 pm = make_float4(pl.x * inv_L, pl.y * inv_L, pl.z * inv_L, pl.w * inv_L);
    gA = gB = cs = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i0 = warp; i0 < L; i0 += 4 * nw) {
      float4 xr[4], d[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * nw;
        if (i < L) {
          xr[u] = __ldg(reinterpret_cast<const float4*>(x + (int64_t)i * E) + lane);
          d[u] = *(reinterpret_cast<const float4*>(dx + (int64_t)i * E) + lane);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * nw;
        if (i < L) {
          const float dA = sA[i], dB = sB[i];
          d[u].x += dA * wA.x + dB * wB.x + pm.x; d[u].y += dA * wA.y + dB * wB.y + pm.y;
          d[u].z += dA * wA.z + dB * wB.z + pm.z; d[u].w += dA * wA.w + dB * wB.w + pm.w;
          *(reinterpret_cast<float4*>(dx + (int64_t)i * E) + lane) = d[u];
          cs.x += d[u].x; cs.y += d[u].y; cs.z += d[u].z; cs.w += d[u].w;
          gA.x = fmaf(dA, xr[u].x, gA.x); gA.y = fmaf(dA, xr[u].y, gA.y); gA.z = fmaf(dA, xr[u].z, gA.z); gA.w = fmaf(dA, xr[u].w, gA.w);
          gB.x = fmaf(dB, xr[u].x, gB.x); gB.y = fmaf(dB, xr[u].y, gB.y); gB.z = fmaf(dB, xr[u].z, gB.z); gB.w = fmaf(dB, xr[u].w, gB.w);
        }
      }
    }
  };
  float* gr = gred + 4 * lane;   // rows: v_query, a_key, a_query, v_key, colsum d_v0, colsum d_a0
  auto put = [&](int row, const float4& v) {
    float* d = gr + row * E;
    atomicAdd(d, v.x); atomicAdd(d + 1, v.y); atomicAdd(d + 2, v.z); atomicAdd(d + 3, v.w);
  };
  side(v0, dv0, T, sd, sd + T, a.wh[0], a.wh[3], dpool, 1.f / (float)T);                        // dA = d_qv (v_query), dB = d_kvs (v_key)
  put(0, gA); put(3, gB); put(4, cs);
  side(a0, da0, Ta, sd + 2 * T, sd + 2 * T + Ta, a.wh[1], a.wh[2], dpool + E, 1.f / (float)Ta);   // dA = d_ka (a_key), dB = d_qa (a_query)
  put(1, gA); put(2, gB); put(5, cs);
  __syncthreads();
  for (int idx = tid; idx < 6 * E; idx += blockDim.x) {
    const int which = idx / E, e = idx - which * E;
    atomicAdd(which < 4 ? a.g_wh[which] + e : which == 4 ? a.g_bv + e : a.g_ba + e, gred[idx]);
  }
}

// ---- TemporalAttentionPooling backward (temporal.py:9-26) ---------------------------------------------------------
// out = Σ_l a_l x_l, a = softmax_l(s), s_l = w2·(GELU(h_l) ⊙ m_l) + b2, h_l = W1 y_l + b1, y_l = LN(x_l)·γ + β.  Nothing but x was
// saved: a first pass over the clip's tokens recomputes the scores and d_a_l = <d_out, x_l>, the softmax backward gives d_s, a
// second pass recomputes y_l, h_l and pushes d_s_l through the score MLP and the LayerNorm:
//   d_x_l = a_l·d_out + LN'(W1ᵀ d_h_l ⊙ γ),  d_h_l = d_s_l·w2 ⊙ m_l ⊙ GELU'(h_l).
// dγ, dβ, dw2, db2 are accumulated here (registers -> shared memory -> one atomic per CTA and entry); y_l and d_h_l are written out
// ([B·L,E], [B·L,H2]) so that dW1 = d_hᵀ·y and db1 = colsum(d_h) run as the batch GEMM / column sum of the Linear helpers.
// One CTA per clip, one warp per token; W1 sits in shared memory twice ([e][j] for h, [j][e] for W1ᵀ d_h — both conflict-free).
struct AttnPoolBwdArgs {
  const float *x, *ln_w, *ln_b, *w1, *b1, *w2, *b2;
  const float* d_out;      // [B, ld_out] gradient of the pooled vector
  int64_t ld_out;
  float* d_x;              // [B,L,E]
  int accumulate;          // d_x += (else =)
  float *y_out, *dh_out;   // [B·L,E], [B·L,H2]
  float *g_ln_w, *g_ln_b, *g_w2, *g_b2;
  int L, E;
  float p_drop;
  uint64_t seed, offset;
  uint32_t rng_stream;
};
__global__ void __launch_bounds__(256) attn_pool_bwd_kernel(AttnPoolBwdArgs a) {
  extern __shared__ float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, E = a.E, H2 = E / 2 > 0 ? E / 2 : 1;
  float* w1t = sm;                    // [E][H2]
  float* w1s = w1t + E * H2;          // [H2][E]
  float* yrow = w1s + E * H2;         // [nw][E]   y_l, then d_x̂_l
  float* xh = yrow + nw * E;          // [nw][E]   x̂_l
  float* dhrow = xh + nw * E;         // [nw][H2]
  float* score = dhrow + nw * H2;     // [L]  s_l, then a_l
  float* da = score + L;              // [L]  <d_out, x_l>, then d_s_l
  float* red = da + L;                // [nw][2E + H2 + 1] partial dγ | dβ | dw2 | db2
  for (int i = tid; i < E * H2; i += blockDim.x) {
    const int j = i / E, e = i - j * E;
    const float w = a.w1[i];
    w1t[e * H2 + j] = w;
    w1s[i] = w;
  }
  __syncthreads();
  const float* xb = a.x + (int64_t)b * L * E;
  const float* dout = a.d_out + (int64_t)b * a.ld_out;
  float* y = yrow + warp * E;
  float* xhat = xh + warp * E;
  float* dh = dhrow + warp * H2;
  // recompute of one token row into y / xhat (returns rstd); h_j for j = lane + 32k is then produced by the caller's loop
  auto ln_row = [&](int l) -> float {
    const float* xr = xb + (int64_t)l * E;
    float sum = 0.f;
    for (int e = lane; e < E; e += 32) sum += xr[e];
    const float mu = warp_sum(sum) / (float)E;
    float var = 0.f;
    for (int e = lane; e < E; e += 32) {
      const float d = xr[e] - mu;
      var = fmaf(d, d, var);
    }
    const float rstd = rsqrtf(warp_sum(var) / (float)E + 1e-5f);
    for (int e = lane; e < E; e += 32) {
      const float xn = (xr[e] - mu) * rstd;
      xhat[e] = xn;
      y[e] = xn * a.ln_w[e] + a.ln_b[e];
    }
    __syncwarp();
    return rstd;
  };
  for (int l = warp; l < L; l += nw) {   // pass 1: scores and d_a
    ln_row(l);
    float sc = 0.f;
    for (int j = lane; j < H2; j += 32) {
      float h = a.b1[j];
      for (int e = 0; e < E; ++e) h = fmaf(w1t[e * H2 + j], y[e], h);
      float g = 0.5f * h * (1.f + erff(h * 0.70710678118654752f));
      if (a.p_drop > 0.f) g *= drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, ((uint64_t)b * L + l) * H2 + j);
      sc = fmaf(a.w2[j], g, sc);
    }
    float d = 0.f;
    for (int e = lane; e < E; e += 32) d = fmaf(dout[e], xb[(int64_t)l * E + e], d);
    sc = warp_sum(sc);
    d = warp_sum(d);
    if (lane == 0) {
      score[l] = sc + a.b2[0];
      da[l] = d;
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == 0) {   // softmax over time and its backward: d_s_l = a_l (d_a_l - Σ_m a_m d_a_m)
    float mx = -INFINITY;
    for (int l = lane; l < L; l += 32) mx = fmaxf(mx, score[l]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int l = lane; l < L; l += 32) {
      const float e = expf(score[l] - mx);
      score[l] = e;
      sum += e;
    }
    const float inv = 1.f / warp_sum(sum);
    float dot = 0.f;
    for (int l = lane; l < L; l += 32) {
      const float p = score[l] * inv;
      score[l] = p;
      dot = fmaf(p, da[l], dot);
    }
    dot = warp_sum(dot);
    for (int l = lane; l < L; l += 32) da[l] = score[l] * (da[l] - dot);
  }
  __syncthreads();
  constexpr int MAXK = 16;   // E <= 512, H2 <= 256
  float gg[MAXK], gb[MAXK], gw2[MAXK / 2];
#pragma unroll
  for (int k = 0; k < MAXK; ++k) gg[k] = gb[k] = 0.f;
#pragma unroll
  for (int k = 0; k < MAXK / 2; ++k) gw2[k] = 0.f;
  float gb2 = 0.f;
  for (int l = warp; l < L; l += nw) {   // pass 2
    const float rstd = ln_row(l);
    const float ds = da[l], al = score[l];
    const int64_t row = (int64_t)b * L + l;
#pragma unroll
    for (int k = 0; k < MAXK / 2; ++k) {
      const int j = lane + 32 * k;
      if (j < H2) {
        float h = a.b1[j];
        for (int e = 0; e < E; ++e) h = fmaf(w1t[e * H2 + j], y[e], h);
        const float cdf = 0.5f * (1.f + erff(h * 0.70710678118654752f));
        const float pdf = 0.3989422804014327f * expf(-0.5f * h * h);
        const float m = a.p_drop > 0.f ? drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)row * H2 + j) : 1.f;
        gw2[k] = fmaf(ds, h * cdf * m, gw2[k]);
        const float d_h = ds * a.w2[j] * m * (cdf + h * pdf);
        dh[j] = d_h;
        a.dh_out[row * H2 + j] = d_h;
      }
    }
    if (lane == 0) gb2 += ds;
    __syncwarp();
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        float dy = 0.f;
        for (int j = 0; j < H2; ++j) dy = fmaf(w1s[j * E + e], dh[j], dy);
        a.y_out[row * E + e] = y[e];
        gg[k] = fmaf(dy, xhat[e], gg[k]);
        gb[k] += dy;
        const float dxn = dy * a.ln_w[e];
        m1 += dxn;
        m2 = fmaf(dxn, xhat[e], m2);
        y[e] = dxn;   // (each lane re-reads only its own entries below)
      }
    }
    m1 = warp_sum(m1) / (float)E;
    m2 = warp_sum(m2) / (float)E;
#pragma unroll
    for (int k = 0; k < MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        const float dx = fmaf(al, dout[e], rstd * (y[e] - m1 - xhat[e] * m2));
        float* p = a.d_x + row * E + e;
        *p = a.accumulate ? *p + dx : dx;
      }
    }
    __syncwarp();
  }
  const int RW = 2 * E + H2 + 1;
  float* r = red + warp * RW;
#pragma unroll
  for (int k = 0; k < MAXK; ++k) {
    const int e = lane + 32 * k;
    if (e < E) { r[e] = gg[k]; r[E + e] = gb[k]; }
  }
#pragma unroll
  for (int k = 0; k < MAXK / 2; ++k) {
    const int j = lane + 32 * k;
    if (j < H2) r[2 * E + j] = gw2[k];
  }
  if (lane == 0) r[2 * E + H2] = gb2;
  __syncthreads();
  for (int i = tid; i < RW; i += blockDim.x) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += red[w * RW + i];
    float* dst = i < E ? a.g_ln_w + i : i < 2 * E ? a.g_ln_b + (i - E) : i < 2 * E + H2 ? a.g_w2 + (i - 2 * E) : a.g_b2;
    atomicAdd(dst, t);
  }
}
static int launch_attn_pool_bwd(cudaStream_t st, const AttnPoolBwdArgs& a, int B) {
  const int E = a.E, H2 = E / 2 > 0 ? E / 2 : 1, nw = 8;
  if (E > 512) XF_FAIL(XATTN_ERR_SHAPE, "attention pooling backward supports d_model <= 512 (got %d)", E);
  const size_t smem = sizeof(float) * (2 * (size_t)E * H2 + 2 * (size_t)nw * E + (size_t)nw * H2 + 2 * (size_t)a.L + (size_t)nw * (2 * E + H2 + 1));
  if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention pooling backward needs %zu B of shared memory (> 200 KiB)", smem);
  XF_TRY(ensure_dyn_smem(attn_pool_bwd_kernel, smem));
  attn_pool_bwd_kernel<<<B, 256, smem, st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// ---- orchestration ------------------------------------------------------------------------------------------------
// `opt.tail_stream`: made to wait for the point of `st` where every gradient EXCEPT the input projections' (the flat buffer
// from XP_V2A_IN_W on) is final, so the caller can all-reduce that part while the input-projection backward still runs.
// `opt.d_video` / `opt.d_audio`: gradients w.r.t. the embeddings (Stage 2 of the reference unfreezes backbone layers,
// train.py:815-830): dX = dY·W through v_in_proj / audio_seq_proj, fp32 [B,T,Dv] / [B,Ta,Da].
template <typename TIO>
static int backward_f32_impl(const xattn_config& c, const float* params, const TIO* xv, const TIO* xa, const float* dlogits,
                             float* grads, float* ws, cudaStream_t st, const BackwardOptions& opt) {
  ParamLayout pl;
  compute_param_layout(c, pl);
  WsF32 w;
  compute_ws_f32(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  auto G = [&](int id) { return grads + pl.off[id]; };
  const int B = c.B, T = c.T, Ta = c.Ta, E = c.E, H = c.H, C = c.C;
  const bool train = c.training != 0;
  const float p_attn = train ? c.p_attn : 0.f, p_head = train ? c.p_head : 0.f, p_prior = train ? c.p_prior : 0.f;
  const bool gated = c.head_kind == XATTN_HEAD_GATED;
  const int HW = gated ? E : c.common_dim;
  const float scale = 1.0f / sqrtf((float)(E / H));

  NvtxRange nvtx_bwd(g_tc_gemm ? "xattn_backward[tf32]" : "xattn_backward[fp32]");
  // d_v0 / d_a0 are final in the prior backward kernels' float4 paths: they leave the column sums in the two bias gradients
  const bool tok_bias_fused =
      c.use_prior && ((E == 128 && sizeof(float) * (2 * E + 8 * E + 2 * (T + Ta) + 4 + c.prior_dim + c.prior_hidden) <= 48 * 1024) ||
                      (E % 4 == 0 && sizeof(float) * (4 + c.prior_dim + c.prior_hidden + 2 * E + 2 * (T + Ta) + 8 * 6 * E) <= 48 * 1024));
  // opt.phase: 0 = the whole backward; 1 = up to the point where every gradient but the input projections' is final; 2 = the
  // input projections only (after a phase-1 call on the same workspace and gradient buffer)
  if (opt.phase != 2) {
    XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * pl.total, st));

    // ---- head (fusion.py:405-411) ----
    if (!gated) {
      linear_bwd_dw<float>(st, dlogits, C, ws + w.hhid, HW, G(XP_HEAD1_W), B, C, HW);
      launch_colsum(st, dlogits, C, B, C, G(XP_HEAD1_B));
      head_bwd_mid_kernel<<<B, 256, 0, st>>>(dlogits, P(XP_HEAD1_W), ws + w.hhid, nullptr, nullptr, nullptr, ws + w.d_hh, nullptr,
                                             nullptr, E, HW, C, 0, 1.f / (1.f - p_head));
      count_launch();
      linear_bwd_dw<float>(st, ws + w.d_hh, HW, ws + w.pooled, 2 * E, G(XP_HEAD0_W), B, HW, 2 * E);
      launch_colsum(st, ws + w.d_hh, HW, B, HW, G(XP_HEAD0_B));
      linear_bwd_dx(st, ws + w.d_hh, HW, P(XP_HEAD0_W), ws + w.d_pooled, B, HW, 2 * E, false);
    } else {
      linear_bwd_dw<float>(st, dlogits, C, ws + w.fusedg, E, G(XP_CLS_W), B, C, E);
      launch_colsum(st, dlogits, C, B, C, G(XP_CLS_B));
      linear_bwd_dx(st, dlogits, C, P(XP_CLS_W), ws + w.d_fused, B, C, E, false);
      head_bwd_mid_kernel<<<B, 256, 0, st>>>(nullptr, P(XP_HEAD1_W), ws + w.hhid, ws + w.pooled, ws + w.gate, ws + w.d_fused,
                                             ws + w.d_hh, ws + w.d_pooled, ws + w.d_z, E, HW, C, 1, 1.f / (1.f - p_head));
      count_launch();
      linear_bwd_dw<float>(st, ws + w.d_z, 1, ws + w.hhid, HW, G(XP_HEAD1_W), B, 1, HW);
      launch_colsum(st, ws + w.d_z, 1, B, 1, G(XP_HEAD1_B));
      linear_bwd_dw<float>(st, ws + w.d_hh, HW, ws + w.pooled, 2 * E, G(XP_HEAD0_W), B, HW, 2 * E);
      launch_colsum(st, ws + w.d_hh, HW, B, HW, G(XP_HEAD0_B));
      linear_bwd_dx(st, ws + w.d_hh, HW, P(XP_HEAD0_W), ws + w.d_pooled, B, HW, 2 * E, true);
    }
    XF_LAUNCH_CHECK();

    const int ln_grid = 4 * num_sms();
    bool bias_fused = false;
    const bool attn_pool = c.pool_mode == XATTN_POOL_ATTN;
    const int H2 = E / 2 > 0 ? E / 2 : 1;
    const float p_pool = train ? c.p_pool : 0.f;
    // learned pooling (temporal.py:9-26): token gradients from attn_pool_bwd_kernel instead of d_pooled/L; its W1 / b1 gradients as
    // a batch GEMM / column sum over the y and d_h rows it wrote into two buffers that are still free at that point
    auto pool_bwd = [&](const float* x, const float* d_out, float* d_x, bool accumulate, float* y_buf, float* dh_buf, int L, int first_id,
                        uint32_t stream) -> int {
      AttnPoolBwdArgs pa;
      pa.x = x; pa.ln_w = P(first_id); pa.ln_b = P(first_id + 1); pa.w1 = P(first_id + 2); pa.b1 = P(first_id + 3);
      pa.w2 = P(first_id + 4); pa.b2 = P(first_id + 5);
      pa.d_out = d_out; pa.ld_out = 2 * E; pa.d_x = d_x; pa.accumulate = accumulate ? 1 : 0; pa.y_out = y_buf; pa.dh_out = dh_buf;
      pa.g_ln_w = G(first_id); pa.g_ln_b = G(first_id + 1); pa.g_w2 = G(first_id + 4); pa.g_b2 = G(first_id + 5);
      pa.L = L; pa.E = E; pa.p_drop = p_pool; pa.seed = c.seed; pa.offset = c.offset; pa.rng_stream = stream;
      XF_TRY(launch_attn_pool_bwd(st, pa, B));
      linear_bwd_dw<float>(st, dh_buf, H2, y_buf, E, G(first_id + 2), B * L, H2, E);
      launch_colsum(st, dh_buf, H2, B * L, H2, G(first_id + 3));
      return XATTN_OK;
    };
    static_assert(XP_APOOL_B2 - XP_APOOL_LN_W == 5 && XP_VPOOL_B2 - XP_VPOOL_LN_W == 5, "pooler parameter ids are consecutive");
    // TemporalTransformerPooling (temporal.py:46-75): the learned pooling's backward on the last layer's output, then the encoder
    // layers in reverse (x2 = x1 + drop(W2·drop(GELU(W1·LN2(x1)))), x1 = x0 + drop(Wo·SA(LN1(x0)))); the sinusoidal positions pass
    // the gradient through.  Leaves the gradient w.r.t. the side's tokens in ws + w.txb_dx.  h = LN(x) is recomputed, everything
    // else comes from the layer's TxSaved block; the forward's scratch set (tx_h, tx_q, tx_kv, tx_f) holds the temporaries.
    const bool tx_pool = c.pool_mode == XATTN_POOL_TRANSFORMER;
    auto transformer_pool_bwd = [&](int side, const float* d_out) -> int {
      const int L = side == 0 ? T : Ta, M = B * L, F = 4 * E, PH = c.pool_heads, layers = pool_layers(c);
      const int64_t n = (int64_t)M * E;
      float *dX = ws + w.txb_dx, *t1 = ws + w.txb_t1, *t2 = ws + w.txb_t2, *th = ws + w.tx_h, *dq = ws + w.tx_q, *dkv = ws + w.tx_kv,
            *df = ws + w.tx_f;
      XF_TRY(pool_bwd(ws + w.tx_out[side], d_out, dX, false, t1, ws + w.txb_dh, L, side == 0 ? XP_VPOOL_LN_W : XP_APOOL_LN_W,
                      side == 0 ? RS_POOLV : RS_POOLA));
      for (int layer = layers - 1; layer >= 0; --layer) {
        const TxSaved sv = tx_saved(ws, w, c, side, layer);
        auto TP = [&](int k) { return P(XATTN_TX_PARAM_ID(side, layer, k)); };
        auto TG = [&](int k) { return G(XATTN_TX_PARAM_ID(side, layer, k)); };
        // ---- feed-forward block ----
        const float* dy = dX;
        if (p_pool > 0.f) {
          XF_TRY(launch_mul_drop(st, t1, dX, n, p_pool, c.seed, c.offset, rs_tx(side, layer, 3)));
          dy = t1;
        }
        linear_bwd_dw<float>(st, dy, E, sv.f_act, F, TG(6), M, E, F);
        launch_colsum(st, dy, E, M, E, TG(7));
        linear_bwd_dx(st, dy, E, TP(6), df, M, E, F, false);
        XF_TRY(launch_gelu_drop_bwd(st, df, sv.f_pre, (int64_t)M * F, p_pool, c.seed, c.offset, rs_tx(side, layer, 2)));
        XF_TRY(launch_ln_rows(st, sv.x_mid, TP(10), TP(11), th, M, E));
        linear_bwd_dw<float>(st, df, F, th, E, TG(4), M, F, E);
        launch_colsum(st, df, F, M, F, TG(5));
        linear_bwd_dx(st, df, F, TP(4), t2, M, F, E, false);
        XF_TRY(launch_ln_rows_bwd(st, sv.x_mid, TP(10), t2, dX, TG(10), TG(11), M, E, true));
        // ---- self-attention block ----
        dy = dX;
        if (p_pool > 0.f) {
          XF_TRY(launch_mul_drop(st, t1, dX, n, p_pool, c.seed, c.offset, rs_tx(side, layer, 1)));
          dy = t1;
        }
        linear_bwd_dw<float>(st, dy, E, sv.o, E, TG(2), M, E, E);
        launch_colsum(st, dy, E, M, E, TG(3));
        linear_bwd_dx(st, dy, E, TP(2), t2, M, E, E, false);
        AttnBwdArgs aa;
        aa.q = sv.q; aa.kv = sv.kv; aa.qs = nullptr; aa.ks = nullptr; aa.bias_scale = nullptr; aa.P = sv.P; aa.dO = t2;
        aa.dq = dq; aa.dkv = dkv; aa.dqs = nullptr; aa.dks = nullptr; aa.dbias_scale = nullptr;
        aa.L = L; aa.S = L; aa.E = E; aa.H = PH; aa.scale = 1.0f / sqrtf((float)(E / PH)); aa.p_drop = p_pool; aa.seed = c.seed;
        aa.offset = c.offset; aa.rng_stream = rs_tx(side, layer, 0);
        aa.dbq = TG(1); aa.dbkv = TG(1) + E;
        bool in_bias_fused = false;
        XF_TRY(launch_attn_bwd(st, aa, B, &in_bias_fused));
        XF_TRY(launch_ln_rows(st, sv.x_in, TP(8), TP(9), th, M, E));
        linear_bwd_dw<float>(st, dq, E, th, E, TG(0), M, E, E);
        linear_bwd_dw<float>(st, dkv, 2 * E, th, E, TG(0) + (int64_t)E * E, M, 2 * E, E);
        if (!in_bias_fused) {
          launch_colsum(st, dq, E, M, E, TG(1));
          launch_colsum(st, dkv, 2 * E, M, 2 * E, TG(1) + E);
        }
        linear_bwd_dx(st, dq, E, TP(0), t1, M, E, E, false);
        linear_bwd_dx(st, dkv, 2 * E, TP(0) + (int64_t)E * E, t1, M, 2 * E, E, true);
        XF_TRY(launch_ln_rows_bwd(st, sv.x_in, TP(8), t1, dX, TG(8), TG(9), M, E, true));
        XF_LAUNCH_CHECK();
      }
      return XATTN_OK;
    };
    if (attn_pool) XF_TRY(pool_bwd(ws + w.a1, ws + w.d_pooled + E, ws + w.d_q2, false, ws + w.d_br2, ws + w.d_o2, Ta, XP_APOOL_LN_W, RS_POOLA));   // (d_q2 is free until the a2v attention backward)
    if (tx_pool) XF_TRY(transformer_pool_bwd(1, ws + w.d_pooled + E));
    const bool tok_pool = attn_pool || tx_pool;   // the a_norm / v_norm outputs receive per-token gradients instead of d_pooled / L
    // ---- a_norm backward (fusion.py:399) + mean pool (temporal.py:108-109) ----
    bool ob_fused = launch_ln_bwd(st, ln_grid, tx_pool ? ws + w.txb_dx : attn_pool ? ws + w.d_q2 : nullptr, tok_pool ? nullptr : ws + w.d_pooled + E, 2 * E,
                                  1.f / (float)Ta, ws + w.s2, ws + w.mean2, ws + w.rstd2,
                                  P(XP_AN_W), ws + w.dp2, ws + w.d_a0, ws + w.d_br2, G(XP_AN_W), G(XP_AN_B), G(XP_A2V_OUT_B), B * Ta, Ta, E);
    // a2v out_proj
    linear_bwd_dw<float>(st, ws + w.d_br2, E, ws + w.o2, E, G(XP_A2V_OUT_W), B * Ta, E, E);
    if (!ob_fused) launch_colsum(st, ws + w.d_br2, E, B * Ta, E, G(XP_A2V_OUT_B));
    linear_bwd_dx(st, ws + w.d_br2, E, P(XP_A2V_OUT_W), ws + w.d_o2, B * Ta, E, E, false);
    XF_LAUNCH_CHECK();
    {
      AttnBwdArgs aa;
      aa.q = ws + w.q2; aa.kv = ws + w.kv2;
      aa.qs = c.use_prior ? ws + w.qa : nullptr; aa.ks = c.use_prior ? ws + w.kvs : nullptr;
      aa.bias_scale = c.use_prior ? P(XP_BIAS_SCALE) : nullptr; aa.P = ws + w.p2; aa.dO = ws + w.d_o2;
      aa.dq = ws + w.d_q2; aa.dkv = ws + w.d_kv2; aa.dqs = ws + w.d_qa; aa.dks = ws + w.d_kvs;
      aa.dbias_scale = c.use_prior ? G(XP_BIAS_SCALE) : nullptr;
      aa.L = Ta; aa.S = T; aa.E = E; aa.H = H; aa.scale = scale; aa.p_drop = p_attn; aa.seed = c.seed; aa.offset = c.offset;
      aa.rng_stream = RS_ATTN2; aa.ldp = p_row_stride(c, T);
      aa.dbq = G(XP_A2V_IN_B); aa.dbkv = G(XP_A2V_IN_B) + E;
      XF_TRY(launch_attn_bwd(st, aa, B, &bias_fused));
    }
    // a2v in_proj: q from a0, k|v from v1
    linear_bwd_dw<float>(st, ws + w.d_q2, E, ws + w.a0, E, G(XP_A2V_IN_W), B * Ta, E, E);
    if (!bias_fused) launch_colsum(st, ws + w.d_q2, E, B * Ta, E, G(XP_A2V_IN_B));
    linear_bwd_dx(st, ws + w.d_q2, E, P(XP_A2V_IN_W), ws + w.d_a0, B * Ta, E, E, true);
    linear_bwd_dw<float>(st, ws + w.d_kv2, 2 * E, ws + w.v1, E, G(XP_A2V_IN_W) + (int64_t)E * E, B * T, 2 * E, E);
    if (!bias_fused) launch_colsum(st, ws + w.d_kv2, 2 * E, B * T, 2 * E, G(XP_A2V_IN_B) + E);
    linear_bwd_dx(st, ws + w.d_kv2, 2 * E, P(XP_A2V_IN_W) + (int64_t)E * E, ws + w.d_v1, B * T, 2 * E, E, false);
    XF_LAUNCH_CHECK();
    if (attn_pool) XF_TRY(pool_bwd(ws + w.v1, ws + w.d_pooled, ws + w.d_v1, true, ws + w.d_br1, ws + w.d_o1, T, XP_VPOOL_LN_W, RS_POOLV));
    if (tx_pool) {
      XF_TRY(transformer_pool_bwd(0, ws + w.d_pooled));
      XF_TRY(launch_add_inplace(st, ws + w.d_v1, ws + w.txb_dx, (int64_t)B * T * E));
    }
    // ---- v_norm backward (fusion.py:395) ----
    ob_fused = launch_ln_bwd(st, ln_grid, ws + w.d_v1, tok_pool ? nullptr : ws + w.d_pooled, 2 * E, 1.f / (float)T, ws + w.s1, ws + w.mean1, ws + w.rstd1,
                             P(XP_VN_W), ws + w.dp1, ws + w.d_v0, ws + w.d_br1, G(XP_VN_W), G(XP_VN_B), G(XP_V2A_OUT_B), B * T, T, E);
    linear_bwd_dw<float>(st, ws + w.d_br1, E, ws + w.o1, E, G(XP_V2A_OUT_W), B * T, E, E);
    if (!ob_fused) launch_colsum(st, ws + w.d_br1, E, B * T, E, G(XP_V2A_OUT_B));
    linear_bwd_dx(st, ws + w.d_br1, E, P(XP_V2A_OUT_W), ws + w.d_o1, B * T, E, E, false);
    XF_LAUNCH_CHECK();
    {
      AttnBwdArgs aa;
      aa.q = ws + w.q1; aa.kv = ws + w.kv1;
      aa.qs = c.use_prior ? ws + w.qv : nullptr; aa.ks = c.use_prior ? ws + w.ka : nullptr;
      aa.bias_scale = c.use_prior ? P(XP_BIAS_SCALE) : nullptr; aa.P = ws + w.p1; aa.dO = ws + w.d_o1;
      aa.dq = ws + w.d_q1; aa.dkv = ws + w.d_kv1; aa.dqs = ws + w.d_qv; aa.dks = ws + w.d_ka;
      aa.dbias_scale = c.use_prior ? G(XP_BIAS_SCALE) : nullptr;
      aa.L = T; aa.S = Ta; aa.E = E; aa.H = H; aa.scale = scale; aa.p_drop = p_attn; aa.seed = c.seed; aa.offset = c.offset;
      aa.rng_stream = RS_ATTN1; aa.ldp = p_row_stride(c, Ta);
      aa.dbq = G(XP_V2A_IN_B); aa.dbkv = G(XP_V2A_IN_B) + E;
      XF_TRY(launch_attn_bwd(st, aa, B, &bias_fused));
    }
    // v2a in_proj: the token gradients first (the prior backward and the input projections wait for them) ...
    linear_bwd_dx(st, ws + w.d_q1, E, P(XP_V2A_IN_W), ws + w.d_v0, B * T, E, E, true);
    linear_bwd_dx(st, ws + w.d_kv1, 2 * E, P(XP_V2A_IN_W) + (int64_t)E * E, ws + w.d_a0, B * Ta, 2 * E, E, true);
    // ... then its weight gradients (HBM-bound tensor-core GEMMs nothing downstream reads) on the high-priority side stream under
    // tf32, beside the latency-bound per-clip prior backward on `st`
    SideFork sf(st);
    if (opt.fork && c.use_prior && g_tc_gemm) XF_TRY(sf.begin());
    linear_bwd_dw<float>(sf.side(), ws + w.d_q1, E, ws + w.v0, E, G(XP_V2A_IN_W), B * T, E, E);
    if (!bias_fused) launch_colsum(sf.side(), ws + w.d_q1, E, B * T, E, G(XP_V2A_IN_B));
    linear_bwd_dw<float>(sf.side(), ws + w.d_kv1, 2 * E, ws + w.a0, E, G(XP_V2A_IN_W) + (int64_t)E * E, B * Ta, 2 * E, E);
    if (!bias_fused) launch_colsum(sf.side(), ws + w.d_kv1, 2 * E, B * Ta, 2 * E, G(XP_V2A_IN_B) + E);
    XF_LAUNCH_CHECK();

    // ---- emotion prior (fusion.py:170-184) ----
    if (c.use_prior) {
      PriorBwdArgs pa;
      pa.v0 = ws + w.v0; pa.a0 = ws + w.a0; pa.phid = ws + w.phid; pa.prior = ws + w.prior;
      pa.w0 = P(XP_PRIOR0_W); pa.w3 = P(XP_PRIOR3_W);
      pa.wh[0] = P(XP_VQ_W); pa.wh[1] = P(XP_AK_W); pa.wh[2] = P(XP_AQ_W); pa.wh[3] = P(XP_VK_W);
      pa.g_wh[0] = G(XP_VQ_W); pa.g_wh[1] = G(XP_AK_W); pa.g_wh[2] = G(XP_AQ_W); pa.g_wh[3] = G(XP_VK_W);
      pa.g_bh[0] = G(XP_VQ_B); pa.g_bh[1] = G(XP_AK_B); pa.g_bh[2] = G(XP_AQ_B); pa.g_bh[3] = G(XP_VK_B);
      pa.d_qv = ws + w.d_qv; pa.d_ka = ws + w.d_ka; pa.d_qa = ws + w.d_qa; pa.d_kvs = ws + w.d_kvs;
      pa.d_v0 = ws + w.d_v0; pa.d_a0 = ws + w.d_a0; pa.d_prior = ws + w.d_prior; pa.d_phid = ws + w.d_phid;
      pa.T = T; pa.Ta = Ta; pa.E = E; pa.P = c.prior_dim; pa.Hd = c.prior_hidden; pa.inv_keep = 1.f / (1.f - p_prior);
      pa.g_bv = G(XP_V_IN_B); pa.g_ba = G(XP_A_IN_B);
      const size_t smem128 = sizeof(float) * (2 * E + 8 * E + 2 * (T + Ta) + 4 + c.prior_dim + c.prior_hidden);
      if (E == 128 && smem128 <= 48 * 1024) {
        pa.vec = 1;
        prior_bwd128_kernel<<<B, 256, smem128, st>>>(pa);
      } else {
        size_t smem = sizeof(float) * (4 + c.prior_dim + c.prior_hidden + 2 * E);
        const size_t smem_vec = smem + sizeof(float) * (2 * (T + Ta) + 8 * 6 * E);
        pa.vec = (E % 4 == 0 && smem_vec <= 48 * 1024) ? 1 : 0;
        if (pa.vec) smem = smem_vec;
        prior_bwd_kernel<<<B, 256, smem, st>>>(pa);
      }
      count_launch();
      linear_bwd_dw<float>(st, ws + w.d_prior, c.prior_dim, ws + w.phid, c.prior_hidden, G(XP_PRIOR3_W), B, c.prior_dim, c.prior_hidden);
      launch_colsum(st, ws + w.d_prior, c.prior_dim, B, c.prior_dim, G(XP_PRIOR3_B));
      linear_bwd_dw<float>(st, ws + w.d_phid, c.prior_hidden, ws + w.pooled0, 2 * E, G(XP_PRIOR0_W), B, c.prior_hidden, 2 * E);
      launch_colsum(st, ws + w.d_phid, c.prior_hidden, B, c.prior_hidden, G(XP_PRIOR0_B));
      XF_LAUNCH_CHECK();
    }
    XF_TRY(sf.join());
  }

  if (opt.tail_stream != nullptr) {   // everything but the input projections' gradients is final from here on
    cudaEvent_t ev = nullptr;
    XF_TRY(device_event(0, &ev));
    XF_CUDA(cudaEventRecord(ev, st));
    XF_CUDA(cudaStreamWaitEvent(opt.tail_stream, ev, 0));
  }
  if (opt.phase == 1) return XATTN_OK;
  NvtxRange nvtx_in("input_projections_backward");
  // ---- input projections (fusion.py:386, :378, :372); dX only on request (Stage 1 freezes the backbones) ----
  linear_bwd_dw<float>(st, ws + w.d_a0, E, ws + w.aseq, E, G(XP_A_IN_W), B * Ta, E, E);
  if (!tok_bias_fused) launch_colsum(st, ws + w.d_a0, E, B * Ta, E, G(XP_A_IN_B));
  linear_bwd_dx(st, ws + w.d_a0, E, P(XP_A_IN_W), ws + w.d_aseq, B * Ta, E, E, false);
  const bool tc16 = g_tc_gemm && std::is_same<TIO, __nv_bfloat16>::value && gemm_cols_ok(B * Ta, E, c.Da, E, c.Da, true) &&
                    gemm_cols_ok(B * T, E, c.Dv, E, c.Dv, true);
  if (tc16) {   // bf16 embeddings: the gradient is rounded to bf16 (into d_o2, dead by now) and both operands go through kind::f16
    __nv_bfloat16* dyb = reinterpret_cast<__nv_bfloat16*>(ws + w.d_o2);
    cvt_f32_to_bf16(st, ws + w.d_aseq, dyb, (int64_t)B * Ta * E);
    XF_TRY(gemm_cols(st, dyb, E, xa, c.Da, true, G(XP_SEQ_W), c.Da, B * Ta, E, c.Da));
  } else {
    linear_bwd_dw<TIO>(st, ws + w.d_aseq, E, xa, c.Da, G(XP_SEQ_W), B * Ta, E, c.Da);
  }
  // colsum(d_aseq) = colsum(d_a0)·W_ain: the audio_seq_proj bias gradient follows from the a_in_proj one by a 128 x 128 product
  // instead of another pass over the [B·Ta, E] gradient
  bias_chain_kernel<<<dim3((E + 127) / 128, (E + 15) / 16), 128, 0, st>>>(G(XP_A_IN_B), P(XP_A_IN_W), G(XP_SEQ_B), E, E);
  count_launch();
  if (tc16) {
    __nv_bfloat16* dyb = reinterpret_cast<__nv_bfloat16*>(ws + w.d_o1);
    cvt_f32_to_bf16(st, ws + w.d_v0, dyb, (int64_t)B * T * E);
    XF_TRY(gemm_cols(st, dyb, E, xv, c.Dv, true, G(XP_V_IN_W), c.Dv, B * T, E, c.Dv));
  } else {
    linear_bwd_dw<TIO>(st, ws + w.d_v0, E, xv, c.Dv, G(XP_V_IN_W), B * T, E, c.Dv);
  }
  if (!tok_bias_fused) launch_colsum(st, ws + w.d_v0, E, B * T, E, G(XP_V_IN_B));
  if (opt.d_audio != nullptr) linear_bwd_dx(st, ws + w.d_aseq, E, P(XP_SEQ_W), opt.d_audio, B * Ta, E, c.Da, false);
  if (opt.d_video != nullptr) linear_bwd_dx(st, ws + w.d_v0, E, P(XP_V_IN_W), opt.d_video, B * T, E, c.Dv, false);
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int backward_f32(const xattn_config& c, const float* params, const void* xv, const void* xa, const float* dlogits, float* grads,
                 void* ws, cudaStream_t st, const BackwardOptions& opt) {
  g_tc_gemm = c.precision == XATTN_PREC_TF32;
  g_tc_status = XATTN_OK;
  int rc;
  if (c.io_dtype == XATTN_DTYPE_BF16)
    rc = backward_f32_impl<__nv_bfloat16>(c, params, (const __nv_bfloat16*)xv, (const __nv_bfloat16*)xa, dlogits, grads, (float*)ws, st, opt);
  else
    rc = backward_f32_impl<float>(c, params, (const float*)xv, (const float*)xa, dlogits, grads, (float*)ws, st, opt);
  g_tc_gemm = false;
  return rc != XATTN_OK ? rc : g_tc_status;
}

}  // namespace xf
