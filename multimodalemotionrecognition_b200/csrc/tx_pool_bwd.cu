// Row-wise backward pieces of TemporalTransformerPooling's encoder layers (temporal.py:59-69: nn.TransformerEncoderLayer,
// norm_first, GELU) and the self-attention backward for token counts whose score matrices do not fit one CTA's shared memory
// (the audio side: L = S = 149).  The orchestration is transformer_pool_bwd in bwd_f32.cu; the Linears go through the shared
// helpers of gemm_simt.cuh.  fp32 SIMT: this pooler is not on a timed path.
#include "kernels_f32.cuh"

namespace xf {

// ---- pre-norm LayerNorm backward: h = LN(x)·γ + β, nothing saved -------------------------------------------------------------
// d_x (+)= rstd·(g - mean(g) - x̂·mean(g ⊙ x̂)), g = d_h ⊙ γ;  dγ += Σ_rows d_h ⊙ x̂;  dβ += Σ_rows d_h.
// One warp per row (grid-stride); the parameter gradients stay in registers over a warp's rows (E <= 512: 16 per lane), are summed
// over the CTA's warps in shared memory and leave as one atomic per CTA and entry.
__global__ void __launch_bounds__(256) ln_rows_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                          const float* __restrict__ d_h, float* __restrict__ d_x,
                                                          float* __restrict__ g_gamma, float* __restrict__ g_beta, int rows, int E,
                                                          int accumulate) {
  extern __shared__ float red[];   // [nw][2E]
  constexpr int MAXK = 16;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float gg[MAXK], gb[MAXK];
#pragma unroll
  for (int k = 0; k < MAXK; ++k) gg[k] = gb[k] = 0.f;
  for (int row = blockIdx.x * nw + warp; row < rows; row += gridDim.x * nw) {
    const float* xr = x + (int64_t)row * E;
    const float* dr = d_h + (int64_t)row * E;
    float sum = 0.f;
    for (int e = lane; e < E; e += 32) sum += xr[e];
    const float mu = warp_sum(sum) / (float)E;
    float var = 0.f;
    for (int e = lane; e < E; e += 32) {
      const float d = xr[e] - mu;
      var = fmaf(d, d, var);
    }
    const float rstd = rsqrtf(warp_sum(var) / (float)E + 1e-5f);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        const float xn = (xr[e] - mu) * rstd, dy = dr[e], g = dy * gamma[e];
        gg[k] = fmaf(dy, xn, gg[k]);
        gb[k] += dy;
        m1 += g;
        m2 = fmaf(g, xn, m2);
      }
    }
    m1 = warp_sum(m1) / (float)E;
    m2 = warp_sum(m2) / (float)E;
#pragma unroll
    for (int k = 0; k < MAXK; ++k) {
      const int e = lane + 32 * k;
      if (e < E) {
        const float xn = (xr[e] - mu) * rstd;
        const float dx = rstd * (dr[e] * gamma[e] - m1 - xn * m2);
        float* p = d_x + (int64_t)row * E + e;
        *p = accumulate ? *p + dx : dx;
      }
    }
  }
  float* r = red + warp * 2 * E;
#pragma unroll
  for (int k = 0; k < MAXK; ++k) {
    const int e = lane + 32 * k;
    if (e < E) { r[e] = gg[k]; r[E + e] = gb[k]; }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += red[w * 2 * E + i];
    atomicAdd(i < E ? g_gamma + i : g_beta + (i - E), t);
  }
}
int launch_ln_rows_bwd(cudaStream_t st, const float* x, const float* gamma, const float* d_h, float* d_x, float* g_gamma,
                       float* g_beta, int rows, int E, bool accumulate) {
  if (E > 512) XF_FAIL(XATTN_ERR_SHAPE, "LayerNorm backward supports d_model <= 512 (got %d)", E);
  int grid = (rows + 7) / 8;
  const int cap = 4 * num_sms();
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  ln_rows_bwd_kernel<<<grid, 256, sizeof(float) * 8 * 2 * E, st>>>(x, gamma, d_h, d_x, g_gamma, g_beta, rows, E, accumulate ? 1 : 0);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// ---- element-wise pieces -------------------------------------------------------------------------------------------------
// act = Dropout(GELU(pre))  =>  d_pre = d_act ⊙ mask ⊙ (Φ(pre) + pre·φ(pre)); in place over d
__global__ void gelu_drop_bwd_kernel(float* __restrict__ d, const float* __restrict__ pre, int64_t n, float p, uint64_t seed,
                                     uint64_t offset, uint32_t stream) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float h = pre[i];
    const float cdf = 0.5f * (1.f + erff(h * 0.70710678118654752f));
    const float pdf = 0.3989422804014327f * expf(-0.5f * h * h);
    d[i] = d[i] * drop_factor(p, seed, offset, stream, (uint64_t)i) * (cdf + h * pdf);
  }
}
// out = x ⊙ mask: the gradient entering a residual branch that ended in a Dropout
__global__ void mul_drop_kernel(float* __restrict__ out, const float* __restrict__ x, int64_t n, float p, uint64_t seed,
                                uint64_t offset, uint32_t stream) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = x[i] * drop_factor(p, seed, offset, stream, (uint64_t)i);
}
__global__ void add_inplace_kernel(float* __restrict__ x, const float* __restrict__ y, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] += y[i];
}
static inline int ew_grid(int64_t n) {
  const int64_t g = (n + 255) / 256, cap = (int64_t)num_sms() * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}
int launch_gelu_drop_bwd(cudaStream_t st, float* d, const float* pre, int64_t n, float p, uint64_t seed, uint64_t offset, uint32_t stream) {
  gelu_drop_bwd_kernel<<<ew_grid(n), 256, 0, st>>>(d, pre, n, p, seed, offset, stream);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
int launch_mul_drop(cudaStream_t st, float* out, const float* x, int64_t n, float p, uint64_t seed, uint64_t offset, uint32_t stream) {
  mul_drop_kernel<<<ew_grid(n), 256, 0, st>>>(out, x, n, p, seed, offset, stream);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}
int launch_add_inplace(cudaStream_t st, float* x, const float* y, int64_t n) {
  add_inplace_kernel<<<ew_grid(n), 256, 0, st>>>(x, y, n);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// ---- attention backward, one CTA per (clip, head), query rows in tiles ------------------------------------------------------
// Same arithmetic as attn_bwd_kernel (bwd_f32.cu) for calls WITHOUT the prior bias: K, V and the dK, dV accumulators of one head
// stay in shared memory; per tile of RT query rows the CTA forms Pd = P ⊙ mask, dS = P ⊙ (dP - rowsum(dP ⊙ P)) with
// dP = (dO Vᵀ) ⊙ mask, adds Pdᵀ dO / dSᵀ (scale·Q) into dV / dK and writes dQ = scale · dS K.  P holds the pre-dropout
// probabilities the fp32 forward kernels saved; the mask is regenerated from the same Philox index ((b·H + h)·L + l)·S + s.
__global__ void __launch_bounds__(256) attn_bwd_rows_kernel(AttnBwdArgs a, int RT) {
  extern __shared__ float sm[];
  const int b = blockIdx.x / a.H, h = blockIdx.x - b * a.H;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const int L = a.L, S = a.S, E = a.E, H = a.H, dh = E / H, dp = dh + 1;
  float* Ks = sm;                   // [S][dp]
  float* Vs = Ks + S * dp;          // [S][dp]
  float* dKs = Vs + S * dp;         // [S][dp]
  float* dVs = dKs + S * dp;        // [S][dp]
  float* Qs = dVs + S * dp;         // [RT][dp]  scale·q
  float* dOs = Qs + RT * dp;        // [RT][dp]
  float* dSs = dOs + RT * dp;       // [RT][S]
  float* Pds = dSs + RT * S;        // [RT][S]
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float* dO = a.dO + (int64_t)b * L * E;
  for (int i = tid; i < S * dh; i += blockDim.x) {
    const int s = i / dh, d = i - s * dh;
    Ks[s * dp + d] = kv[(int64_t)s * 2 * E + h * dh + d];
    Vs[s * dp + d] = kv[(int64_t)s * 2 * E + E + h * dh + d];
    dKs[s * dp + d] = 0.f;
    dVs[s * dp + d] = 0.f;
  }
  for (int l0 = 0; l0 < L; l0 += RT) {
    const int rt = min(RT, L - l0);
    for (int i = tid; i < rt * dh; i += blockDim.x) {
      const int r = i / dh, d = i - r * dh;
      Qs[r * dp + d] = q[(int64_t)(l0 + r) * E + h * dh + d] * a.scale;
      dOs[r * dp + d] = dO[(int64_t)(l0 + r) * E + h * dh + d];
    }
    __syncthreads();
    for (int r = warp; r < rt; r += nw) {
      const int64_t pbase = (((int64_t)b * H + h) * L + (l0 + r)) * S;
      float rowdot = 0.f;
      for (int s = lane; s < S; s += 32) {
        const float p = a.P[pbase + s];
        const float f = a.p_drop > 0.f ? drop_factor(a.p_drop, a.seed, a.offset, a.rng_stream, (uint64_t)(pbase + s)) : 1.f;
        float dpd = 0.f;
        for (int d = 0; d < dh; ++d) dpd = fmaf(dOs[r * dp + d], Vs[s * dp + d], dpd);
        const float dP = dpd * f;
        Pds[r * S + s] = p * f;
        dSs[r * S + s] = dP;   // dP until the row dot is known
        rowdot = fmaf(dP, p, rowdot);
      }
      rowdot = warp_sum(rowdot);
      for (int s = lane; s < S; s += 32) dSs[r * S + s] = a.P[pbase + s] * (dSs[r * S + s] - rowdot);
    }
    __syncthreads();
    for (int i = tid; i < S * dh; i += blockDim.x) {   // each thread owns its (s, d) accumulators
      const int s = i / dh, d = i - s * dh;
      float dv = dVs[s * dp + d], dk = dKs[s * dp + d];
      for (int r = 0; r < rt; ++r) {
        dv = fmaf(Pds[r * S + s], dOs[r * dp + d], dv);
        dk = fmaf(dSs[r * S + s], Qs[r * dp + d], dk);
      }
      dVs[s * dp + d] = dv;
      dKs[s * dp + d] = dk;
    }
    for (int i = tid; i < rt * dh; i += blockDim.x) {
      const int r = i / dh, d = i - r * dh;
      float dq = 0.f;
      for (int s = 0; s < S; ++s) dq = fmaf(dSs[r * S + s], Ks[s * dp + d], dq);
      a.dq[((int64_t)b * L + l0 + r) * E + h * dh + d] = dq * a.scale;
    }
    __syncthreads();
  }
  for (int i = tid; i < S * dh; i += blockDim.x) {
    const int s = i / dh, d = i - s * dh;
    a.dkv[((int64_t)b * S + s) * 2 * E + h * dh + d] = dKs[s * dp + d];
    a.dkv[((int64_t)b * S + s) * 2 * E + E + h * dh + d] = dVs[s * dp + d];
  }
}
int launch_attn_bwd_rows(cudaStream_t st, const AttnBwdArgs& a, int B) {
  if (a.qs != nullptr) XF_FAIL(XATTN_ERR_SHAPE, "the row-tiled attention backward has no prior-bias path (L=%d S=%d)", a.L, a.S);
  const int dh = a.E / a.H, dp = dh + 1;
  int RT = a.L < 32 ? a.L : 32;
  auto bytes = [&](int rt) { return sizeof(float) * (4 * (size_t)a.S * dp + 2 * (size_t)rt * dp + 2 * (size_t)rt * a.S); };
  while (RT > 1 && bytes(RT) > 200 * 1024) RT >>= 1;
  const size_t smem = bytes(RT);
  if (smem > 200 * 1024) XF_FAIL(XATTN_ERR_SHAPE, "attention backward needs %zu B of shared memory (> 200 KiB) at S=%d", smem, a.S);
  XF_TRY(ensure_dyn_smem(attn_bwd_rows_kernel, smem));
  attn_bwd_rows_kernel<<<B * a.H, 256, smem, st>>>(a, RT);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
