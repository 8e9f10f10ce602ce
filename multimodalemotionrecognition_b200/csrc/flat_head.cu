// Flat concat / gated fusion heads (BASELINE config 5; reference src/models/fusion.py:226-260, :413-435) with the
// CLIP-style alignment projections (:127-139).  The temporal mean that ends audio_model.encode /
// video_model.encode (:414-415) is fused in as the first kernel, so the inputs are the same [B,T,Dv] /
// [B,Ta,Da] embeddings the xattn head reads — an HBM-bound streaming reduction with 128-bit loads.
// Forward in eval and training mode (hidden-layer Dropout by Philox, ModalityDropout decisions from the caller), backward of the
// head and of the alignment loss (precision F32 / TF32); the bf16 tensor-core forward is the inference path.
#include "gemm_simt.cuh"
#include "gemm_tc.cuh"

namespace xf {

struct FlatLayout {
  int64_t off[FP_COUNT], size[FP_COUNT], total;
};
static inline int64_t al4(int64_t n) { return (n + 3) & ~(int64_t)3; }

static void flat_layout(const flat_head_config& c, FlatLayout& L) {
  const int64_t A = c.align_dim, Cm = c.common_dim;
  const int64_t in_a = A > 0 ? A : c.Da, in_v = A > 0 ? A : c.Dv;
  int64_t sz[FP_COUNT] = {0};
  if (A > 0) {
    sz[FP_ALIGN_A_W] = A * c.Da; sz[FP_ALIGN_A_B] = A; sz[FP_ALIGN_V_W] = A * c.Dv; sz[FP_ALIGN_V_B] = A;
  }
  sz[FP_AUDIO_W] = Cm * in_a; sz[FP_AUDIO_B] = Cm; sz[FP_VIDEO_W] = Cm * in_v; sz[FP_VIDEO_B] = Cm;
  sz[FP_H0_W] = Cm * 2 * Cm; sz[FP_H0_B] = Cm;
  sz[FP_H1_W] = c.gated ? Cm : (int64_t)c.C * Cm; sz[FP_H1_B] = c.gated ? 1 : c.C;
  sz[FP_CLS_W] = c.gated ? (int64_t)c.C * Cm : 0; sz[FP_CLS_B] = c.gated ? c.C : 0;
  int64_t off = 0;
  for (int i = 0; i < FP_COUNT; ++i) {
    L.off[i] = off;
    L.size[i] = sz[i];
    off += al4(sz[i]);
  }
  L.total = off;
}

static int flat_validate(const flat_head_config* c) {
  if (!c) XF_FAIL(XATTN_ERR_NULL, "cfg is NULL");
  if (c->B < 0 || c->T <= 0 || c->Ta <= 0 || c->Dv <= 0 || c->Da <= 0 || c->common_dim <= 0 || c->C <= 0 || c->align_dim < 0)
    XF_FAIL(XATTN_ERR_SHAPE, "bad flat head dimensions");
  if (c->io_dtype != XATTN_DTYPE_F32 && c->io_dtype != XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "unknown io_dtype %d", c->io_dtype);
  if (c->precision != XATTN_DTYPE_F32 && c->precision != XATTN_DTYPE_BF16 && c->precision != XATTN_PREC_TF32)
    XF_FAIL(XATTN_ERR_DTYPE, "unknown precision %d", c->precision);
  if (c->training && c->precision == XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "flat heads: training runs with precision f32 / tf32");
  if (c->training && !(c->p_drop >= 0.f && c->p_drop < 1.f)) XF_FAIL(XATTN_ERR_SHAPE, "flat heads: p_drop %f outside [0, 1)", (double)c->p_drop);
  if (c->precision == XATTN_DTYPE_BF16) {   // tcgen05 projection GEMMs: bf16 embeddings, K %% 64 == 0, N %% 128 == 0
    const int in_a = c->align_dim > 0 ? c->align_dim : c->Da, in_v = c->align_dim > 0 ? c->align_dim : c->Dv;
    if (c->io_dtype != XATTN_DTYPE_BF16 || c->Da % 64 || c->Dv % 64 || c->align_dim % 128 || c->common_dim % 128 || in_a % 64 || in_v % 64)
      XF_FAIL(XATTN_ERR_SHAPE, "flat heads, precision = BF16: bf16 embeddings with Da, Dv %% 64 == 0 and align_dim, common_dim %% 128 == 0 only");
  }
  return XATTN_OK;
}

struct FlatWs {
  int64_t a_pool, v_pool, a_al, v_al, cat, hid, z, total;   // offsets in floats
  int64_t wbf;                                               // precision = BF16: bf16 copy of the parameter buffer
  int64_t d_cat, d_hid, d_z, d_fused, d_a_al, d_v_al, d_a_pool, d_v_pool;   // backward scratch (precision F32 / TF32)
};
static void flat_ws(const flat_head_config& c, FlatWs& w) {
  const int64_t B = c.B, A = c.align_dim > 0 ? c.align_dim : 1, Cm = c.common_dim;
  int64_t off = 0;
  auto take = [&](int64_t n) { int64_t o = off; off += al4(n); return o; };
  w.a_pool = take(B * c.Da); w.v_pool = take(B * c.Dv); w.a_al = take(B * A); w.v_al = take(B * A);
  w.cat = take(B * 2 * Cm); w.hid = take(B * Cm); w.z = take(B);
  w.wbf = off;
  if (c.precision == XATTN_DTYPE_BF16) {   // (the activation slots above are reused as bf16: half of each is spare)
    FlatLayout pl;
    flat_layout(c, pl);
    take((pl.total + 1) / 2);
    w.d_cat = w.d_hid = w.d_z = w.d_fused = w.d_a_al = w.d_v_al = w.d_a_pool = w.d_v_pool = off;
  } else {
    w.d_cat = take(B * 2 * Cm); w.d_hid = take(B * Cm); w.d_z = take(B); w.d_fused = take(B * Cm);
    w.d_a_al = take(B * A); w.d_v_al = take(B * A); w.d_a_pool = take(B * c.Da); w.d_v_pool = take(B * c.Dv);
  }
  w.total = off;
}

__device__ __forceinline__ void store_out(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_out(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// Temporal mean over [B,L,D] -> [B,D]; each thread owns VEC consecutive channels (one 128-bit load per row),
// a CTA covers one clip's channel slab; rows unrolled x4 for memory-level parallelism.
template <typename TIO, typename TOUT>
__global__ void __launch_bounds__(128) temporal_mean_kernel(const TIO* __restrict__ x, TOUT* __restrict__ out, int L, int D) {
  constexpr int VEC = 16 / sizeof(TIO);
  const int b = blockIdx.y;
  const int c0 = (blockIdx.x * blockDim.x + threadIdx.x) * VEC;
  if (c0 >= D) return;
  const TIO* base = x + (int64_t)b * L * D + c0;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
  if (c0 + VEC <= D && (D % VEC) == 0) {
    int l = 0;
    for (; l + 4 <= L; l += 4) {
      uint4 r[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) r[u] = __ldg(reinterpret_cast<const uint4*>(base + (int64_t)(l + u) * D));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const TIO* e = reinterpret_cast<const TIO*>(&r[u]);
#pragma unroll
        for (int i = 0; i < VEC; ++i) acc[i] += to_f32(e[i]);
      }
    }
    for (; l < L; ++l) {
      const uint4 r = __ldg(reinterpret_cast<const uint4*>(base + (int64_t)l * D));
      const TIO* e = reinterpret_cast<const TIO*>(&r);
#pragma unroll
      for (int i = 0; i < VEC; ++i) acc[i] += to_f32(e[i]);
    }
  } else {
    for (int l = 0; l < L; ++l)
      for (int i = 0; i < VEC && c0 + i < D; ++i) acc[i] += to_f32(base[(int64_t)l * D + i]);
  }
  const float inv = 1.f / (float)L;
  for (int i = 0; i < VEC && c0 + i < D; ++i) store_out(out + (int64_t)b * D + c0 + i, acc[i] * inv);
}

template <typename TIO, typename TOUT>
static void launch_mean(cudaStream_t st, const TIO* x, TOUT* out, int B, int L, int D) {
  constexpr int VEC = 16 / sizeof(TIO);
  const int threads_needed = (D + VEC - 1) / VEC;
  dim3 grid((threads_needed + 127) / 128, B);
  temporal_mean_kernel<TIO, TOUT><<<grid, 128, 0, st>>>(x, out, L, D);
  count_launch();
}

// gated blend + classifier: g = sigmoid(z); fused = g*a + (1-g)*v (fusion.py:433-435)
__global__ void __launch_bounds__(256) flat_gate_cls_kernel(const float* __restrict__ cat, const float* __restrict__ z,
                                                            const float* __restrict__ wc, const float* __restrict__ bc,
                                                            float* __restrict__ logits, int Cm, int C) {
  extern __shared__ float fused[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
  const float g = 1.f / (1.f + expf(-z[b]));
  for (int e = tid; e < Cm; e += blockDim.x)
    fused[e] = g * cat[(int64_t)b * 2 * Cm + e] + (1.f - g) * cat[(int64_t)b * 2 * Cm + Cm + e];
  __syncthreads();
  for (int c = warp; c < C; c += nw) {
    float s = 0.f;
    for (int k = lane; k < Cm; k += 32) s = fmaf(wc[(int64_t)c * Cm + k], fused[k], s);
    s = warp_sum(s);
    if (lane == 0) logits[(int64_t)b * C + c] = s + bc[c];
  }
}

static inline int ew_blocks(int64_t n) {
  const int64_t g = (n + 255) / 256, cap = (int64_t)num_sms() * 16;
  return (int)(g < cap ? (g > 0 ? g : 1) : cap);
}
// cat = [a | v] per row: zero the audio and / or the video half
__global__ void zero_halves_kernel(float* __restrict__ cat, int64_t n, int Cm, int drop_a, int drop_v) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const bool first = (i % (2 * Cm)) < Cm;
    if (first ? drop_a != 0 : drop_v != 0) cat[i] = 0.f;
  }
}
__global__ void drop_inplace_kernel(float* __restrict__ x, int64_t n, float p, uint64_t seed, uint64_t offset) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    x[i] *= drop_factor(p, seed, offset, RS_FLAT, (uint64_t)i);
}
// d_pre = d_hid where the saved hidden unit is alive (ReLU on and not dropped: hid > 0), times 1/keep; in place
__global__ void relu_drop_bwd_kernel(float* __restrict__ d, const float* __restrict__ hid, int64_t n, float inv_keep) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    d[i] = hid[i] > 0.f ? d[i] * inv_keep : 0.f;
}
// gated blend backward (fusion.py:433-434): fused = g·a + (1-g)·v, g = sigmoid(z).  Per clip: d_z = g(1-g)·<d_fused, a - v>;
// d_cat = [g·d_fused | (1-g)·d_fused] (written; the gate MLP's contribution is accumulated onto it afterwards).
__global__ void __launch_bounds__(256) gate_blend_bwd_kernel(const float* __restrict__ cat, const float* __restrict__ z,
                                                             const float* __restrict__ d_fused, float* __restrict__ d_cat,
                                                             float* __restrict__ d_z, int Cm) {
  __shared__ float red[8];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float g = 1.f / (1.f + expf(-z[b]));
  const float* a = cat + (int64_t)b * 2 * Cm;
  const float* v = a + Cm;
  const float* df = d_fused + (int64_t)b * Cm;
  float s = 0.f;
  for (int e = tid; e < Cm; e += blockDim.x) {
    const float d = df[e];
    s = fmaf(d, a[e] - v[e], s);
    d_cat[(int64_t)b * 2 * Cm + e] = g * d;
    d_cat[(int64_t)b * 2 * Cm + Cm + e] = (1.f - g) * d;
  }
  s = warp_sum(s);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) t += red[w2];
    d_z[b] = t * g * (1.f - g);
  }
}
// fused[b, e] = g·a + (1-g)·v (recomputed for the classifier's weight gradient)
__global__ void gate_blend_kernel(const float* __restrict__ cat, const float* __restrict__ z, float* __restrict__ fused, int64_t n, int Cm) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / Cm;
    const int e = (int)(i - b * Cm);
    const float g = 1.f / (1.f + expf(-z[b]));
    fused[i] = g * cat[b * 2 * Cm + e] + (1.f - g) * cat[b * 2 * Cm + Cm + e];
  }
}
// out[b, l, :] = d_pooled[b, :] / L: the temporal mean's backward
__global__ void mean_bwd_kernel(const float* __restrict__ d_pooled, float* __restrict__ out, int64_t n, int L, int D) {
  const float inv = 1.f / (float)L;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / ((int64_t)L * D);
    out[i] = d_pooled[b * D + (i % D)] * inv;
  }
}
template <typename TIO>
static int flat_forward_impl(const flat_head_config& c, const float* params, const TIO* xv, const TIO* xa, float* logits, float* ws,
                             cudaStream_t st) {
  FlatLayout pl;
  flat_layout(c, pl);
  FlatWs w;
  flat_ws(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  const int B = c.B, Cm = c.common_dim, A = c.align_dim;
  launch_mean<TIO, float>(st, xa, ws + w.a_pool, B, c.Ta, c.Da);
  launch_mean<TIO, float>(st, xv, ws + w.v_pool, B, c.T, c.Dv);
  g_tc_gemm = c.precision == XATTN_PREC_TF32;   // (linear_fwd: tcgen05 tf32 where the dimensions allow; launch_gemm stays SIMT)
  g_tc_status = XATTN_OK;
  const float* a_in = ws + w.a_pool;
  const float* v_in = ws + w.v_pool;
  int in_a = c.Da, in_v = c.Dv;
  if (A > 0) {  // ClipStyleAlignment projections, fusion.py:138-139
    linear_fwd<float>(st, a_in, P(FP_ALIGN_A_W), P(FP_ALIGN_A_B), ws + w.a_al, B, A, c.Da);
    linear_fwd<float>(st, v_in, P(FP_ALIGN_V_W), P(FP_ALIGN_V_B), ws + w.v_al, B, A, c.Dv);
    a_in = ws + w.a_al; v_in = ws + w.v_al; in_a = in_v = A;
  }
  // audio_proj / video_proj written side by side: cat = [a | v] (fusion.py:421-425)
  launch_gemm<float, float>(st, a_in, in_a, 1, P(FP_AUDIO_W), 1, in_a, ws + w.cat, 2 * Cm, B, Cm, in_a, P(FP_AUDIO_B), 0);
  launch_gemm<float, float>(st, v_in, in_v, 1, P(FP_VIDEO_W), 1, in_v, ws + w.cat + Cm, 2 * Cm, B, Cm, in_v, P(FP_VIDEO_B), 0);
  if (c.training && c.gated && (c.drop_audio || c.drop_video)) {   // ModalityDropout (fusion.py:47-53): a whole modality set to zero
    zero_halves_kernel<<<ew_blocks((int64_t)B * 2 * Cm), 256, 0, st>>>(ws + w.cat, (int64_t)B * 2 * Cm, Cm, c.drop_audio, c.drop_video);
    count_launch();
  }
  linear_fwd<float>(st, ws + w.cat, P(FP_H0_W), P(FP_H0_B), ws + w.hid, B, Cm, 2 * Cm, true);
  if (c.training && c.p_drop > 0.f) {   // nn.Dropout(0.2) behind the ReLU (fusion.py:246 / :255)
    drop_inplace_kernel<<<ew_blocks((int64_t)B * Cm), 256, 0, st>>>(ws + w.hid, (int64_t)B * Cm, c.p_drop, c.seed, c.offset);
    count_launch();
  }
  if (!c.gated) {
    linear_fwd<float>(st, ws + w.hid, P(FP_H1_W), P(FP_H1_B), logits, B, c.C, Cm);
  } else {
    linear_fwd<float>(st, ws + w.hid, P(FP_H1_W), P(FP_H1_B), ws + w.z, B, 1, Cm);
    flat_gate_cls_kernel<<<B, 256, sizeof(float) * Cm, st>>>(ws + w.cat, ws + w.z, P(FP_CLS_W), P(FP_CLS_B), logits, Cm, c.C);
    count_launch();
  }
  g_tc_gemm = false;
  XF_LAUNCH_CHECK();
  return g_tc_status;
}

// Backward of flat_forward_impl from the activations it left in the workspace (a_pool, v_pool, a_al, v_al, cat, hid, z).
static int flat_backward_impl(const flat_head_config& c, const float* params, const float* dlogits, float* grads, float* d_xv, float* d_xa,
                              float* ws, cudaStream_t st) {
  FlatLayout pl;
  flat_layout(c, pl);
  FlatWs w;
  flat_ws(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  auto G = [&](int id) { return grads + pl.off[id]; };
  const int B = c.B, Cm = c.common_dim, A = c.align_dim, C = c.C;
  const float inv_keep = (c.training && c.p_drop > 0.f) ? 1.f / (1.f - c.p_drop) : 1.f;
  XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * pl.total, st));
  g_tc_gemm = c.precision == XATTN_PREC_TF32;
  g_tc_status = XATTN_OK;
  const float* a_in = A > 0 ? ws + w.a_al : ws + w.a_pool;
  const float* v_in = A > 0 ? ws + w.v_al : ws + w.v_pool;
  const int in_a = A > 0 ? A : c.Da, in_v = A > 0 ? A : c.Dv;
  float *d_cat = ws + w.d_cat, *d_hid = ws + w.d_hid;
  const int64_t nh = (int64_t)B * Cm;
  if (!c.gated) {   // logits = W1·drop(relu(W0·cat + b0)) + b1   (fusion.py:243-248)
    linear_bwd_dw<float>(st, dlogits, C, ws + w.hid, Cm, G(FP_H1_W), B, C, Cm);
    launch_colsum(st, dlogits, C, B, C, G(FP_H1_B));
    linear_bwd_dx(st, dlogits, C, P(FP_H1_W), d_hid, B, C, Cm, false);
    relu_drop_bwd_kernel<<<ew_blocks(nh), 256, 0, st>>>(d_hid, ws + w.hid, nh, inv_keep);
    count_launch();
    linear_bwd_dw<float>(st, d_hid, Cm, ws + w.cat, 2 * Cm, G(FP_H0_W), B, Cm, 2 * Cm);
    launch_colsum(st, d_hid, Cm, B, Cm, G(FP_H0_B));
    linear_bwd_dx(st, d_hid, Cm, P(FP_H0_W), d_cat, B, Cm, 2 * Cm, false);
  } else {          // logits = Wc·(g·a + (1-g)·v) + bc, g = sigmoid(w1·drop(relu(W0·cat + b0)) + b1)   (fusion.py:252-259, :431-435)
    float *fused = d_hid, *d_fused = ws + w.d_fused, *d_z = ws + w.d_z;   // (d_hid is free until the gate MLP's backward)
    gate_blend_kernel<<<ew_blocks(nh), 256, 0, st>>>(ws + w.cat, ws + w.z, fused, nh, Cm);
    count_launch();
    linear_bwd_dw<float>(st, dlogits, C, fused, Cm, G(FP_CLS_W), B, C, Cm);
    launch_colsum(st, dlogits, C, B, C, G(FP_CLS_B));
    linear_bwd_dx(st, dlogits, C, P(FP_CLS_W), d_fused, B, C, Cm, false);
    gate_blend_bwd_kernel<<<B, 256, 0, st>>>(ws + w.cat, ws + w.z, d_fused, d_cat, d_z, Cm);
    count_launch();
    linear_bwd_dw<float>(st, d_z, 1, ws + w.hid, Cm, G(FP_H1_W), B, 1, Cm);
    launch_colsum(st, d_z, 1, B, 1, G(FP_H1_B));
    linear_bwd_dx(st, d_z, 1, P(FP_H1_W), d_hid, B, 1, Cm, false);
    relu_drop_bwd_kernel<<<ew_blocks(nh), 256, 0, st>>>(d_hid, ws + w.hid, nh, inv_keep);
    count_launch();
    linear_bwd_dw<float>(st, d_hid, Cm, ws + w.cat, 2 * Cm, G(FP_H0_W), B, Cm, 2 * Cm);
    launch_colsum(st, d_hid, Cm, B, Cm, G(FP_H0_B));
    linear_bwd_dx(st, d_hid, Cm, P(FP_H0_W), d_cat, B, Cm, 2 * Cm, true);
    if (c.training && (c.drop_audio || c.drop_video)) {   // a dropped modality was replaced by zeros: no gradient flows back
      zero_halves_kernel<<<ew_blocks((int64_t)B * 2 * Cm), 256, 0, st>>>(d_cat, (int64_t)B * 2 * Cm, Cm, c.drop_audio, c.drop_video);
      count_launch();
    }
  }
  // audio_proj / video_proj (fusion.py:421-422): cat = [a | v], row stride 2·Cm
  linear_bwd_dw<float>(st, d_cat, 2 * Cm, a_in, in_a, G(FP_AUDIO_W), B, Cm, in_a);
  launch_colsum(st, d_cat, 2 * Cm, B, Cm, G(FP_AUDIO_B));
  linear_bwd_dw<float>(st, d_cat + Cm, 2 * Cm, v_in, in_v, G(FP_VIDEO_W), B, Cm, in_v);
  launch_colsum(st, d_cat + Cm, 2 * Cm, B, Cm, G(FP_VIDEO_B));
  if (A > 0) {   // ClipStyleAlignment projections (fusion.py:138-139)
    linear_bwd_dx(st, d_cat, 2 * Cm, P(FP_AUDIO_W), ws + w.d_a_al, B, Cm, A, false);
    linear_bwd_dx(st, d_cat + Cm, 2 * Cm, P(FP_VIDEO_W), ws + w.d_v_al, B, Cm, A, false);
    linear_bwd_dw<float>(st, ws + w.d_a_al, A, ws + w.a_pool, c.Da, G(FP_ALIGN_A_W), B, A, c.Da);
    launch_colsum(st, ws + w.d_a_al, A, B, A, G(FP_ALIGN_A_B));
    linear_bwd_dw<float>(st, ws + w.d_v_al, A, ws + w.v_pool, c.Dv, G(FP_ALIGN_V_W), B, A, c.Dv);
    launch_colsum(st, ws + w.d_v_al, A, B, A, G(FP_ALIGN_V_B));
    if (d_xa) linear_bwd_dx(st, ws + w.d_a_al, A, P(FP_ALIGN_A_W), ws + w.d_a_pool, B, A, c.Da, false);
    if (d_xv) linear_bwd_dx(st, ws + w.d_v_al, A, P(FP_ALIGN_V_W), ws + w.d_v_pool, B, A, c.Dv, false);
  } else {
    if (d_xa) linear_bwd_dx(st, d_cat, 2 * Cm, P(FP_AUDIO_W), ws + w.d_a_pool, B, Cm, c.Da, false);
    if (d_xv) linear_bwd_dx(st, d_cat + Cm, 2 * Cm, P(FP_VIDEO_W), ws + w.d_v_pool, B, Cm, c.Dv, false);
  }
  if (d_xa) {
    const int64_t n = (int64_t)B * c.Ta * c.Da;
    mean_bwd_kernel<<<ew_blocks(n), 256, 0, st>>>(ws + w.d_a_pool, d_xa, n, c.Ta, c.Da);
    count_launch();
  }
  if (d_xv) {
    const int64_t n = (int64_t)B * c.T * c.Dv;
    mean_bwd_kernel<<<ew_blocks(n), 256, 0, st>>>(ws + w.d_v_pool, d_xv, n, c.T, c.Dv);
    count_launch();
  }
  g_tc_gemm = false;
  XF_LAUNCH_CHECK();
  return g_tc_status;
}

// ---- precision = BF16: the same op sequence with every Linear of width >= 128 on the tcgen05 projection GEMM (gemm_tc.cu) ----
__global__ void __launch_bounds__(256) cvt_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
  const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 4 <= n) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    *reinterpret_cast<uint2*>(y + i) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
  } else {
    for (int64_t k = i; k < n; ++k) y[k] = __float2bfloat16_rn(x[k]);
  }
}

// Last layer(s), one warp per clip (fusion.py:427-435): concat: logits = W1·hid + b1.  gated: z = w1·hid + b1, g = sigmoid(z),
// fused = g·a + (1-g)·v with [a | v] = cat (bf16), logits = Wc·fused + bc.
__global__ void __launch_bounds__(256) flat_out_kernel(const float* __restrict__ hid, const __nv_bfloat16* __restrict__ cat,
                                                       const float* __restrict__ w1, const float* __restrict__ b1,
                                                       const float* __restrict__ wc, const float* __restrict__ bc,
                                                       float* __restrict__ logits, int B, int Cm, int C, int gated) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  const float* h = hid + (int64_t)b * Cm;
  if (!gated) {
    for (int c = 0; c < C; ++c) {
      float s = 0.f;
      for (int k = lane; k < Cm; k += 32) s = fmaf(__ldg(w1 + (int64_t)c * Cm + k), h[k], s);
      s = warp_sum(s);
      if (lane == 0) logits[(int64_t)b * C + c] = s + __ldg(b1 + c);
    }
    return;
  }
  float z = 0.f;
  for (int k = lane; k < Cm; k += 32) z = fmaf(__ldg(w1 + k), h[k], z);
  z = warp_sum(z) + __ldg(b1);
  const float g = 1.f / (1.f + expf(-z));
  const __nv_bfloat16* cr = cat + (int64_t)b * 2 * Cm;
  for (int c = 0; c < C; ++c) {
    float s = 0.f;
    for (int k = lane; k < Cm; k += 32) {
      const float f = g * __bfloat162float(cr[k]) + (1.f - g) * __bfloat162float(cr[Cm + k]);
      s = fmaf(__ldg(wc + (int64_t)c * Cm + k), f, s);
    }
    s = warp_sum(s);
    if (lane == 0) logits[(int64_t)b * C + c] = s + __ldg(bc + c);
  }
}

static int flat_forward_tc(const flat_head_config& c, const float* params, const __nv_bfloat16* xv, const __nv_bfloat16* xa, float* logits,
                           float* ws, cudaStream_t st) {
  FlatLayout pl;
  flat_layout(c, pl);
  FlatWs w;
  flat_ws(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  __nv_bfloat16* wbf = reinterpret_cast<__nv_bfloat16*>(ws + w.wbf);
  auto PB = [&](int id) { return wbf + pl.off[id]; };
  auto BF = [&](int64_t off) { return reinterpret_cast<__nv_bfloat16*>(ws + off); };
  const int B = c.B, Cm = c.common_dim, A = c.align_dim;
  for (int id : {FP_ALIGN_A_W, FP_ALIGN_V_W, FP_AUDIO_W, FP_VIDEO_W, FP_H0_W})
    if (pl.off[id] % 8 != 0) XF_FAIL(XATTN_ERR_ALIGN, "flat heads, precision = BF16: weight %d is not 16-byte aligned in the bf16 copy", id);
  cvt_f32_bf16_kernel<<<(unsigned)((pl.total / 4 + 255) / 256 + 1), 256, 0, st>>>(params, wbf, pl.total);
  count_launch();
  launch_mean<__nv_bfloat16, __nv_bfloat16>(st, xa, BF(w.a_pool), B, c.Ta, c.Da);
  launch_mean<__nv_bfloat16, __nv_bfloat16>(st, xv, BF(w.v_pool), B, c.T, c.Dv);
  const __nv_bfloat16* a_in = BF(w.a_pool);
  const __nv_bfloat16* v_in = BF(w.v_pool);
  int in_a = c.Da, in_v = c.Dv;
  if (A > 0) {  // ClipStyleAlignment projections, fusion.py:138-139
    XF_TRY(gemm_tc_bf16(st, a_in, PB(FP_ALIGN_A_W), P(FP_ALIGN_A_B), BF(w.a_al), true, A, B, A, c.Da));
    XF_TRY(gemm_tc_bf16(st, v_in, PB(FP_ALIGN_V_W), P(FP_ALIGN_V_B), BF(w.v_al), true, A, B, A, c.Dv));
    a_in = BF(w.a_al); v_in = BF(w.v_al); in_a = in_v = A;
  }
  // audio_proj / video_proj written side by side: cat = [a | v] (fusion.py:421-425)
  XF_TRY(gemm_tc_bf16(st, a_in, PB(FP_AUDIO_W), P(FP_AUDIO_B), BF(w.cat), true, 2 * Cm, B, Cm, in_a));
  XF_TRY(gemm_tc_bf16(st, v_in, PB(FP_VIDEO_W), P(FP_VIDEO_B), BF(w.cat) + Cm, true, 2 * Cm, B, Cm, in_v));
  XF_TRY(gemm_tc_bf16(st, BF(w.cat), PB(FP_H0_W), P(FP_H0_B), ws + w.hid, false, Cm, B, Cm, 2 * Cm, nullptr, nullptr, nullptr, 0, true));
  flat_out_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.hid, BF(w.cat), P(FP_H1_W), P(FP_H1_B), P(FP_CLS_W), P(FP_CLS_B), logits, B, Cm,
                                              c.C, c.gated);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// ---- ClipStyleAlignment loss (fusion.py:141-149): 0.5·(CE(s·â v̂ᵀ, arange) + CE(s·v̂ âᵀ, arange)), s = min(exp(logit_scale), 100) ----
struct AlignWs {
  int64_t a_pool, v_pool, a_al, v_al, L, terms, total;   // floats
  int64_t inv_a, inv_v, d_an, d_vn, d_a_pool, d_v_pool, ds;   // backward (inside the same total: the forward ignores them)
};
static void align_ws(const flat_head_config& c, AlignWs& w) {
  const int64_t B = c.B, A = c.align_dim > 0 ? c.align_dim : 1;
  int64_t off = 0;
  auto take = [&](int64_t n) { int64_t o = off; off += al4(n); return o; };
  w.a_pool = take(B * c.Da); w.v_pool = take(B * c.Dv); w.a_al = take(B * A); w.v_al = take(B * A);
  w.L = take(B * B); w.terms = take(2 * B);
  w.inv_a = take(B); w.inv_v = take(B); w.d_an = take(B * A); w.d_vn = take(B * A);
  w.d_a_pool = take(B * c.Da); w.d_v_pool = take(B * c.Dv); w.ds = take(4);
  w.total = off;
}

// x_r /= max(||x_r||, 1e-12)   (F.normalize, fusion.py:141-142); one warp per row
__global__ void __launch_bounds__(256) l2norm_rows_kernel(float* __restrict__ x, int rows, int D, float* __restrict__ inv_out = nullptr) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= rows) return;
  float* p = x + (int64_t)r * D;
  float s = 0.f;
  for (int k = lane; k < D; k += 32) s = fmaf(p[k], p[k], s);
  s = warp_sum(s);
  const float inv = 1.f / fmaxf(sqrtf(s), 1e-12f);
  for (int k = lane; k < D; k += 32) p[k] *= inv;
  if (inv_out != nullptr && lane == 0) inv_out[r] = inv;
}

// out[i] = logsumexp_j(s·L[i,j]) - s·L[i,i]: the cross-entropy term of row i against target i; one warp per row, two passes
__global__ void __launch_bounds__(256) clip_row_terms_kernel(const float* __restrict__ L, int B, const float* __restrict__ logit_scale,
                                                             float* __restrict__ out) {
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= B) return;
  const float s = fminf(expf(__ldg(logit_scale)), 100.f);
  const float* row = L + (int64_t)i * B;
  float mx = -INFINITY;
  for (int j = lane; j < B; j += 32) mx = fmaxf(mx, row[j]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < B; j += 32) sum += expf(s * (row[j] - mx));
  sum = warp_sum(sum);
  if (lane == 0) out[i] = s * mx + logf(sum) - s * row[i];
}

// loss = 0.5·(mean(terms[0:B]) + mean(terms[B:2B])), summed in a fixed order by one CTA
__global__ void __launch_bounds__(256) clip_loss_final_kernel(const float* __restrict__ terms, int B, float* __restrict__ loss) {
  __shared__ float red[256];
  float s = 0.f;
  for (int i = threadIdx.x; i < 2 * B; i += 256) s += terms[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss[0] = 0.5f * red[0] / (float)B;
}

template <typename TIO>
static int alignment_loss_impl(const flat_head_config& c, const float* params, const float* logit_scale, const TIO* xv, const TIO* xa,
                               float* loss, float* ws, cudaStream_t st) {
  FlatLayout pl;
  flat_layout(c, pl);
  AlignWs w;
  align_ws(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  const int B = c.B, A = c.align_dim;
  launch_mean<TIO, float>(st, xa, ws + w.a_pool, B, c.Ta, c.Da);
  launch_mean<TIO, float>(st, xv, ws + w.v_pool, B, c.T, c.Dv);
  g_tc_gemm = c.precision != XATTN_DTYPE_F32;   // tensor-core Linears (tf32 products) where the dimensions allow
  g_tc_status = XATTN_OK;
  linear_fwd<float>(st, ws + w.a_pool, P(FP_ALIGN_A_W), P(FP_ALIGN_A_B), ws + w.a_al, B, A, c.Da);
  linear_fwd<float>(st, ws + w.v_pool, P(FP_ALIGN_V_W), P(FP_ALIGN_V_B), ws + w.v_al, B, A, c.Dv);
  l2norm_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.a_al, B, A);
  l2norm_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.v_al, B, A);
  count_launch(2);
  // the B x B cosine matrix once per direction (the second product is the transpose: rows stay contiguous for the row-wise pass)
  linear_fwd<float>(st, ws + w.a_al, ws + w.v_al, nullptr, ws + w.L, B, B, A);
  clip_row_terms_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.L, B, logit_scale, ws + w.terms);
  linear_fwd<float>(st, ws + w.v_al, ws + w.a_al, nullptr, ws + w.L, B, B, A);
  clip_row_terms_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.L, B, logit_scale, ws + w.terms + B);
  clip_loss_final_kernel<<<1, 256, 0, st>>>(ws + w.terms, B, loss);
  count_launch(3);
  g_tc_gemm = false;
  XF_LAUNCH_CHECK();
  return g_tc_status;
}

// ---- backward of the alignment loss -------------------------------------------------------------------------------------------
// With Z = s·â v̂ᵀ: d loss / d Z_ij = (0.5 / B)·(softmax_j(Z_i·)_j + softmax_i(Z_·j)_i − 2 δ_ij) =: G_ij (times the upstream gradient);
// d â = s·G v̂, d v̂ = s·Gᵀ â, d s = Σ G ⊙ (â v̂ᵀ), d logit_scale = d s · s where exp(logit_scale) < 100 (the clamp), and through
// F.normalize  d x = (d y − y <y, d y>) / ||x||.
// lse[i] = logsumexp_j(s·L[i,j]); one warp per row
__global__ void __launch_bounds__(256) clip_lse_rows_kernel(const float* __restrict__ L, int B, const float* __restrict__ logit_scale,
                                                            float* __restrict__ lse) {
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (i >= B) return;
  const float s = fminf(expf(__ldg(logit_scale)), 100.f);
  const float* row = L + (int64_t)i * B;
  float mx = -INFINITY;
  for (int j = lane; j < B; j += 32) mx = fmaxf(mx, row[j]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < B; j += 32) sum += expf(s * (row[j] - mx));
  sum = warp_sum(sum);
  if (lane == 0) lse[i] = s * mx + logf(sum);
}
// lse[j] = logsumexp_i(s·L[i,j]); one thread per column (coalesced across the threads of a warp)
__global__ void __launch_bounds__(256) clip_lse_cols_kernel(const float* __restrict__ L, int B, const float* __restrict__ logit_scale,
                                                            float* __restrict__ lse) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= B) return;
  const float s = fminf(expf(__ldg(logit_scale)), 100.f);
  float mx = -INFINITY;
  for (int i = 0; i < B; ++i) mx = fmaxf(mx, L[(int64_t)i * B + j]);
  float sum = 0.f;
  for (int i = 0; i < B; ++i) sum += expf(s * (L[(int64_t)i * B + j] - mx));
  lse[j] = s * mx + logf(sum);
}
// L <- G (in place); *ds += Σ G ⊙ L
__global__ void __launch_bounds__(256) clip_grad_kernel(float* __restrict__ L, int B, const float* __restrict__ logit_scale,
                                                        const float* __restrict__ lse_row, const float* __restrict__ lse_col,
                                                        const float* __restrict__ dloss, float* __restrict__ ds) {
  __shared__ float red[8];
  const float s = fminf(expf(__ldg(logit_scale)), 100.f);
  const float cst = __ldg(dloss) * 0.5f / (float)B;
  const int64_t n = (int64_t)B * B;
  float acc = 0.f;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    const int i = (int)(k / B), j = (int)(k - (int64_t)i * B);
    const float l = L[k], z = s * l;
    const float g = cst * (expf(z - lse_row[i]) + expf(z - lse_col[j]) - (i == j ? 2.f : 0.f));
    acc = fmaf(g, l, acc);
    L[k] = g;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w2 = 0; w2 < (int)(blockDim.x >> 5); ++w2) t += red[w2];
    atomicAdd(ds, t);
  }
}
// d <- s·(d − y <y, d>)·inv  (row-wise; y = the normalised row, inv = 1 / ||x||); one warp per row
__global__ void __launch_bounds__(256) l2norm_bwd_kernel(float* __restrict__ d, const float* __restrict__ y, const float* __restrict__ inv,
                                                         int rows, int D, const float* __restrict__ logit_scale) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= rows) return;
  const float s = fminf(expf(__ldg(logit_scale)), 100.f);
  float* dr = d + (int64_t)r * D;
  const float* yr = y + (int64_t)r * D;
  float dot = 0.f;
  for (int k = lane; k < D; k += 32) dot = fmaf(yr[k], dr[k], dot);
  dot = warp_sum(dot);
  const float f = s * inv[r];
  for (int k = lane; k < D; k += 32) dr[k] = f * (dr[k] - yr[k] * dot);
}
__global__ void logit_scale_grad_kernel(const float* __restrict__ ds, const float* __restrict__ logit_scale, float* __restrict__ out) {
  const float e = expf(logit_scale[0]);
  out[0] = e < 100.f ? ds[0] * e : 0.f;   // clamp(max=100): no gradient where it is active
}

template <typename TIO>
static int alignment_backward_impl(const flat_head_config& c, const float* params, const float* logit_scale, const TIO* xv, const TIO* xa,
                                   const float* dloss, float* grads, float* d_logit_scale, float* d_xv, float* d_xa, float* ws,
                                   cudaStream_t st) {
  FlatLayout pl;
  flat_layout(c, pl);
  AlignWs w;
  align_ws(c, w);
  auto P = [&](int id) { return params + pl.off[id]; };
  auto G = [&](int id) { return grads + pl.off[id]; };
  const int B = c.B, A = c.align_dim;
  float *an = ws + w.a_al, *vn = ws + w.v_al, *Lm = ws + w.L, *lse_r = ws + w.terms, *lse_c = ws + w.terms + B;
  XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * pl.total, st));
  XF_CUDA(cudaMemsetAsync(ws + w.ds, 0, sizeof(float) * 4, st));
  XF_CUDA(cudaMemsetAsync(ws + w.d_vn, 0, sizeof(float) * (size_t)B * A, st));
  launch_mean<TIO, float>(st, xa, ws + w.a_pool, B, c.Ta, c.Da);
  launch_mean<TIO, float>(st, xv, ws + w.v_pool, B, c.T, c.Dv);
  g_tc_gemm = c.precision != XATTN_DTYPE_F32;
  g_tc_status = XATTN_OK;
  linear_fwd<float>(st, ws + w.a_pool, P(FP_ALIGN_A_W), P(FP_ALIGN_A_B), an, B, A, c.Da);
  linear_fwd<float>(st, ws + w.v_pool, P(FP_ALIGN_V_W), P(FP_ALIGN_V_B), vn, B, A, c.Dv);
  l2norm_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(an, B, A, ws + w.inv_a);
  l2norm_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(vn, B, A, ws + w.inv_v);
  linear_fwd<float>(st, an, vn, nullptr, Lm, B, B, A);
  clip_lse_rows_kernel<<<(B + 7) / 8, 256, 0, st>>>(Lm, B, logit_scale, lse_r);
  clip_lse_cols_kernel<<<(B + 255) / 256, 256, 0, st>>>(Lm, B, logit_scale, lse_c);
  clip_grad_kernel<<<ew_blocks((int64_t)B * B), 256, 0, st>>>(Lm, B, logit_scale, lse_r, lse_c, dloss, ws + w.ds);
  count_launch(5);
  linear_bwd_dx(st, Lm, B, vn, ws + w.d_an, B, B, A, false);            // d â / s = G v̂
  linear_bwd_dw<float>(st, Lm, B, an, A, ws + w.d_vn, B, B, A);        // d v̂ / s = Gᵀ â  (into the zeroed buffer)
  l2norm_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.d_an, an, ws + w.inv_a, B, A, logit_scale);
  l2norm_bwd_kernel<<<(B + 7) / 8, 256, 0, st>>>(ws + w.d_vn, vn, ws + w.inv_v, B, A, logit_scale);
  logit_scale_grad_kernel<<<1, 1, 0, st>>>(ws + w.ds, logit_scale, d_logit_scale);
  count_launch(3);
  linear_bwd_dw<float>(st, ws + w.d_an, A, ws + w.a_pool, c.Da, G(FP_ALIGN_A_W), B, A, c.Da);
  launch_colsum(st, ws + w.d_an, A, B, A, G(FP_ALIGN_A_B));
  linear_bwd_dw<float>(st, ws + w.d_vn, A, ws + w.v_pool, c.Dv, G(FP_ALIGN_V_W), B, A, c.Dv);
  launch_colsum(st, ws + w.d_vn, A, B, A, G(FP_ALIGN_V_B));
  if (d_xa) {
    linear_bwd_dx(st, ws + w.d_an, A, P(FP_ALIGN_A_W), ws + w.d_a_pool, B, A, c.Da, false);
    const int64_t n = (int64_t)B * c.Ta * c.Da;
    mean_bwd_kernel<<<ew_blocks(n), 256, 0, st>>>(ws + w.d_a_pool, d_xa, n, c.Ta, c.Da);
    count_launch();
  }
  if (d_xv) {
    linear_bwd_dx(st, ws + w.d_vn, A, P(FP_ALIGN_V_W), ws + w.d_v_pool, B, A, c.Dv, false);
    const int64_t n = (int64_t)B * c.T * c.Dv;
    mean_bwd_kernel<<<ew_blocks(n), 256, 0, st>>>(ws + w.d_v_pool, d_xv, n, c.T, c.Dv);
    count_launch();
  }
  g_tc_gemm = false;
  XF_LAUNCH_CHECK();
  return g_tc_status;
}

}  // namespace xf

using namespace xf;

extern "C" XATTN_API int flat_head_param_layout(const flat_head_config* cfg, int64_t* offsets, int64_t* sizes, int64_t* total_floats) {
  XF_TRY(flat_validate(cfg));
  FlatLayout L;
  flat_layout(*cfg, L);
  for (int i = 0; i < FP_COUNT; ++i) {
    if (offsets) offsets[i] = L.off[i];
    if (sizes) sizes[i] = L.size[i];
  }
  if (total_floats) *total_floats = L.total;
  return XATTN_OK;
}

extern "C" XATTN_API int flat_head_workspace_bytes(const flat_head_config* cfg, int64_t* bytes) {
  XF_TRY(flat_validate(cfg));
  if (!bytes) XF_FAIL(XATTN_ERR_NULL, "bytes is NULL");
  FlatWs w;
  flat_ws(*cfg, w);
  *bytes = w.total * (int64_t)sizeof(float);
  return XATTN_OK;
}

extern "C" XATTN_API int flat_head_forward(const flat_head_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                                 float* logits, void* workspace, int64_t workspace_bytes, void* stream) {
  XF_TRY(flat_validate(cfg));
  if (!params || !video_emb || !audio_emb || !logits || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (((uintptr_t)video_emb | (uintptr_t)audio_emb | (uintptr_t)params | (uintptr_t)workspace) & 15)
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  FlatWs w;
  flat_ws(*cfg, w);
  if (workspace_bytes < w.total * (int64_t)sizeof(float))
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)(w.total * 4));
  if (cfg->B == 0) return XATTN_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->precision == XATTN_DTYPE_BF16)
    return flat_forward_tc(*cfg, params, (const __nv_bfloat16*)video_emb, (const __nv_bfloat16*)audio_emb, logits, (float*)workspace, st);
  if (cfg->io_dtype == XATTN_DTYPE_BF16)
    return flat_forward_impl<__nv_bfloat16>(*cfg, params, (const __nv_bfloat16*)video_emb, (const __nv_bfloat16*)audio_emb, logits,
                                            (float*)workspace, st);
  return flat_forward_impl<float>(*cfg, params, (const float*)video_emb, (const float*)audio_emb, logits, (float*)workspace, st);
}

extern "C" XATTN_API int flat_head_alignment_workspace_bytes(const flat_head_config* cfg, int64_t* bytes) {
  XF_TRY(flat_validate(cfg));
  if (!bytes) XF_FAIL(XATTN_ERR_NULL, "bytes is NULL");
  AlignWs w;
  align_ws(*cfg, w);
  *bytes = w.total * (int64_t)sizeof(float);
  return XATTN_OK;
}

extern "C" XATTN_API int flat_head_alignment_loss(const flat_head_config* cfg, const float* params, const float* logit_scale,
                                                  const void* video_emb, const void* audio_emb, float* loss, void* workspace,
                                                  int64_t workspace_bytes, void* stream) {
  XF_TRY(flat_validate(cfg));
  if (!params || !logit_scale || !video_emb || !audio_emb || !loss || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (cfg->align_dim <= 0) XF_FAIL(XATTN_ERR_SHAPE, "the alignment loss needs align_dim > 0 (fusion_align_mode=\"clip\")");
  if (cfg->B <= 0) XF_FAIL(XATTN_ERR_SHAPE, "the alignment loss needs at least one clip");
  if (((uintptr_t)video_emb | (uintptr_t)audio_emb | (uintptr_t)params | (uintptr_t)workspace) & 15)
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  AlignWs w;
  align_ws(*cfg, w);
  if (workspace_bytes < w.total * (int64_t)sizeof(float))
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)(w.total * 4));
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->io_dtype == XATTN_DTYPE_BF16)
    return alignment_loss_impl<__nv_bfloat16>(*cfg, params, logit_scale, (const __nv_bfloat16*)video_emb, (const __nv_bfloat16*)audio_emb, loss,
                                              (float*)workspace, st);
  return alignment_loss_impl<float>(*cfg, params, logit_scale, (const float*)video_emb, (const float*)audio_emb, loss, (float*)workspace, st);
}

extern "C" XATTN_API int flat_head_backward(const flat_head_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                                            const float* dlogits, float* grads, float* d_video_emb, float* d_audio_emb, void* workspace,
                                            int64_t workspace_bytes, void* stream) {
  (void)video_emb; (void)audio_emb;   // (the pooled embeddings the weight gradients need are in the workspace)
  XF_TRY(flat_validate(cfg));
  if (!params || !dlogits || !grads || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (((uintptr_t)params | (uintptr_t)grads | (uintptr_t)workspace | (uintptr_t)d_video_emb | (uintptr_t)d_audio_emb) & 15)
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  if (cfg->precision == XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "flat_head_backward runs with precision f32 / tf32");
  XF_TRY(check_device_sm100());
  FlatWs w;
  flat_ws(*cfg, w);
  if (workspace_bytes < w.total * (int64_t)sizeof(float))
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)(w.total * 4));
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->B == 0) {
    FlatLayout pl;
    flat_layout(*cfg, pl);
    XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * pl.total, st));
    return XATTN_OK;
  }
  return flat_backward_impl(*cfg, params, dlogits, grads, d_video_emb, d_audio_emb, (float*)workspace, st);
}

extern "C" XATTN_API int flat_head_alignment_backward(const flat_head_config* cfg, const float* params, const float* logit_scale,
                                                      const void* video_emb, const void* audio_emb, const float* dloss, float* grads,
                                                      float* d_logit_scale, float* d_video_emb, float* d_audio_emb, void* workspace,
                                                      int64_t workspace_bytes, void* stream) {
  XF_TRY(flat_validate(cfg));
  if (!params || !logit_scale || !video_emb || !audio_emb || !dloss || !grads || !d_logit_scale || !workspace)
    XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (cfg->align_dim <= 0) XF_FAIL(XATTN_ERR_SHAPE, "the alignment loss needs align_dim > 0 (fusion_align_mode=\"clip\")");
  if (cfg->B <= 0) XF_FAIL(XATTN_ERR_SHAPE, "the alignment loss needs at least one clip");
  if (((uintptr_t)video_emb | (uintptr_t)audio_emb | (uintptr_t)params | (uintptr_t)grads | (uintptr_t)workspace | (uintptr_t)d_video_emb |
       (uintptr_t)d_audio_emb) & 15)
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  AlignWs w;
  align_ws(*cfg, w);
  if (workspace_bytes < w.total * (int64_t)sizeof(float))
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)(w.total * 4));
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->io_dtype == XATTN_DTYPE_BF16)
    return alignment_backward_impl<__nv_bfloat16>(*cfg, params, logit_scale, (const __nv_bfloat16*)video_emb, (const __nv_bfloat16*)audio_emb,
                                                  dloss, grads, d_logit_scale, d_video_emb, d_audio_emb, (float*)workspace, st);
  return alignment_backward_impl<float>(*cfg, params, logit_scale, (const float*)video_emb, (const float*)audio_emb, dloss, grads,
                                        d_logit_scale, d_video_emb, d_audio_emb, (float*)workspace, st);
}
