// Layouts of the bf16 path and xattn_pack_weights: derives the bf16 operand set from the flat fp32 parameters.
//  * W_fold = a_in_proj.weight · audio_seq_proj.weight, b_fold = a_in_proj.weight · audio_seq_proj.bias + a_in_proj.bias
//    (fusion.py:378 then :386 are two Linears with nothing in between — SURVEY 3.1 fact 2); folded in fp32, rounded once.
//  * the eight 128x128 attention matrices (slices of in_proj_weight / out_proj.weight, fusion.py:276-281) as bf16 with
//    a 136-element row pitch; * head layer 0 transposed for coalesced reads.
#include "bf16_layout.cuh"
#include "gemm_simt.cuh"
#include "gemm_tc.cuh"
#include "kernels_bf16.cuh"

namespace xf {

static inline int64_t al256(int64_t n) { return (n + 255) & ~(int64_t)255; }

int bf16_supported(const xattn_config& c) {
  if (c.E != BE || c.H != BH) XF_FAIL(XATTN_ERR_SHAPE, "bf16 path is specialised for d_model=128, num_heads=4 (got E=%d H=%d)", c.E, c.H);
  if (c.T > BTP || c.Ta > BTAP)
    XF_FAIL(XATTN_ERR_SHAPE, "bf16 path holds the whole score tile on chip: needs T <= 16 and Ta <= 160 (got T=%d Ta=%d)", c.T, c.Ta);
  if (c.Dv % 64 != 0 || c.Da % 64 != 0) XF_FAIL(XATTN_ERR_SHAPE, "bf16 path needs Dv, Da multiples of 64 (got %d, %d)", c.Dv, c.Da);
  if (c.io_dtype != XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "bf16 path reads bf16 embeddings (io_dtype must be bf16)");
  const int HW = c.head_kind == XATTN_HEAD_GATED ? c.E : c.common_dim;
  if (HW > 256 || HW % 64 != 0 || c.C > 16)
    XF_FAIL(XATTN_ERR_SHAPE, "bf16 path head supports hidden widths 64/128/192/256 and <= 16 classes (got %d, %d)", HW, c.C);
  if (c.use_prior && (c.prior_hidden > 128 || c.prior_dim > 32))
    XF_FAIL(XATTN_ERR_SHAPE, "bf16 path supports prior_hidden <= 128, prior_dim <= 32");
  if (c.pool_mode != XATTN_POOL_MEAN && c.pool_mode != XATTN_POOL_ATTN)
    XF_FAIL(XATTN_ERR_SHAPE, "bf16 path pools by mean or by the learned attention pooling; temporal_pooling=transformer runs in the "
                             "fp32 kernel family (precision f32 / tf32)");
  return XATTN_OK;
}

void compute_pack_bf16(const xattn_config& c, PackBf16& p) {
  const int64_t E = c.E;
  const int64_t HW = c.head_kind == XATTN_HEAD_GATED ? E : c.common_dim;
  int64_t off = 0;
  auto take = [&](int64_t bytes) {
    int64_t o = off;
    off += al256(bytes);
    return o;
  };
  p.Wv = take(E * c.Dv * 2);
  p.Wfold = take(E * c.Da * 2);
  p.bfold = take(E * 4);
  p.Wq1 = take(E * BEP * 2); p.Wk1 = take(E * BEP * 2); p.Wv1 = take(E * BEP * 2); p.Wo1 = take(E * BEP * 2);
  p.Wk2 = take(E * BEP * 2); p.Wv2 = take(E * BEP * 2); p.Wq2 = take(E * BEP * 2); p.Wo2 = take(E * BEP * 2);
  p.W0b = take(HW * 2 * E * 2);
  p.pW0 = take((int64_t)(c.use_prior ? (c.prior_hidden + 7) / 8 * 8 : 0) * 2 * E * 2);
  p.scw = take(2 * E * 4);
  p.Gmix = take(E * 4 * E * 2); p.bmix = take(E * 4);
  p.Fkv = take(8 * E * E * 2); p.bkv = take(8 * E * 4);
  p.wcv = take((4 * E + 4) * 4);
  p.scratch = take((E * c.Da > 8 * E * E ? E * c.Da : 8 * E * E) * 4);
  p.total = off;
}

void compute_ws_bf16(const xattn_config& c, WsBf16& w) {
  const int64_t B = c.B, E = c.E;
  int64_t off = 0;
  auto take = [&](int64_t bytes) {
    int64_t o = off;
    off += al256(bytes);
    return o;
  };
  w.v0 = take(B * c.T * E * 4);
  w.a0 = take(B * c.Ta * BOP * 2);
  w.qt = take(B * 64 * BOP * 2);
  w.meanv = take(B * E * 4);
  w.tqv = take(B * 16 * 4);
  w.tkv = take(B * 16 * 4);
  w.u = take(B * 64 * E * 2);
  w.tok = take(B * BTAP * 2 * 4);
  w.acol = take(B * kColsumSlots * E * 4);
  w.pc = take(B * 4 * 4);
  w.kt = take(B * 64 * BOP * 2);
  w.vt = take(B * 64 * BOP * 2);
  w.cv = take(B * 64 * 4);
  w.pooled = take(B * 2 * E * 4);
  w.poolp = take(B * 8 * E * 4);
  w.v1 = take(B * 16 * E * 2);
  w.xhat = take(c.pool_mode == XATTN_POOL_ATTN ? B * c.Ta * E * 2 : 0);
  w.total = off + 256;
}

int packed_bytes_bf16(const xattn_config& c, int64_t* bytes) {
  if (c.E != BE || c.H != BH) {   // unsupported configs never reach the bf16 kernels; a token size keeps callers simple
    *bytes = 256;
    return XATTN_OK;
  }
  PackBf16 p;
  compute_pack_bf16(c, p);
  *bytes = p.total;
  return XATTN_OK;
}

int workspace_bytes_bf16(const xattn_config& c, int64_t* bytes) {
  XF_TRY(bf16_supported(c));
  WsBf16 w;
  compute_ws_bf16(c, w);
  *bytes = w.total;
  return XATTN_OK;
}

// dst[r, 0..cols) (pitch dpitch, bf16) = src[r, 0..cols) (pitch spitch, fp32); pad columns zeroed.
__global__ void cvt_rows_bf16_kernel(const float* __restrict__ src, int64_t spitch, __nv_bfloat16* __restrict__ dst,
                                     int64_t dpitch, int rows, int cols, float scale) {
  const int64_t n = (int64_t)rows * dpitch;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / dpitch, cidx = i - r * dpitch;
    dst[i] = __float2bfloat16(cidx < cols ? scale * src[r * spitch + cidx] : 0.f);
  }
}
// out[n] = alpha · sum_{d<D} x[d]·M[d*rs + n*cs]  (+ beta·add[n])      — the bias / vector halves of the folded projections
__global__ void fold_vec_kernel(const float* __restrict__ x, const float* __restrict__ Mx, int64_t rs, int64_t cs, int D, int N,
                                float alpha, const float* __restrict__ add, float beta, float* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float s = 0.f;
  for (int d = 0; d < D; ++d) s = fmaf(x[d], Mx[d * rs + n * cs], s);
  out[n] = alpha * s + (add != nullptr ? beta * add[n] : 0.f);
}
// W[N,K] (fp32, row pitch K) -> bf16 in mma.m16n8k16 B-fragment order: for n-tile nt (8 rows), k-step pair kp (32 columns) and
// lane (g = lane / 4, t = lane % 4) the four registers {b0, b1 of k-step 2kp, b0, b1 of k-step 2kp+1} are one uint4 at
// [(nt * K/32 + kp) * 32 + lane]: a warp's fetch is 512 contiguous bytes instead of 8 rows x 16 bytes.  Rows past N repeat N-1.
__global__ void frag_pack_kernel(const float* __restrict__ W, int N, int K, uint4* __restrict__ out) {
  const int ntiles = (N + 7) / 8, kps = K / 32;
  const int total = ntiles * kps * 32;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int lane = i & 31, kp = (i >> 5) % kps, nt = (i >> 5) / kps, g = lane >> 2, t = lane & 3;
    const float* w = W + (int64_t)min(nt * 8 + g, N - 1) * K + kp * 32 + 2 * t;
    auto pk = [&](int k) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(w[k], w[k + 1]);
      return *reinterpret_cast<const uint32_t*>(&h);
    };
    out[i] = make_uint4(pk(0), pk(8), pk(16), pk(24));
  }
}
// out[e] = W_in[e,:]·b_seq + b_in[e]
__global__ void fold_bias_kernel(const float* __restrict__ w_in, const float* __restrict__ b_seq, const float* __restrict__ b_in,
                                 float* __restrict__ out, int E) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  float s = b_in[e];
  for (int m = 0; m < E; ++m) s = fmaf(w_in[(int64_t)e * E + m], b_seq[m], s);
  out[e] = s;
}
int pack_weights_bf16(const xattn_config& c, const float* params, void* packed_v, cudaStream_t st) {
  if (c.E != BE || c.H != BH) XF_FAIL(XATTN_ERR_SHAPE, "bf16 path is specialised for d_model=128, num_heads=4 (got E=%d H=%d)", c.E, c.H);
  ParamLayout pl;
  compute_param_layout(c, pl);
  PackBf16 pk;
  compute_pack_bf16(c, pk);
  uint8_t* packed = reinterpret_cast<uint8_t*>(packed_v);
  auto P = [&](int id) { return params + pl.off[id]; };
  auto BF = [&](int64_t off) { return reinterpret_cast<__nv_bfloat16*>(packed + off); };
  const int E = c.E;
  auto cvt = [&](const float* src, int64_t spitch, int64_t dst_off, int64_t dpitch, int rows, int cols, float scale = 1.f) {
    const int64_t n = (int64_t)rows * dpitch;
    const int blocks = (int)((n + 255) / 256 < 1024 ? (n + 255) / 256 : 1024);
    cvt_rows_bf16_kernel<<<blocks, 256, 0, st>>>(src, spitch, BF(dst_off), dpitch, rows, cols, scale);
    count_launch();
  };
  auto fvec = [&](const float* x, const float* Mx, int64_t rs, int64_t cs, int D, int N, float alpha, const float* add, float beta,
                  float* out) {
    fold_vec_kernel<<<(N + 127) / 128, 128, 0, st>>>(x, Mx, rs, cs, D, N, alpha, add, beta, out);
    count_launch();
  };
  cvt(P(XP_V_IN_W), c.Dv, pk.Wv, c.Dv, E, c.Dv);
  // W_fold[e,k] = sum_m W_in[e,m] W_seq[m,k]   (fp32 SIMT GEMM: A(m,k)=W_in[e*E+m], B(k,n)=W_seq[m*Da+n])
  float* scratch = reinterpret_cast<float*>(packed + pk.scratch);
  launch_gemm<float, float>(st, P(XP_A_IN_W), E, 1, P(XP_SEQ_W), c.Da, 1, scratch, c.Da, E, c.Da, E, nullptr, 0);
  cvt(scratch, c.Da, pk.Wfold, c.Da, E, c.Da);
  fold_bias_kernel<<<(E + 127) / 128, 128, 0, st>>>(P(XP_A_IN_W), P(XP_SEQ_B), P(XP_A_IN_B), reinterpret_cast<float*>(packed + pk.bfold), E);
  count_launch();
  const float* in1 = P(XP_V2A_IN_W);
  const float* in2 = P(XP_A2V_IN_W);
  cvt(in1, E, pk.Wq1, BEP, E, E);
  cvt(in1 + (int64_t)E * E, E, pk.Wk1, BEP, E, E);
  cvt(in1 + (int64_t)2 * E * E, E, pk.Wv1, BEP, E, E);
  cvt(P(XP_V2A_OUT_W), E, pk.Wo1, BEP, E, E);
  cvt(in2, E, pk.Wq2, BEP, E, E);
  cvt(in2 + (int64_t)E * E, E, pk.Wk2, BEP, E, E);
  cvt(in2 + (int64_t)2 * E * E, E, pk.Wv2, BEP, E, E);
  cvt(P(XP_A2V_OUT_W), E, pk.Wo2, BEP, E, E);
  {
    // ---- the video-side update as two GEMMs: per-head folds in fp32, rounded to bf16 once (fusion.py:394-398) ----
    const float scale = 1.0f / sqrtf((float)BDH);
    const float *Wv1 = in1 + (int64_t)2 * E * E, *bv1 = P(XP_V2A_IN_B) + 2 * E, *Wo1 = P(XP_V2A_OUT_W), *bo1 = P(XP_V2A_OUT_B);
    const float *Wq2 = in2, *bq2 = P(XP_A2V_IN_B), *Wk2 = in2 + (int64_t)E * E, *bk2 = bq2 + E, *Wv2 = in2 + (int64_t)2 * E * E,
                *bv2 = bq2 + 2 * E, *Wo2 = P(XP_A2V_OUT_W), *bo2 = P(XP_A2V_OUT_B);
    // G[n][128h + e] = sum_d Wo1[n][32h+d]·Wv1[32h+d][e]: O_h = U_h·Wv1_h^T + bv1_h (softmax rows sum to one), then out_proj
    for (int h = 0; h < BH; ++h)
      launch_gemm<float, float>(st, Wo1 + 32 * h, E, 1, Wv1 + (int64_t)32 * h * E, E, 1, scratch + h * E, 4 * E, E, E, BDH, nullptr, 0);
    cvt(scratch, 4 * E, pk.Gmix, 4 * E, E, 4 * E);
    fvec(bv1, Wo1, 1, E, E, E, 1.f, bo1, 1.f, reinterpret_cast<float*>(packed + pk.bmix));
    // Fk_h[e'][k] = scale·sum_d Wq2[32h+d][e']·Wk2[32h+d][k]   (Kt_h = scale·K2_h·Wq2_h with K2_h = v1·Wk2_h^T + bk2_h)
    // Fv_h[e'][k] =       sum_d Wo2[e'][32h+d]·Wv2[32h+d][k]   (Vt_h = V2_h·Wo2_h^T + bo2/4)
    float* bkv = reinterpret_cast<float*>(packed + pk.bkv);
    float* wcv = reinterpret_cast<float*>(packed + pk.wcv);
    for (int h = 0; h < BH; ++h) {
      const int64_t ho = (int64_t)32 * h * E;
      launch_gemm<float, float>(st, Wq2 + ho, 1, E, Wk2 + ho, E, 1, scratch + (int64_t)h * E * E, E, E, E, BDH, nullptr, 0);
      launch_gemm<float, float>(st, Wo2 + 32 * h, E, 1, Wv2 + ho, E, 1, scratch + (int64_t)(4 + h) * E * E, E, E, E, BDH, nullptr, 0);
      fvec(bk2 + 32 * h, Wq2 + ho, E, 1, BDH, E, scale, nullptr, 0.f, bkv + h * E);
      fvec(bv2 + 32 * h, Wo2 + 32 * h, 1, E, BDH, E, 1.f, bo2, 0.25f, bkv + (4 + h) * E);
      fvec(bq2 + 32 * h, Wk2 + ho, E, 1, BDH, E, scale, nullptr, 0.f, wcv + h * E);            // wc_h[k]
      fvec(bk2 + 32 * h, bq2 + 32 * h, 1, 0, BDH, 1, scale, nullptr, 0.f, wcv + 4 * E + h);    // cc_h
    }
    cvt(scratch, E, pk.Fkv, E, 4 * E, E, scale);
    cvt(scratch + (int64_t)4 * E * E, E, pk.Fkv + (int64_t)4 * E * E * 2, E, 4 * E, E);
  }
  const int HW = c.head_kind == XATTN_HEAD_GATED ? E : c.common_dim;
  frag_pack_kernel<<<32, 256, 0, st>>>(P(XP_HEAD0_W), HW, 2 * E, reinterpret_cast<uint4*>(packed + pk.W0b));
  count_launch();
  XF_CUDA(cudaMemsetAsync(packed + pk.scw, 0, 2 * E * 4, st));
  if (c.use_prior) {
    frag_pack_kernel<<<32, 256, 0, st>>>(P(XP_PRIOR0_W), c.prior_hidden, 2 * E, reinterpret_cast<uint4*>(packed + pk.pW0));
    count_launch();
    XF_CUDA(cudaMemcpyAsync(packed + pk.scw, P(XP_AK_W), E * 4, cudaMemcpyDeviceToDevice, st));           // a_key_bias.weight[:E]
    XF_CUDA(cudaMemcpyAsync(packed + pk.scw + E * 4, P(XP_AQ_W), E * 4, cudaMemcpyDeviceToDevice, st));   // a_query_bias.weight[:E]
  }
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
