// Buffers and kernel entry points of the bf16 tensor-core inference path (precision == XATTN_DTYPE_BF16).
//
// Data flow per clip (reference src/models/fusion.py:372-411), B clips per launch:
//   gemm_tc   v0  = video_emb · Wv^T + b                      [T,128]  fp32           :372
//   gemm_tc   a0  = audio_emb · (W_in·W_seq)^T + b_fold         [Ta,128] bf16                      :378 + :386 folded
//               + epilogue extras: token scores (w_ak·a0_j, w_aq·a0_j) and per-clip column sums of a0 (-> mean_t a0, :180)
//   video_pre   Qt_h = (scale·(v0·Wq1^T + bq1))_h · Wk1_h       [64,128] bf16 — so that  Q_h·K_h^T = Qt_h · a0^T   (:394)
//               + mean_t v0, w_vq·v0_i, w_vk·v0_i  (emotion prior inputs, :171-181); tcgen05, 8 clips per M=128 tile
//   prior       prior_net([mean_t v0 ; mean_t a0]) and the prior parts of the four bias heads, 16 clips per CTA  (:158-181)
//   v2a_attn    S = Qt·a0^T + tanh(qv_i+ka_j)·s, softmax, U_h = P_h · a0              [16, 4x128] bf16 (row i, head-major columns)
//   video_upd   two GEMMs over weights folded per head at pack time (exact algebra: no nonlinearity in between):
//     video_mix   v1 = LN(v0 + sum_h U_h·(Wo1_h·Wv1_h)^T + (Wo1·bv1 + bo1))  (:394-395) ; pooled_v ; c_hi = scale·K2_h[i]·bq2_h
//     video_kv    Kt_h = v1·(scale·Wq2_h^T·Wk2_h)^T + scale·bk2_h·Wq2_h ,  Vt_h = v1·(Wo2_h·Wv2_h)^T + Wo2_h·bv2_h + bo2/4
//                                                                                                      [64,128] bf16 each
//   a2v_attn    S = a0·Kt^T + c + tanh(qa_j+kv_i)·s, softmax over the 16 keys of each head, out = P·Vt + bo2,
//               a1 = LN(a0 + out) (:399), pooled_a = mean_t a1 (:402)
//   head        concat MLP / gated head on [pooled_v ; pooled_a]  (:405-411)
// The K/V/Q/out projections of the AUDIO tokens never run as token GEMMs: they are folded into the per-clip 64x128
// operands Qt / Kt / Vt (exact algebra), which removes 19.5 of the 55.7 MFLOP/clip.
#pragma once
#include "common.cuh"

namespace xf {

constexpr int BE = 128;    // d_model the bf16 kernels are specialised for
constexpr int BH = 4;      // heads
constexpr int BDH = 32;    // head dim
constexpr int BEP = 136;   // bf16 row pitch (elements) of the 128-wide WEIGHT tiles kept in smem: 272 B rows => conflict-free ldmatrix
constexpr int BOP = 128;   // row pitch of the per-clip operands in HBM (a0, Qt, Kt, Vt): dense 256 B rows, so every 128 B
                           // half-row a TMA box fetches is one aligned line
constexpr int BTP = 16;    // video tokens padded to one m16 tile
constexpr int BTAP = 160;  // audio tokens padded to ten m16 tiles

// ---- packed (derived) weights, byte offsets --------------------------------------------------------------------------
struct PackBf16 {
  int64_t Wv, Wfold, bfold;                          // bf16 [E,Dv], bf16 [E,Da], f32 [E]
  int64_t Wq1, Wk1, Wv1, Wo1, Wk2, Wv2, Wq2, Wo2;    // bf16 [E,BEP] each (torch [out,in] layout, padded pitch)
  int64_t W0b;                                       // bf16 [HW,2E] in mma B-fragment order (bf16_pack.cu: frag_pack_kernel): head layer 0
                                                     // (xattn_mlp.0 / xattn_gate.0) weight
  int64_t pW0;                                       // same order, [ceil8(Hd),2E]: prior_net.0.weight (0 bytes without the prior)
  int64_t scw;                                       // f32 [2,E]: a_key_bias / a_query_bias weights, token part (first E entries)
  // video-side update folded into two GEMMs (gemm_tc.cuh: video_mix, video_kv)
  int64_t Gmix, bmix;                                // bf16 [E,4E]: G[n][128h+e] = sum_d Wo1[n][32h+d]·Wv1[32h+d][e]; f32 [E]: Wo1·bv1 + bo1
  int64_t Fkv, bkv;                                  // bf16 [8E,E]: rows 128h..   Fk_h[e'][k] = scale·sum_d Wq2[32h+d][e']·Wk2[32h+d][k]
                                                     //              rows 512+128h Fv_h[e'][k] = sum_d Wo2[e'][32h+d]·Wv2[32h+d][k]
                                                     // f32 [8E]: scale·bk2_h·Wq2_h | Wo2_h·bv2_h + bo2/4
  int64_t wcv;                                       // f32 [4E + 4]: wc_h[k] = scale·sum_d Wk2[32h+d][k]·bq2[32h+d]; cc_h = scale·bk2_h·bq2_h
  int64_t scratch;                                   // f32 [E,Da] temp of the fold
  int64_t total;
};
void compute_pack_bf16(const xattn_config& c, PackBf16& p);

// ---- workspace, byte offsets -----------------------------------------------------------------------------------------
struct WsBf16 {
  int64_t v0;       // f32  [B,T,E]
  int64_t a0;       // bf16 [B*Ta, BOP]
  int64_t qt;       // bf16 [B,64,BOP]
  int64_t meanv;    // f32  [B,E]
  int64_t tqv, tkv; // f32  [B,16]
  int64_t u;        // bf16 [B*16, 4E]: row (clip, i), columns 128h.. = U_h[i] = softmax(S_h)[i]·a0
  int64_t tok;      // f32  [2,B,160]  w_ak·a0_j | w_aq·a0_j from the a0 GEMM epilogue
  int64_t acol;     // f32  [B,6,E]    column sums of a0 per clip and 32-row output group (gemm_tc.cuh)
  int64_t pc;       // f32  [B,4]
  int64_t kt, vt;   // bf16 [B,64,BOP]  (Vt rows carry + bo2/4: every head's probabilities sum to one)
  int64_t cv;       // f32  [B,64]
  int64_t pooled;   // f32  [B,2E]
  int64_t poolp;    // f32  [B,8,E]    per-warp partial sums of LayerNorm-ed audio rows (a2v_tc -> head)
  int64_t v1;       // bf16 [B*16, E]  updated video tokens (fusion.py:395), 16 slots per clip
  int64_t xhat;     // bf16 [B*Ta, E]  pool_mode = ATTN only: the audio tokens after a_norm without its affine (a2v_tc -> attn_pool)
  int64_t total;
};
void compute_ws_bf16(const xattn_config& c, WsBf16& w);

int bf16_supported(const xattn_config& c);   // XATTN_OK or XATTN_ERR_SHAPE with the reason in last_error

// ---- kernel launchers (each returns an xattn_status) ---------------------------------------------------------------------
struct Bf16Ctx {
  const xattn_config* cfg;
  const float* params;          // flat fp32 parameters
  ParamLayout pl;
  const uint8_t* packed;
  PackBf16 pk;
  uint8_t* ws;
  WsBf16 w;
  cudaStream_t st;
};
int launch_video_pre(const Bf16Ctx& x);
int launch_prior(const Bf16Ctx& x);
int launch_v2a_tc(const Bf16Ctx& x);
int launch_a2v_tc(const Bf16Ctx& x);     // leaves pooled_a as per-warp partial sums in ws.poolp (finalised by the head kernel);
                                         // pool_mode = ATTN: also the normalised audio tokens in ws.xhat
int launch_learned_pool(const Bf16Ctx& x);   // pool_mode = ATTN: TemporalAttentionPooling of v1 and of the audio tokens into ws.pooled
int launch_video_update(const Bf16Ctx& x);   // the video_mix + video_kv GEMMs (gemm_tc.cuh)
int launch_head_pooled(const Bf16Ctx& x, float* logits);

}  // namespace xf
