// Argument blocks + entry declarations of the fp32 kernels (fwd_f32.cu / bwd_f32.cu).
#pragma once
#include "common.cuh"

namespace xf {

struct PriorArgs {
  const float *v0, *a0;                 // [B,T,E], [B,Ta,E] pre-attention tokens
  const float *w0, *b0, *w3, *b3;       // prior_net.0 / .3
  const float* wh[4];                   // score heads: 0 v_query, 1 a_key, 2 a_query, 3 v_key   ([E+P])
  const float* bh[4];
  float *pooled0, *phid, *prior;        // [B,2E], [B,Hd] (post ReLU+dropout), [B,P]
  float *qv, *ka, *qa, *kvs;            // [B,T], [B,Ta], [B,Ta], [B,T]
  int T, Ta, E, P, Hd;
  float p_drop;
  uint64_t seed, offset;
};

struct AttnArgs {
  const float* q;        // [B,L,E]
  const float* kv;       // [B,S,2E]  (K | V)
  const float* qs;       // [B,L] query-side prior score or nullptr
  const float* ks;       // [B,S]
  const float* bias_scale;
  float* P;              // [B,H,L,S] softmax probabilities (pre-dropout) or nullptr
  int ldp = 0;           // row stride of P for the tensor-core kernels (attn_mma.cu; 0 = S); the SIMT kernels always use S
  float* O;              // [B,L,E]
  int L, S, E, H;
  float scale, p_drop;
  uint64_t seed, offset;
  uint32_t rng_stream;
};

struct AttnBwdArgs {
  const float *q, *kv, *qs, *ks, *bias_scale, *P;
  int ldp = 0;           // as AttnArgs::ldp
  const float* dO;       // [B,L,E]
  float* dq;             // [B,L,E]
  float* dkv;            // [B,S,2E]
  float* dqs;            // [B,L]  (nullptr without prior)
  float* dks;            // [B,S]
  float* dbias_scale;    // scalar, atomically accumulated
  float* dbq = nullptr;  // [E]  += column sums of dq  (in_proj bias gradient; attn_bwd32_kernel only, see launch_attn_bwd)
  float* dbkv = nullptr; // [2E] += column sums of dkv
  int L, S, E, H;
  float scale, p_drop;
  uint64_t seed, offset;
  uint32_t rng_stream;
};

struct HeadArgs {
  const float *v1, *a1;                 // [B,T,E], [B,Ta,E] post-LN tokens
  const float *w0, *b0, *w1, *b1, *wc, *bc;
  float *pooled, *hhid, *gate, *fusedg; // saved: [B,2E], [B,HW], [B], [B,E]
  float* logits;
  int T, Ta, E, HW, C, gated;
  int stage = 0;                        // pool_head_fwd_kernel: 0 = whole head, 1 = temporal means only, 2 = from a.pooled / a.hhid
  float p_drop;
  uint64_t seed, offset;
};

size_t attn_smem_bytes(int L, int S, int dh, int nwarps);
int launch_attn_fwd(cudaStream_t st, const AttnArgs& a, int B);
// tensor-core (mma.sync tf32) versions for the shapes of the fusion block (attn_mma.cu); taken when precision = TF32
bool attn_mma_ok(int L, int S, int E, int H);
int launch_attn_fwd_mma(cudaStream_t st, const AttnArgs& a, int B);
struct AttnBwdArgs;
int launch_attn_bwd_mma(cudaStream_t st, const AttnBwdArgs& a, int B);

int launch_attn_pool_bf16(cudaStream_t st, const void* x_bf16, int64_t clip_stride, const float* pre_g, const float* pre_b,
                          const float* const* p, float* out, int64_t ld_out, int B, int L, int E);   // fwd_f32.cu
// row-wise pieces of the transformer pooler (fwd_f32.cu / tx_pool_bwd.cu)
int launch_ln_rows(cudaStream_t st, const float* x, const float* gamma, const float* beta, float* y, int rows, int E);
int launch_ln_rows_bwd(cudaStream_t st, const float* x, const float* gamma, const float* d_h, float* d_x, float* g_gamma,
                       float* g_beta, int rows, int E, bool accumulate);
int launch_gelu_drop_bwd(cudaStream_t st, float* d, const float* pre, int64_t n, float p, uint64_t seed, uint64_t offset, uint32_t stream);
int launch_mul_drop(cudaStream_t st, float* out, const float* x, int64_t n, float p, uint64_t seed, uint64_t offset, uint32_t stream);
int launch_add_inplace(cudaStream_t st, float* x, const float* y, int64_t n);
int launch_attn_bwd_rows(cudaStream_t st, const AttnBwdArgs& a, int B);   // no prior bias; any L, S that fit a [32, S] tile

int forward_f32(const xattn_config& c, const float* params, const void* xv, const void* xa, float* logits, void* ws,
                cudaStream_t st);
struct BackwardOptions {
  cudaStream_t tail_stream = nullptr;   // see backward_f32_impl
  float* d_video = nullptr;             // [B,T,Dv] fp32 or nullptr
  float* d_audio = nullptr;             // [B,Ta,Da] fp32 or nullptr
  bool fork = true;                     // the library may use its per-device side stream (SideFork)
  int phase = 0;                        // 0 whole backward; 1 up to the tail point; 2 the input projections only (backward_f32_impl)
};
int backward_f32(const xattn_config& c, const float* params, const void* xv, const void* xa, const float* dlogits,
                 float* grads, void* ws, cudaStream_t st, const BackwardOptions& opt = BackwardOptions());

}  // namespace xf
