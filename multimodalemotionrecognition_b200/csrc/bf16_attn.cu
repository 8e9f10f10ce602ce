// The two fused per-clip attention kernels of the bf16 path.  One CTA owns one clip; the clip's audio tokens a0
// [Ta,128] (bf16) are pulled into shared memory once by a bulk async copy (cp.async.bulk + mbarrier) and serve as BOTH
// the key and the value operand, because the K/V projections are folded into the small per-clip operands Qt / Kt / Vt.
// The whole score tile lives in registers: bias add (tanh of the emotion-prior token scores, fusion.py:175-176),
// warp-shuffle softmax, and the P·V product run in one pass with bf16 mma.sync (fp32 accumulate, fp32 softmax/LayerNorm).
//   v2a_attn_kernel : video queries audio (fusion.py:394)
//   a2v_attn_kernel : audio queries the updated video (:398), residual + LayerNorm (:399), temporal mean (:402)
#include "bf16_layout.cuh"
#include "frag.cuh"

namespace xf {

namespace {

constexpr int kA0Bytes = BTAP * BEP * 2;   // 43,520
constexpr int kOpBytes = 64 * BEP * 2;     // 17,408  (Qt / Kt / Vt)

struct V2AArgs {
  const __nv_bfloat16* a0;   // [B*Ta, BEP]
  const __nv_bfloat16* qt;   // [B,64,BEP]
  const float *tqv, *tok, *pc;   // [B,16] w_vq·v0_i ; [2,B,160] w_ak·a0_j | w_aq·a0_j ; [B,4] prior parts of the bias heads
  const float* bias_scale;
  __nv_bfloat16* u;          // [B,64,E]
  int B, Ta, T, use_prior;
};

__global__ void __launch_bounds__(128, 3) v2a_attn_kernel(V2AArgs a) {
  extern __shared__ __align__(128) uint8_t sm[];
  __nv_bfloat16* a0s = reinterpret_cast<__nv_bfloat16*>(sm);
  __nv_bfloat16* qts = reinterpret_cast<__nv_bfloat16*>(sm + kA0Bytes);
  float* fsm = reinterpret_cast<float*>(sm + kA0Bytes + kOpBytes);
  float* ka_s = fsm;             // [160] a_key token scores + prior part
  float* qv_s = ka_s + BTAP;     // [16]  v_query token scores + prior part
  uint64_t* bar = reinterpret_cast<uint64_t*>(qv_s + 16);
  const int c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int Ta = a.Ta;
  if (tid == 0) {
    tc::mbar_init(bar, 1);
    tc::fence_barrier_init();
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t nb = (uint32_t)Ta * BEP * 2;
    tc::mbar_expect_tx(bar, nb + kOpBytes);
    tc::bulk_load_1d(a0s, a.a0 + (int64_t)c * Ta * BEP, nb, bar);
    tc::bulk_load_1d(qts, a.qt + (int64_t)c * 64 * BEP, kOpBytes, bar);
  }
  {   // zero the padding token rows while the copy is in flight
    uint4* z = reinterpret_cast<uint4*>(a0s + Ta * BEP);
    const int nz = (BTAP - Ta) * BEP * 2 / 16;
    for (int i = tid; i < nz; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
    if (a.use_prior) {
      const float pq = a.pc[(int64_t)c * 4 + 0], pk = a.pc[(int64_t)c * 4 + 1];
      for (int i = tid; i < Ta; i += blockDim.x) ka_s[i] = a.tok[(int64_t)c * BTAP + i] + pk;
      if (tid < 16) qv_s[tid] = a.tqv[(int64_t)c * 16 + tid] + pq;
    }
  }
  __syncthreads();
  tc::mbar_wait(bar, 0);
  const uint32_t a0b = tc::smem_u32(a0s), qtb = tc::smem_u32(qts);

  // ---- scores of head `warp`: S[i, j] = Qt_h[i,:]·a0[j,:]  (+ bias), i < 16, j < 160 ----
  const int h = warp;
  uint32_t pa[BTAP / 16][4];   // softmax numerators as A fragments (bf16), filled below
  float inv0, inv1;
  {
    float s[BTAP / 8][4];
    uint32_t aq[8][4];
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) fr::ldA(qtb, 16 * h, ks, lane, aq[ks]);
#pragma unroll
    for (int nt = 0; nt < BTAP / 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 8; ks += 2) {
        uint32_t b[4];
        fr::ldB_nk(a0b, nt, ks, lane, b);
        tc::mma_bf16(s[nt], aq[ks], b[0], b[1]);
        tc::mma_bf16(s[nt], aq[ks + 1], b[2], b[3]);
      }
    }
    float m0 = -INFINITY, m1 = -INFINITY;
    const float bsc = a.use_prior ? a.bias_scale[0] : 0.f;
    const float q0 = a.use_prior ? qv_s[g] : 0.f, q1 = a.use_prior ? qv_s[g + 8] : 0.f;
#pragma unroll
    for (int nt = 0; nt < BTAP / 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int j = nt * 8 + 2 * t + e;
        if (j < Ta) {
          if (a.use_prior) {
            const float kj = ka_s[j];
            s[nt][e] += tc::tanh_fast(q0 + kj) * bsc;
            s[nt][2 + e] += tc::tanh_fast(q1 + kj) * bsc;
          }
        } else {
          s[nt][e] = -INFINITY;
          s[nt][2 + e] = -INFINITY;
        }
        m0 = fmaxf(m0, s[nt][e]);
        m1 = fmaxf(m1, s[nt][2 + e]);
      }
    }
    m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
    m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < BTAP / 8; ++nt) {
      s[nt][0] = __expf(s[nt][0] - m0); s[nt][1] = __expf(s[nt][1] - m0);
      s[nt][2] = __expf(s[nt][2] - m1); s[nt][3] = __expf(s[nt][3] - m1);
      l0 += s[nt][0] + s[nt][1];
      l1 += s[nt][2] + s[nt][3];
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    inv0 = 1.f / l0;
    inv1 = 1.f / l1;
#pragma unroll
    for (int kk = 0; kk < BTAP / 16; ++kk) fr::c2a(s[2 * kk], s[2 * kk + 1], pa[kk]);
  }
  // ---- U_h = softmax(S_h) · a0   [16,128] ----
  float o[16][4];
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
  for (int kk = 0; kk < BTAP / 16; ++kk) {
#pragma unroll
    for (int nt2 = 0; nt2 < 8; ++nt2) {
      uint32_t b[4];
      fr::ldB_kn(a0b, 16 * kk, nt2, lane, b);
      tc::mma_bf16(o[2 * nt2], pa[kk], b[0], b[1]);
      tc::mma_bf16(o[2 * nt2 + 1], pa[kk], b[2], b[3]);
    }
  }
  // U leaves in A-FRAGMENT order: lane (g,t) of head h writes the 32 registers it would feed to mma.sync as the A operand
  // of the next product (video_update reads them back with eight 16-byte loads): [clip][head][ks][lane][4] x bf16x2 (a warp-wide 16-byte access is 512 contiguous bytes).
  uint4* up = reinterpret_cast<uint4*>(a.u) + ((int64_t)c * BH + h) * 256 + lane;
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) {
    uint4 v;
    v.x = tc::pack_bf16(o[2 * ks][0] * inv0, o[2 * ks][1] * inv0);
    v.y = tc::pack_bf16(o[2 * ks][2] * inv1, o[2 * ks][3] * inv1);
    v.z = tc::pack_bf16(o[2 * ks + 1][0] * inv0, o[2 * ks + 1][1] * inv0);
    v.w = tc::pack_bf16(o[2 * ks + 1][2] * inv1, o[2 * ks + 1][3] * inv1);
    up[ks * 32] = v;
  }
}

struct A2VArgs {
  const __nv_bfloat16 *a0, *kt, *vt;   // [B*Ta,BEP], [B,64,BEP] x2
  const float *tok, *tkv, *pc, *cv;    // [2,B,160], [B,16], [B,4], [B,64]
  const float *bo2, *gam, *bet, *bias_scale;
  float* pooled;                       // [B,2E] (second E = pooled_a)
  int B, Ta, T, use_prior;
};

__global__ void __launch_bounds__(160, 2) a2v_attn_kernel(A2VArgs a) {
  extern __shared__ __align__(128) uint8_t sm[];
  __nv_bfloat16* a0s = reinterpret_cast<__nv_bfloat16*>(sm);
  __nv_bfloat16* kts = reinterpret_cast<__nv_bfloat16*>(sm + kA0Bytes);
  __nv_bfloat16* vts = reinterpret_cast<__nv_bfloat16*>(sm + kA0Bytes + kOpBytes);
  float* fsm = reinterpret_cast<float*>(sm + kA0Bytes + 2 * kOpBytes);
  float* qa_s = fsm;            // [160] a_query token scores + prior part
  float* kv_s = qa_s + BTAP;    // [16]  v_key token scores + prior part
  float* cv_s = kv_s + 16;      // [64]  scale * K2_h[i]·bq2_h
  float* bo_s = cv_s + 64;      // [128]
  float* ga_s = bo_s + BE;      // [128]
  float* be_s = ga_s + BE;      // [128]
  float* pool_s = be_s + BE;    // [5][128] per-warp partial sums (summed in a fixed order: deterministic)
  uint64_t* bar = reinterpret_cast<uint64_t*>(pool_s + 5 * BE);
  const int c = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int Ta = a.Ta, T = a.T;
  if (tid == 0) {
    tc::mbar_init(bar, 1);
    tc::fence_barrier_init();
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t nb = (uint32_t)Ta * BEP * 2;
    tc::mbar_expect_tx(bar, nb + 2 * kOpBytes);
    tc::bulk_load_1d(a0s, a.a0 + (int64_t)c * Ta * BEP, nb, bar);
    tc::bulk_load_1d(kts, a.kt + (int64_t)c * 64 * BEP, kOpBytes, bar);
    tc::bulk_load_1d(vts, a.vt + (int64_t)c * 64 * BEP, kOpBytes, bar);
  }
  {
    uint4* z = reinterpret_cast<uint4*>(a0s + Ta * BEP);
    const int nz = (BTAP - Ta) * BEP * 2 / 16;
    for (int i = tid; i < nz; i += blockDim.x) z[i] = make_uint4(0, 0, 0, 0);
    const float pq = a.use_prior ? a.pc[(int64_t)c * 4 + 2] : 0.f, pk = a.use_prior ? a.pc[(int64_t)c * 4 + 3] : 0.f;
    for (int i = tid; i < BTAP; i += blockDim.x) qa_s[i] = (a.use_prior && i < Ta) ? a.tok[((int64_t)a.B + c) * BTAP + i] + pq : 0.f;
    if (tid < 16) kv_s[tid] = a.use_prior ? a.tkv[(int64_t)c * 16 + tid] + pk : 0.f;
    if (tid < 64) cv_s[tid] = a.cv[(int64_t)c * 64 + tid];
    for (int i = tid; i < BE; i += blockDim.x) {
      bo_s[i] = a.bo2[i]; ga_s[i] = a.gam[i]; be_s[i] = a.bet[i];
    }
  }
  __syncthreads();
  tc::mbar_wait(bar, 0);
  const uint32_t a0b = tc::smem_u32(a0s), ktb = tc::smem_u32(kts), vtb = tc::smem_u32(vts);
  const float bsc = a.use_prior ? a.bias_scale[0] : 0.f;
  float pool[16][2];
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) pool[nt][0] = pool[nt][1] = 0.f;
  float poolc = 0.f;   // sum over this thread's rows of mu_j * r_j (same for every column)

#pragma unroll 1
  for (int mt = 2 * warp; mt < 2 * warp + 2; ++mt) {
    const int j0 = 16 * mt + g, j1 = j0 + 8;
    // ---- S[j, (h,i)] = a0[j,:]·Kt[(h,i),:] + c_hi + bias ----
    uint32_t pf[4][4];
    {
      float s[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 8; ks += 2) {
        uint32_t x0[4], x1[4];
        fr::ldA(a0b, 16 * mt, ks, lane, x0);
        fr::ldA(a0b, 16 * mt, ks + 1, lane, x1);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          uint32_t b[4];
          fr::ldB_nk(ktb, nt, ks, lane, b);
          tc::mma_bf16(s[nt], x0, b[0], b[1]);
          tc::mma_bf16(s[nt], x1, b[2], b[3]);
        }
      }
      // the prior bias tanh(qa_j + kv_i)·scale is the same for every head (fusion.py:351-354): once per (row, key)
      float bia0[4], bia1[4];
      {
        const float qa0 = qa_s[j0], qa1 = qa_s[j1];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = (q >> 1) * 8 + 2 * t + (q & 1);
          const float kvv = kv_s[i];
          bia0[q] = a.use_prior ? tc::tanh_fast(qa0 + kvv) * bsc : 0.f;
          bia1[q] = a.use_prior ? tc::tanh_fast(qa1 + kvv) * bsc : 0.f;
          if (i >= T) bia0[q] = bia1[q] = -INFINITY;   // padded video tokens never receive probability mass
        }
      }
      constexpr float kL2E = 1.4426950408889634f;
#pragma unroll
      for (int hh = 0; hh < BH; ++hh) {
        float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int nt = 2 * hh + (q >> 1), e = q & 1;
          const float cc = cv_s[hh * 16 + (q >> 1) * 8 + 2 * t + e];
          s[nt][e] += cc + bia0[q];
          s[nt][2 + e] += cc + bia1[q];
          m0 = fmaxf(m0, s[nt][e]);
          m1 = fmaxf(m1, s[nt][2 + e]);
        }
        m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)); m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
        m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)); m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
        const float n0 = -m0 * kL2E, n1 = -m1 * kL2E;
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int n2 = 0; n2 < 2; ++n2) {
          const int nt = 2 * hh + n2;
          s[nt][0] = exp2f(fmaf(s[nt][0], kL2E, n0)); s[nt][1] = exp2f(fmaf(s[nt][1], kL2E, n0));
          s[nt][2] = exp2f(fmaf(s[nt][2], kL2E, n1)); s[nt][3] = exp2f(fmaf(s[nt][3], kL2E, n1));
          l0 += s[nt][0] + s[nt][1];
          l1 += s[nt][2] + s[nt][3];
        }
        l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const float i0 = __frcp_rn(l0), i1 = __frcp_rn(l1);
#pragma unroll
        for (int n2 = 0; n2 < 2; ++n2) {
          const int nt = 2 * hh + n2;
          s[nt][0] *= i0; s[nt][1] *= i0; s[nt][2] *= i1; s[nt][3] *= i1;
        }
        fr::c2a(s[2 * hh], s[2 * hh + 1], pf[hh]);
      }
    }
    // ---- out = P · Vt  (out_proj already folded into Vt) ----
    float o[16][4];
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) o[nt][0] = o[nt][1] = o[nt][2] = o[nt][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int nt2 = 0; nt2 < 8; ++nt2) {
        uint32_t b[4];
        fr::ldB_kn(vtb, 16 * kk, nt2, lane, b);
        tc::mma_bf16(o[2 * nt2], pf[kk], b[0], b[1]);
        tc::mma_bf16(o[2 * nt2 + 1], pf[kk], b[2], b[3]);
      }
    }
    // ---- a1 = LayerNorm(a0 + out + bo2) (fusion.py:399), accumulated into the temporal mean (:402) ----
    // mean_j xhat_j = sum_j r_j s_j - sum_j r_j mu_j : the rows are never normalised element by element.
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      const int col = nt * 8 + 2 * t;
      const __nv_bfloat162 r0 = *reinterpret_cast<const __nv_bfloat162*>(a0s + j0 * BEP + col);
      const __nv_bfloat162 r1 = *reinterpret_cast<const __nv_bfloat162*>(a0s + j1 * BEP + col);
      const float2 bo = *reinterpret_cast<const float2*>(bo_s + col);
      o[nt][0] += bo.x + __low2float(r0);
      o[nt][1] += bo.y + __high2float(r0);
      o[nt][2] += bo.x + __low2float(r1);
      o[nt][3] += bo.y + __high2float(r1);
      s0 += o[nt][0] + o[nt][1];
      s1 += o[nt][2] + o[nt][3];
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
    const float mu0 = s0 * (1.f / BE), mu1 = s1 * (1.f / BE);
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {   // two-pass variance (as torch LayerNorm): sum (x - mu)^2
      float d;
      d = o[nt][0] - mu0; q0 = fmaf(d, d, q0);
      d = o[nt][1] - mu0; q0 = fmaf(d, d, q0);
      d = o[nt][2] - mu1; q1 = fmaf(d, d, q1);
      d = o[nt][3] - mu1; q1 = fmaf(d, d, q1);
    }
    q0 += __shfl_xor_sync(0xffffffffu, q0, 1); q0 += __shfl_xor_sync(0xffffffffu, q0, 2);
    q1 += __shfl_xor_sync(0xffffffffu, q1, 1); q1 += __shfl_xor_sync(0xffffffffu, q1, 2);
    const float r0 = (j0 < Ta) ? rsqrtf(q0 * (1.f / BE) + 1e-5f) : 0.f;   // padded token rows contribute nothing
    const float r1 = (j1 < Ta) ? rsqrtf(q1 * (1.f / BE) + 1e-5f) : 0.f;
    poolc += mu0 * r0 + mu1 * r1;
#pragma unroll
    for (int nt = 0; nt < 16; ++nt) {
      pool[nt][0] = fmaf(o[nt][0], r0, fmaf(o[nt][2], r1, pool[nt][0]));
      pool[nt][1] = fmaf(o[nt][1], r0, fmaf(o[nt][3], r1, pool[nt][1]));
    }
  }
  // pooled_a[e] = gamma[e] * mean_j xhat[j,e] + beta[e]
#pragma unroll
  for (int o2 = 4; o2 <= 16; o2 <<= 1) poolc += __shfl_xor_sync(0xffffffffu, poolc, o2);
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) {
    float p0 = pool[nt][0], p1 = pool[nt][1];
#pragma unroll
    for (int o2 = 4; o2 <= 16; o2 <<= 1) {
      p0 += __shfl_xor_sync(0xffffffffu, p0, o2);
      p1 += __shfl_xor_sync(0xffffffffu, p1, o2);
    }
    if (g == 0) {
      pool_s[warp * BE + nt * 8 + 2 * t] = p0 - poolc;
      pool_s[warp * BE + nt * 8 + 2 * t + 1] = p1 - poolc;
    }
  }
  __syncthreads();
  if (tid < BE) {
    const float tot = (((pool_s[tid] + pool_s[BE + tid]) + pool_s[2 * BE + tid]) + pool_s[3 * BE + tid]) + pool_s[4 * BE + tid];
    a.pooled[(int64_t)c * 2 * BE + BE + tid] = ga_s[tid] * (tot / (float)Ta) + be_s[tid];
  }
}

}  // namespace

int launch_v2a_attn(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  V2AArgs a;
  a.a0 = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.a0);
  a.qt = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.qt);
  a.tqv = reinterpret_cast<const float*>(x.ws + x.w.tqv);
  a.tok = reinterpret_cast<const float*>(x.ws + x.w.tok);
  a.pc = reinterpret_cast<const float*>(x.ws + x.w.pc);
  a.bias_scale = P(XP_BIAS_SCALE);
  a.u = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.u);
  a.B = c.B; a.Ta = c.Ta; a.T = c.T; a.use_prior = c.use_prior;
  const int smem = kA0Bytes + kOpBytes + (BTAP + 16) * 4 + 16;
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(v2a_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  v2a_attn_kernel<<<c.B, 128, smem, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int launch_a2v_attn(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  A2VArgs a;
  a.a0 = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.a0);
  a.kt = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.kt);
  a.vt = reinterpret_cast<const __nv_bfloat16*>(x.ws + x.w.vt);
  a.tok = reinterpret_cast<const float*>(x.ws + x.w.tok);
  a.tkv = reinterpret_cast<const float*>(x.ws + x.w.tkv);
  a.pc = reinterpret_cast<const float*>(x.ws + x.w.pc);
  a.cv = reinterpret_cast<const float*>(x.ws + x.w.cv);
  a.bo2 = P(XP_A2V_OUT_B); a.gam = P(XP_AN_W); a.bet = P(XP_AN_B);
  a.bias_scale = P(XP_BIAS_SCALE);
  a.pooled = reinterpret_cast<float*>(x.ws + x.w.pooled);
  a.B = c.B; a.Ta = c.Ta; a.T = c.T; a.use_prior = c.use_prior;
  const int smem = kA0Bytes + 2 * kOpBytes + (BTAP + 16 + 64 + 8 * BE) * 4 + 16;
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(a2v_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  a2v_attn_kernel<<<c.B, 160, smem, x.st>>>(a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
