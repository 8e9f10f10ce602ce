// bf16 tensor-core forward of the xattn fusion head (precision == XATTN_DTYPE_BF16): launch sequence + classifier head.
// Stage map against reference src/models/fusion.py is in bf16_layout.cuh.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "bf16_layout.cuh"
#include "gemm_tc.cuh"
#include "kernels_bf16.cuh"
#include "kernels_f32.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kHeadClips = 32;   // clips per CTA of the head kernel (two m16 tiles): 128 CTAs at B = 4096 — every CTA reads the whole
                                 // layer-0 weight from L2, and with more, smaller CTAs those identical reads queue up on the same lines

struct HeadPooledArgs {
  float* pooled;         // [B,2E] = [pooled_v ; pooled_a]
  const float* poolp;    // [B,8,E] partial sums of the normalised audio rows (tcgen05 a2v) or nullptr (pooled_a already final)
  const float *an_w, *an_b;   // a_norm affine, applied here when poolp is set: mean_t LN(x) = gamma * mean xhat + beta
  int Ta;
  const __nv_bfloat16* w0b;   // [HW,2E] layer 0 weight, bf16 in mma B-fragment order (frag_pack_kernel)
  const float *b0, *w1, *b1, *wc, *bc;
  float* logits;
  int B, HW, C, gated;
};

// concat head: Linear(2E->HW), ReLU, (Dropout eval), Linear(HW->C)                       (fusion.py:405-407)
// gated head : g = sigmoid(Linear(E->1)(ReLU(Linear(2E->E)))), fused = g*v + (1-g)*a, Linear(E->C)   (:408-411)
// Layer 0 runs on bf16 mma.sync with M = the clips of the CTA (m16 tiles).  The fp32 activations are split once into a hi + lo bf16
// pair (two MMAs per step) so only the weights are rounded, as everywhere else on this path.  Row pitches are chosen so
// that the eight rows a fragment load touches fall in different banks.
constexpr int kXP = 2 * BE + 8;    // bf16 pitch of the split activations (528 B: row g -> bank 4g)
constexpr int kHP = 256 + 4;       // fp32 pitch of hid / final-layer weights
constexpr int kHeadThreads = 512;   // 16 warps: the n-tiles of layer 0 spread thin, the kernel is one latency chain per CTA
constexpr int kHeadRegion = 2 * kHeadClips * kXP * 2 > 16 * kHP * 4 ? 2 * kHeadClips * kXP * 2 : 16 * kHP * 4;
constexpr int kHeadSmem = kHeadClips * 8 * BE * 4 + kHeadRegion + kHeadClips * kHP * 4;
__global__ void __launch_bounds__(kHeadThreads, 1) head_pooled_kernel(HeadPooledArgs a) {
  // dynamic smem: [pstage: the clip block's partial sums][region][hid].  xhi/xlo (layer-0 operands) and wl (final linear, [C][K])
  // are live in different phases and share `region`.
  extern __shared__ __align__(16) unsigned char hsm[];
  float* pstage = reinterpret_cast<float*>(hsm);   // [kHeadClips][8][BE] (only with poolp)
  unsigned char* region = hsm + kHeadClips * 8 * BE * 4;
  auto hid = reinterpret_cast<float(*)[kHP]>(region + kHeadRegion);
  auto xhi = reinterpret_cast<__nv_bfloat16(*)[kXP]>(region);
  auto xlo = reinterpret_cast<__nv_bfloat16(*)[kXP]>(region + kHeadClips * kXP * 2);
  auto wl = reinterpret_cast<float(*)[kHP]>(region);
  __shared__ float gate[kHeadClips];
  const int tid = threadIdx.x, c0 = blockIdx.x * kHeadClips;
  const int nc = min(kHeadClips, a.B - c0);
  const float* wfin = a.gated ? a.wc : a.w1;
  const float* bfin = a.gated ? a.bc : a.b1;
  const int K = a.gated ? BE : a.HW;
  // The clip block's partial sums (16 clips x 8 slots x 128 floats, contiguous) arrive as ONE bulk async copy: one memory round
  // trip instead of one per loop iteration of dependent loads.
  __shared__ uint64_t pbar;
  tc::pdl_wait();
  tc::pdl_launch();
  if (a.poolp != nullptr && tid == 0) {
    tc::mbar_init(&pbar, 1);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(&pbar, (uint32_t)(nc * 8 * BE * 4));
    tc::bulk_load_1d(pstage, a.poolp + (int64_t)c0 * 8 * BE, (uint32_t)(nc * 8 * BE * 4), &pbar);
  }
  // The weight fragments of a layer-0 tile (fragment order: 8 x 512 contiguous bytes per warp) are fetched one tile ahead of the
  // MMAs that use them; the first tile's right here, so their L2 round trip hides behind the activation loads.
  const int warp = tid >> 5, lane = tid & 31;
  const int ntot = a.HW / 8;   // n-tiles of layer 0, dealt round-robin to the 16 warps
  auto fetch = [&](uint4 (&bw)[2 * BE / 32], int nt) {
    const uint4* wf = reinterpret_cast<const uint4*>(a.w0b) + ((int64_t)nt * (2 * BE / 32)) * 32 + lane;
#pragma unroll
    for (int kp = 0; kp < 2 * BE / 32; ++kp) bw[kp] = __ldg(wf + kp * 32);
  };
  uint4 b[2 * BE / 32];
  if (warp < ntot) fetch(b, warp);
  float2 xv[kHeadClips * BE / 2 / kHeadThreads];   // pooled_v, in flight meanwhile
#pragma unroll
  for (int r = 0; r < kHeadClips * BE / 2 / kHeadThreads; ++r) {
    const int i = tid + r * kHeadThreads, cc = i / (BE / 2), k = (i % (BE / 2)) * 2;
    xv[r] = cc < nc ? *reinterpret_cast<const float2*>(a.pooled + (int64_t)(c0 + cc) * 2 * BE + k) : make_float2(0.f, 0.f);
  }
  float4 wpre[2];   // final-layer weights, fetched now and parked in smem once layer 0 is done with the region (C*K <= 4096)
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int i = tid + r * kHeadThreads;
    wpre[r] = i < a.C * K / 4 ? __ldg(reinterpret_cast<const float4*>(wfin) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  auto put = [&](int cc, int k, float2 x) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(x.x, x.y);
    *reinterpret_cast<__nv_bfloat162*>(&xhi[cc][k]) = h;
    *reinterpret_cast<__nv_bfloat162*>(&xlo[cc][k]) = __floats2bfloat162_rn(x.x - __low2float(h), x.y - __high2float(h));
  };
#pragma unroll
  for (int r = 0; r < kHeadClips * BE / 2 / kHeadThreads; ++r) {
    const int i = tid + r * kHeadThreads;
    put(i / (BE / 2), (i % (BE / 2)) * 2, xv[r]);
  }
  if (a.poolp != nullptr) {
    __syncthreads();   // the barrier is initialised
    tc::mbar_wait(&pbar, 0);
  }
  for (int i = tid; i < kHeadClips * BE / 2; i += kHeadThreads) {
    const int cc = i / (BE / 2), k = (i % (BE / 2)) * 2;
    float2 x = make_float2(0.f, 0.f);
    if (cc < nc) {
      const int64_t c = c0 + cc;
      if (a.poolp != nullptr) {   // add the clip's partials in slot order (even clips own 5 slots, odd clips 7)
        const int ns = (c & 1) ? 7 : 5;
        const float* pp = pstage + cc * 8 * BE + k;
#pragma unroll
        for (int s2 = 0; s2 < 7; ++s2) {   // fixed order
          if (s2 < ns) {
            const float2 v = *reinterpret_cast<const float2*>(pp + s2 * BE);
            x.x += v.x;
            x.y += v.y;
          }
        }
        const float inv = 1.f / (float)a.Ta;
        x.x = a.an_w[k] * (x.x * inv) + a.an_b[k];
        x.y = a.an_w[k + 1] * (x.y * inv) + a.an_b[k + 1];
        *reinterpret_cast<float2*>(a.pooled + c * 2 * BE + BE + k) = x;
      } else {
        x = *reinterpret_cast<const float2*>(a.pooled + c * 2 * BE + BE + k);
      }
    }
    put(cc, BE + k, x);
  }
  __syncthreads();
  {
    const int g = lane >> 2, t = lane & 3;
    uint4 bn[2 * BE / 32];
    const uint32_t ahi = tc::smem_u32(&xhi[(lane & 7) + 8 * ((lane >> 3) & 1)][8 * (lane >> 4)]);
    const uint32_t alo = tc::smem_u32(&xlo[(lane & 7) + 8 * ((lane >> 3) & 1)][8 * (lane >> 4)]);
    for (int nt = warp; nt < ntot; nt += kHeadThreads / 32) {
      if (nt + kHeadThreads / 32 < ntot) fetch(bn, nt + kHeadThreads / 32);
      float d[kHeadClips / 16][4], e[kHeadClips / 16][4];
#pragma unroll
      for (int mt = 0; mt < kHeadClips / 16; ++mt)
#pragma unroll
        for (int r = 0; r < 4; ++r) d[mt][r] = e[mt][r] = 0.f;
#pragma unroll
      for (int ks = 0; ks < 2 * BE / 16; ++ks) {
        const uint32_t b0 = (ks & 1) ? b[ks >> 1].z : b[ks >> 1].x, b1 = (ks & 1) ? b[ks >> 1].w : b[ks >> 1].y;
#pragma unroll
        for (int mt = 0; mt < kHeadClips / 16; ++mt) {
          uint32_t hi[4], lo[4];
          tc::ldsm_x4(ahi + mt * 16 * kXP * 2 + ks * 32, hi[0], hi[1], hi[2], hi[3]);
          tc::ldsm_x4(alo + mt * 16 * kXP * 2 + ks * 32, lo[0], lo[1], lo[2], lo[3]);
          tc::mma_bf16(d[mt], hi, b0, b1);
          tc::mma_bf16(e[mt], lo, b0, b1);
        }
      }
      const int col = nt * 8 + 2 * t;
      const float c0b = a.b0[col], c1b = a.b0[col + 1];
#pragma unroll
      for (int mt = 0; mt < kHeadClips / 16; ++mt) {
        hid[16 * mt + g][col] = fmaxf(d[mt][0] + e[mt][0] + c0b, 0.f);
        hid[16 * mt + g][col + 1] = fmaxf(d[mt][1] + e[mt][1] + c1b, 0.f);
        hid[16 * mt + g + 8][col] = fmaxf(d[mt][2] + e[mt][2] + c0b, 0.f);
        hid[16 * mt + g + 8][col + 1] = fmaxf(d[mt][3] + e[mt][3] + c1b, 0.f);
      }
      if (nt + kHeadThreads / 32 < ntot) {
#pragma unroll
        for (int kp = 0; kp < 2 * BE / 32; ++kp) b[kp] = bn[kp];
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int i = tid + r * kHeadThreads;
    if (i < a.C * K / 4) *reinterpret_cast<float4*>(&wl[(i * 4) / K][(i * 4) % K]) = wpre[r];
  }
  if (a.gated) {
    if (tid < kHeadClips) {
      float s = a.b1[0];
      for (int k = 0; k < a.HW; ++k) s = fmaf(a.w1[k], hid[tid][k], s);
      gate[tid] = 1.f / (1.f + expf(-s));
    }
    __syncthreads();
    for (int i = tid; i < kHeadClips * BE; i += blockDim.x) {   // fused overwrites hid[:, :E]
      const int cc = i / BE, e = i % BE;
      float pv = 0.f, pa = 0.f;
      if (cc < nc) {
        pv = a.pooled[(int64_t)(c0 + cc) * 2 * BE + e];
        pa = a.pooled[(int64_t)(c0 + cc) * 2 * BE + BE + e];
      }
      hid[cc][e] = gate[cc] * pv + (1.f - gate[cc]) * pa;
    }
  }
  __syncthreads();
  for (int o = tid; o < nc * a.C; o += blockDim.x) {
    const int cc = o / a.C, cls = o % a.C;
    float s0 = bfin[cls], s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int k = 0; k < K; k += 4) {
      const float4 wv = *reinterpret_cast<const float4*>(&wl[cls][k]);
      const float4 hv = *reinterpret_cast<const float4*>(&hid[cc][k]);
      s0 = fmaf(wv.x, hv.x, s0); s1 = fmaf(wv.y, hv.y, s1); s2 = fmaf(wv.z, hv.z, s2); s3 = fmaf(wv.w, hv.w, s3);
    }
    a.logits[(int64_t)(c0 + cc) * a.C + cls] = (s0 + s1) + (s2 + s3);
  }
}

}  // namespace

int launch_head_pooled(const Bf16Ctx& x, float* logits) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  HeadPooledArgs a;
  a.pooled = reinterpret_cast<float*>(x.ws + x.w.pooled);
  // learned pooling: pooled_a is final (launch_learned_pool); mean pooling: the a2v kernel's partials are summed here
  a.poolp = c.pool_mode == XATTN_POOL_ATTN ? nullptr : reinterpret_cast<const float*>(x.ws + x.w.poolp);
  a.an_w = P(XP_AN_W); a.an_b = P(XP_AN_B);
  a.Ta = c.Ta;
  a.w0b = reinterpret_cast<const __nv_bfloat16*>(x.packed + x.pk.W0b);
  a.b0 = P(XP_HEAD0_B); a.w1 = P(XP_HEAD1_W); a.b1 = P(XP_HEAD1_B); a.wc = P(XP_CLS_W); a.bc = P(XP_CLS_B);
  a.logits = logits;
  a.B = c.B; a.C = c.C; a.gated = c.head_kind == XATTN_HEAD_GATED;
  a.HW = a.gated ? c.E : c.common_dim;
  XF_TRY(ensure_dyn_smem(head_pooled_kernel, kHeadSmem));
  XF_CUDA(launch_chain(head_pooled_kernel, dim3((c.B + kHeadClips - 1) / kHeadClips), dim3(kHeadThreads), kHeadSmem, x.st, a));
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

// TemporalAttentionPooling (temporal.py:9-26) on the tensor-core path: the SIMT pooling kernel of the fp32 family over the bf16
// token rows this path leaves — v1 (a_norm's twin v_norm applied by video_mix) and the audio tokens as a2v_tc normalised them
// (a_norm's affine is applied on load) — into the two halves of the pooled buffer the head reads.
int launch_learned_pool(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  const float* pv[6] = {P(XP_VPOOL_LN_W), P(XP_VPOOL_LN_B), P(XP_VPOOL_W1), P(XP_VPOOL_B1), P(XP_VPOOL_W2), P(XP_VPOOL_B2)};
  const float* pa[6] = {P(XP_APOOL_LN_W), P(XP_APOOL_LN_B), P(XP_APOOL_W1), P(XP_APOOL_B1), P(XP_APOOL_W2), P(XP_APOOL_B2)};
  float* pooled = reinterpret_cast<float*>(x.ws + x.w.pooled);
  XF_TRY(launch_attn_pool_bf16(x.st, x.ws + x.w.v1, 16 * BE, nullptr, nullptr, pv, pooled, 2 * BE, c.B, c.T, BE));
  return launch_attn_pool_bf16(x.st, x.ws + x.w.xhat, (int64_t)c.Ta * BE, P(XP_AN_W), P(XP_AN_B), pa, pooled + BE, 2 * BE, c.B, c.Ta, BE);
}

// Per-device resources of the forked video branch: one private non-blocking stream and the fork / join events, created once
// (under a lock: two threads may call into the same device for the first time together) and reused by every forward.
// Re-recording an event that an earlier forward's cudaStreamWaitEvent already consumed is safe: a wait captures the event's
// state at the time it is enqueued.
struct ForkRes {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  std::mutex enqueue;   // held while one forward enqueues its fork .. join: the events are shared by every caller of the device
};
static int fork_resources(ForkRes** out) {
  static ForkRes res[64];
  static std::mutex mu;
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) XF_FAIL(XATTN_ERR_CUDA, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  ForkRes& r = res[dev];
  if (r.side == nullptr) XF_CUDA(cudaStreamCreateWithFlags(&r.side, cudaStreamNonBlocking));
  if (r.fork == nullptr) XF_CUDA(cudaEventCreateWithFlags(&r.fork, cudaEventDisableTiming));
  if (r.join == nullptr) XF_CUDA(cudaEventCreateWithFlags(&r.join, cudaEventDisableTiming));
  *out = &r;
  return XATTN_OK;
}

// fusion.py:394-398, video side: v1 = LN(v0 + out_proj(attn)) and the per-clip operands of a2v_attn, as two tensor-core GEMMs
// over per-head folded weights (bf16_pack.cu).
int launch_video_update(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  auto F32 = [&](int64_t off) { return reinterpret_cast<const float*>(x.packed + off); };
  XF_TRY(gemm_tc_video_mix(x.st, x.ws + x.w.u, x.packed + x.pk.Gmix, F32(x.pk.bmix), reinterpret_cast<const float*>(x.ws + x.w.v0),
                           P(XP_VN_W), P(XP_VN_B), F32(x.pk.wcv), x.ws + x.w.v1, reinterpret_cast<float*>(x.ws + x.w.cv),
                           reinterpret_cast<float*>(x.ws + x.w.pooled), c.B, c.T));
  return gemm_tc_video_kv(x.st, x.ws + x.w.v1, x.packed + x.pk.Fkv, F32(x.pk.bkv), x.ws + x.w.kt, x.ws + x.w.vt, c.B);
}

int forward_bf16(const xattn_config& c, const float* params, const void* packed, const void* xv, const void* xa, float* logits,
                 void* ws, cudaStream_t st, bool fork) {
  XF_TRY(bf16_supported(c));
  NvtxRange nvtx_fwd("xattn_forward[bf16]");
  Bf16Ctx x;
  x.cfg = &c;
  x.params = params;
  compute_param_layout(c, x.pl);
  x.packed = reinterpret_cast<const uint8_t*>(packed);
  compute_pack_bf16(c, x.pk);
  x.ws = reinterpret_cast<uint8_t*>(ws);
  compute_ws_bf16(c, x.w);
  x.st = st;
  // The video branch up to Qt (v_in_proj :372 -> video_pre) only needs the video embeddings, is latency-bound and small;
  // the a0 projection (a_in_proj∘audio_seq_proj :378, :386) is one long HBM stream.  They run side by side: the video
  // branch is forked onto a private stream and joined before the first kernel that needs both.  With per-kernel timing on
  // everything stays on the caller's stream so the brackets measure one kernel each.
  cudaStream_t side = st;
  ForkRes* fr = nullptr;
  std::unique_lock<std::mutex> enqueue_lock;
  static const bool no_fork_env = getenv("XATTN_NO_FORK") != nullptr;
  if (fork && !no_fork_env && !g_stage_timing) {
    XF_TRY(fork_resources(&fr));
    enqueue_lock = std::unique_lock<std::mutex>(fr->enqueue);
    side = fr->side;
    XF_CUDA(cudaEventRecord(fr->fork, st));
    XF_CUDA(cudaStreamWaitEvent(side, fr->fork, 0));
  }
  int rc = XATTN_OK;
  {
    Bf16Ctx xs = x;
    xs.st = side;
    {
      StageScope t(ST_GEMM_V0, side);
      rc = gemm_tc_bf16(side, xv, x.packed + x.pk.Wv, params + x.pl.off[XP_V_IN_B], x.ws + x.w.v0, false, BE, c.B * c.T, BE, c.Dv);
    }
    if (rc == XATTN_OK) {
      StageScope t(ST_VIDEO_PRE, side);
      rc = launch_video_pre(xs);
    }
  }
  if (rc == XATTN_OK) {
    StageScope t(ST_GEMM_A0, st);
    rc = gemm_tc_bf16(st, xa, x.packed + x.pk.Wfold, reinterpret_cast<const float*>(x.packed + x.pk.bfold), x.ws + x.w.a0, true,
                      BOP, c.B * c.Ta, BE, c.Da, c.use_prior ? reinterpret_cast<const float*>(x.packed + x.pk.scw) : nullptr,
                      reinterpret_cast<float*>(x.ws + x.w.tok), reinterpret_cast<float*>(x.ws + x.w.acol), c.Ta);
  }
  if (fr != nullptr) {   // always rejoin, also on an error above, so the caller's stream stays ordered after the side work
    cudaEventRecord(fr->join, side);
    cudaStreamWaitEvent(st, fr->join, 0);
    enqueue_lock.unlock();
  }
  XF_TRY(rc);
  if (c.use_prior) { StageScope t(ST_PRIOR, st); XF_TRY(launch_prior(x)); }
  { StageScope t(ST_V2A, st); XF_TRY(launch_v2a_tc(x)); }
  { StageScope t(ST_VIDEO_UPD, st); XF_TRY(launch_video_update(x)); }
  { StageScope t(ST_A2V, st); XF_TRY(launch_a2v_tc(x)); }
  {
    StageScope t(ST_HEAD, st);
    if (c.pool_mode == XATTN_POOL_ATTN) XF_TRY(launch_learned_pool(x));
    XF_TRY(launch_head_pooled(x, logits));
  }
  return XATTN_OK;
}

}  // namespace xf
