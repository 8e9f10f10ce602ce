// tcgen05 versions of the two per-clip attention kernels of the bf16 path (reference src/models/fusion.py:394-402).
// Both are persistent, warp-specialised kernels (one CTA per SM) that stream clips through shared memory with TMA and keep
// every score / output tile in TMEM; the register file only ever holds one 32-column strip of a row.
//
// v2a_tc_kernel — video queries audio (fusion.py:394), two clips per pipeline step:
//   S_c = Qt_c[64,128] · a0_c[Ta,128]^T      M=64 (4 heads x 16 queries), N=160, K=128; A, B K-major from TMA boxes
//   P_c = exp(S_c + tanh(qv_i + ka_j)·s - m)   one thread per (clip, head, query) row, straight from TMEM (fusion.py:175)
//   U_c = P_c[64,Ta] · a0_c[Ta,128]          M=64, N=128, K=160; A = P written to smem in the swizzled K-major layout,
//                                            B = THE SAME a0 tile read as an MN-major operand — a0 is loaded once
//   The two clips' accumulators share TMEM columns: clip 0 in lanes 0-15, clip 1 in lanes 16-31 of every sub-partition
//   (the M=64 data-path layout), so a warp's 32 lanes are 32 independent softmax rows and no row statistic crosses lanes.
//   U leaves in mma A-fragment order (16x256b TMEM loads return exactly that distribution) for video_update.
//   warp 0: TMA producer (a0 double-buffered one pair ahead, Qt)   warp 1: MMA issuer   warps 2-5: softmax + U epilogue
#include <cuda.h>

#include "bf16_layout.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kBox = 64 * 128;             // 8 KB: 64 rows x 64 bf16, 128-byte swizzle
constexpr int kA0Box = BTAP * 128;         // 20,480 B: 160 token rows x 64 features
constexpr int kA0Clip = 2 * kA0Box;        // features 0-63 | 64-127
constexpr int kA0Stage = 2 * kA0Clip;      // one pair of clips
constexpr int kPClip = 3 * kBox;           // P [64 rows][192 keys] = 3 boxes; Qt [64][128] = the first two (aliased)
constexpr int kV2AThreads = 192;
constexpr int kKaPitch = 176;              // floats; 176 % 32 == 16: the two clips' key scores sit in different banks
constexpr uint32_t kTmemS = 0, kTmemU = 256;
constexpr float kL2E = 1.4426950408889634f;

struct V2ATcArgs {
  const float *tqv, *tok, *pc;   // [B,16] w_vq·v0_i ; [B*Ta,2] (w_ak·a0_j, w_aq·a0_j) ; [B,4]
  const float* bias_scale;
  __nv_bfloat16* u;              // [B,4,32,8] x uint4: mma A-fragment order
  int B, Ta, use_prior;
};

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kV2AThreads, 1) v2a_tc_kernel(const __grid_constant__ CUtensorMap tmA0,
                                                                 const __grid_constant__ CUtensorMap tmQt, V2ATcArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* a0s = smem;                              // [2 stages][2 clips][2 boxes]
  uint8_t* qps = smem + 2 * kA0Stage;               // [2 clips] Qt (2 boxes) / P (3 boxes)
  float* ka_s = reinterpret_cast<float*>(qps + 2 * kPClip);   // [2][kKaPitch]
  float* qv_s = ka_s + 2 * kKaPitch;                // [2][16]
  float* invl_s = qv_s + 32;                        // [4 heads][2 clips][16]
  uint64_t* bars = reinterpret_cast<uint64_t*>(invl_s + 128);
  uint64_t *a0_full = bars, *a0_empty = bars + 2, *qt_full = bars + 4, *qp_free = bars + 5, *s_full = bars + 6,
           *p_full = bars + 7, *u_full = bars + 8, *u_empty = bars + 9;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ta = a.Ta;
  const int npairs = (a.B + 1) / 2;

  if (threadIdx.x == 0) {
    tc::tma_prefetch_desc(&tmA0);
    tc::tma_prefetch_desc(&tmQt);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&a0_full[i], 1);
      tc::mbar_init(&a0_empty[i], 1);
    }
    tc::mbar_init(qt_full, 1);
    tc::mbar_init(qp_free, 1);
    tc::mbar_init(s_full, 1);
    tc::mbar_init(p_full, 128);
    tc::mbar_init(u_full, 1);
    tc::mbar_init(u_empty, 128);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int p = blockIdx.x; p < npairs; p += gridDim.x, ++it) {
        const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
        tc::mbar_wait(&a0_empty[s], ph ^ 1u);
        tc::mbar_expect_tx(&a0_full[s], kA0Stage);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = a0s + s * kA0Stage + i * kA0Clip;
          const int row = (2 * p + i) * Ta;   // rows past the clip / the tensor are harmless: P is zero there, OOB reads as zero
          tc::tma_load_2d(dst, &tmA0, &a0_full[s], 0, row);
          tc::tma_load_2d(dst + kA0Box, &tmA0, &a0_full[s], 64, row);
        }
        tc::mbar_wait(qp_free, (it & 1u) ^ 1u);   // previous pair's U product (the last reader of P) has retired
        tc::mbar_expect_tx(qt_full, 4 * kBox);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = qps + i * kPClip;
          const int row = (2 * p + i) * 64;
          tc::tma_load_2d(dst, &tmQt, qt_full, 0, row);
          tc::tma_load_2d(dst + kBox, &tmQt, qt_full, 64, row);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = tc::umma_idesc_bf16_ex(64, BTAP, false, false);
      constexpr uint32_t idesc_u = tc::umma_idesc_bf16_ex(64, BE, false, true);
      uint32_t it = 0;
      for (int p = blockIdx.x; p < npairs; p += gridDim.x, ++it) {
        const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
        tc::mbar_wait(qt_full, it & 1u);
        tc::mbar_wait(&a0_full[s], ph);
        tc::tc_fence_after();
        const uint32_t a0b = tc::smem_u32(a0s + s * kA0Stage), qpb = tc::smem_u32(qps);
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemS + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t da = tc::umma_desc_sw128(qpb + i * kPClip + (ks >> 2) * kBox + (ks & 3) * 32, 16, 1024);
            const uint64_t db = tc::umma_desc_sw128(a0b + i * kA0Clip + (ks >> 2) * kA0Box + (ks & 3) * 32, 16, 1024);
            tc::umma_bf16(d, da, db, idesc_s, ks != 0);
          }
        }
        tc::umma_commit(s_full);
        tc::mbar_wait(p_full, it & 1u);           // all 128 rows of P are in smem (and S has been read)
        tc::mbar_wait(u_empty, (it & 1u) ^ 1u);   // previous pair's U has left TMEM
        tc::tc_fence_after();
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemU + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < BTAP / 16; ++ks) {
            const uint64_t da = tc::umma_desc_sw128(qpb + i * kPClip + (ks >> 2) * kBox + (ks & 3) * 32, 16, 1024);
            const uint64_t db = tc::umma_desc_sw128(a0b + i * kA0Clip + ks * 16 * 128, kA0Box, 1024);   // MN-major: K = token rows
            tc::umma_bf16(d, da, db, idesc_u, ks != 0);
          }
        }
        tc::umma_commit(u_full);
        tc::umma_commit(&a0_empty[s]);
        tc::umma_commit(qp_free);
      }
    }
  } else {
    // ---- softmax + epilogue warps.  Warp w may touch TMEM lanes [32*(w%4), +32): sub-partition == head. ----
    const int h = warp & 3, wt = threadIdx.x - 64;
    const int ci = lane >> 4, q = lane & 15;          // this thread's row: clip ci of the pair, head h, query q
    const int prow = 16 * h + q;                      // row of the 64-row P tile
    const uint32_t tlane = (uint32_t)(32 * h) << 16;
    const float bsc = a.use_prior ? __ldg(a.bias_scale) : 0.f;
    uint32_t it = 0;
    for (int p = blockIdx.x; p < npairs; p += gridDim.x, ++it) {
      // key-side bias inputs of the pair: ka_j = w_ak·a0_j + prior part (fusion.py:173), qv_i likewise (:171)
      if (a.use_prior) {
        for (int i = 0; i < 2; ++i) {
          const int c = min(2 * p + i, a.B - 1);
          const float pk = __ldg(a.pc + (int64_t)c * 4 + 1);
          for (int j = wt; j < Ta; j += 128) ka_s[i * kKaPitch + j] = __ldg(a.tok + ((int64_t)c * Ta + j) * 2) + pk;
        }
        if (wt < 32) {
          const int c = min(2 * p + (wt >> 4), a.B - 1);
          qv_s[wt] = __ldg(a.tqv + (int64_t)c * 16 + (wt & 15)) + __ldg(a.pc + (int64_t)c * 4);
        }
      }
      tc::named_bar_sync(1, 128);
      const float qv = a.use_prior ? qv_s[ci * 16 + q] : 0.f;
      const float* ka = ka_s + ci * kKaPitch;
      tc::mbar_wait(s_full, it & 1u);
      tc::tc_fence_after();
      // pass 1: row maximum of the raw scores.  tanh is bounded, so m = max + |s| bounds the biased row from above and
      // the softmax is shift-invariant; the row never has to be held in registers.
      float m = -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < BTAP; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem + kTmemS + tlane + (uint32_t)c0, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c0 + i < Ta) m = fmaxf(m, __uint_as_float(r[i]));
      }
      const float nm = -(m + fabsf(bsc)) * kL2E;
      // pass 2: numerators -> bf16 -> the swizzled K-major A tile of the U product; row sum for the epilogue
      float l = 0.f;
      uint8_t* prow_base = qps + ci * kPClip;
#pragma unroll 1
      for (int c0 = 0; c0 < BTAP; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem + kTmemS + tlane + (uint32_t)c0, r);
        tc::tmem_ld_wait();
        float e[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int j = c0 + i;
          float sc = __uint_as_float(r[i]);
          if (a.use_prior) sc = fmaf(tc::tanh_fast(qv + ka[min(j, BTAP - 1)]), bsc, sc);
          e[i] = j < Ta ? ex2_fast(fmaf(sc, kL2E, nm)) : 0.f;
          l += e[i];
        }
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          uint4 pk;
          pk.x = tc::pack_bf16(e[8 * v + 0], e[8 * v + 1]);
          pk.y = tc::pack_bf16(e[8 * v + 2], e[8 * v + 3]);
          pk.z = tc::pack_bf16(e[8 * v + 4], e[8 * v + 5]);
          pk.w = tc::pack_bf16(e[8 * v + 6], e[8 * v + 7]);
          const int col = c0 + 8 * v;
          *reinterpret_cast<uint4*>(prow_base + (col >> 6) * kBox + tc::sw128_off(prow, col & 63)) = pk;
        }
      }
      invl_s[(h * 2 + ci) * 16 + q] = 1.f / l;
      tc::fence_proxy_async();   // generic-proxy writes of P -> visible to the tensor core (async proxy)
      tc::tc_fence_before();
      tc::mbar_arrive(p_full);
      // ---- epilogue: U rows scaled by 1/l, bf16, mma A-fragment order ----
      tc::mbar_wait(u_full, it & 1u);
      tc::tc_fence_after();
      __syncwarp();
      const int g = lane >> 2, t = lane & 3;
      (void)t;
#pragma unroll 1
      for (int i = 0; i < 2; ++i) {
        uint32_t r[64];
        tc::tmem_ld_16x256_x16(tmem + kTmemU + tlane + ((uint32_t)(16 * i) << 16), r);
        tc::tmem_ld_wait();
        const int c = 2 * p + i;
        if (c < a.B) {
          const float inv0 = invl_s[(h * 2 + i) * 16 + g], inv1 = invl_s[(h * 2 + i) * 16 + g + 8];
          uint4* up = reinterpret_cast<uint4*>(a.u) + (((int64_t)c * BH + h) * 32 + lane) * 8;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            uint4 v;
            v.x = tc::pack_bf16(__uint_as_float(r[8 * ks + 0]) * inv0, __uint_as_float(r[8 * ks + 1]) * inv0);
            v.y = tc::pack_bf16(__uint_as_float(r[8 * ks + 2]) * inv1, __uint_as_float(r[8 * ks + 3]) * inv1);
            v.z = tc::pack_bf16(__uint_as_float(r[8 * ks + 4]) * inv0, __uint_as_float(r[8 * ks + 5]) * inv0);
            v.w = tc::pack_bf16(__uint_as_float(r[8 * ks + 6]) * inv1, __uint_as_float(r[8 * ks + 7]) * inv1);
            up[ks] = v;
          }
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive(u_empty);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

constexpr int kV2ASmem = 2 * kA0Stage + 2 * kPClip + (2 * kKaPitch + 32 + 128) * 4 + 16 * 8 + 1024;

}  // namespace

int launch_v2a_tc(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  V2ATcArgs a;
  a.tqv = reinterpret_cast<const float*>(x.ws + x.w.tqv);
  a.tok = reinterpret_cast<const float*>(x.ws + x.w.tok);
  a.pc = reinterpret_cast<const float*>(x.ws + x.w.pc);
  a.bias_scale = P(XP_BIAS_SCALE);
  a.u = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.u);
  a.B = c.B; a.Ta = c.Ta; a.use_prior = c.use_prior;
  alignas(64) CUtensorMap tmA0, tmQt;
  XF_TRY(make_tmap_bf16_2d(&tmA0, x.ws + x.w.a0, (int64_t)c.B * c.Ta, BE, BTAP, 64, BEP));
  XF_TRY(make_tmap_bf16_2d(&tmQt, x.ws + x.w.qt, (int64_t)c.B * 64, BE, 64, 64, BEP));
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(v2a_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kV2ASmem));
    configured = true;
  }
  const int npairs = (c.B + 1) / 2;
  const int grid = npairs < num_sms() ? npairs : num_sms();
  v2a_tc_kernel<<<grid, kV2AThreads, kV2ASmem, x.st>>>(tmA0, tmQt, a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
