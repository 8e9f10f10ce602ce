// tcgen05 versions of the two per-clip attention kernels of the bf16 path (reference src/models/fusion.py:394-402).
// Both are persistent, warp-specialised kernels (one CTA per SM) that stream clips through shared memory with TMA and keep
// every score / probability / output tile in TMEM; the register file only ever holds one 32-column strip of a row.
//
// v2a_tc_kernel — video queries audio (fusion.py:394), two clips per pipeline step:
//   S_c = Qt_c[64,128] · a0_c[Ta,128]^T      M=64 (4 heads x 16 queries), N=160, K=128; A, B K-major from TMA boxes
//   P_c = exp(S_c + tanh(qv_i + ka_j)·s - m)   one thread per (clip, head, query) row, straight from TMEM (fusion.py:175);
//                                            the bf16 numerators are written back over S with tcgen05.st
//   U_c = P_c[64,Ta] · a0_c[Ta,128]          M=64, N=128, K=160; A = P read from TENSOR MEMORY, B = THE SAME a0 tile read
//                                            as an MN-major operand — a0 is loaded once, P never touches shared memory
//   The two clips' accumulators share TMEM columns: clip 0 in lanes 0-15, clip 1 in lanes 16-31 of every sub-partition
//   (the M=64 data-path layout), so a warp's 32 lanes are 32 independent softmax rows and no row statistic crosses lanes.
//   U leaves in mma A-fragment order (16x256b TMEM loads return exactly that distribution) for video_update.
//   warp 0: TMA producer (a0 and Qt rings, one pair ahead)    warp 1: MMA issuer — S of pair i+1 is issued before U of pair i
//   warps 2-5 / 6-9: two softmax + epilogue groups that alternate pairs, each with its own S/P accumulator.
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
constexpr int kQtClip = 2 * kBox;          // Qt [64][128]
constexpr int kQtStage = 2 * kQtClip;
constexpr int kV2AThreads = 64 + 256;
constexpr uint32_t kTmemS = 0, kTmemSStride = 160, kTmemU = 320;   // S/P of group 0 | S/P of group 1 | U
constexpr float kL2E = 1.4426950408889634f;

struct V2ATcArgs {
  const float *tqv, *tok, *pc;   // [B,16] w_vq·v0_i ; [2,B*Ta] w_ak·a0_j | w_aq·a0_j ; [B,4]
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
  uint8_t* qts = smem + 2 * kA0Stage;               // [2 stages][2 clips][2 boxes]
  uint64_t* bars = reinterpret_cast<uint64_t*>(qts + 2 * kQtStage);
  uint64_t *a0_full = bars, *a0_empty = bars + 2, *qt_full = bars + 4, *qt_empty = bars + 6, *s_full = bars + 8,
           *p_full = bars + 10, *u_full = bars + 12, *u_empty = bars + 14;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ta = a.Ta;
  const int npairs = (a.B + 1) / 2;

  if (threadIdx.x == 0) {
    tc::tma_prefetch_desc(&tmA0);
    tc::tma_prefetch_desc(&tmQt);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&a0_full[i], 1);
      tc::mbar_init(&a0_empty[i], 1);
      tc::mbar_init(&qt_full[i], 1);
      tc::mbar_init(&qt_empty[i], 1);
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&p_full[i], 128);
      tc::mbar_init(&u_full[i], 1);      // per group: a group sees every second pair, so a shared barrier's parity would alias
      tc::mbar_init(&u_empty[i], 128);
    }
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
        tc::mbar_wait(&qt_empty[s], ph ^ 1u);
        tc::mbar_expect_tx(&qt_full[s], kQtStage);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = qts + s * kQtStage + i * kQtClip;
          const int row = (2 * p + i) * 64;
          tc::tma_load_2d(dst, &tmQt, &qt_full[s], 0, row);
          tc::tma_load_2d(dst + kBox, &tmQt, &qt_full[s], 64, row);
        }
        tc::mbar_wait(&a0_empty[s], ph ^ 1u);
        tc::mbar_expect_tx(&a0_full[s], kA0Stage);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = a0s + s * kA0Stage + i * kA0Clip;
          const int row = (2 * p + i) * Ta;   // rows past the clip / the tensor are harmless: P is zero there, OOB reads as zero
          tc::tma_load_2d(dst, &tmA0, &a0_full[s], 0, row);
          tc::tma_load_2d(dst + kA0Box, &tmA0, &a0_full[s], 64, row);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = tc::umma_idesc_bf16_ex(64, BTAP, false, false);
      constexpr uint32_t idesc_u = tc::umma_idesc_bf16_ex(64, BE, false, true);
      const int n_local = (npairs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      auto s_ready = [&](uint32_t it) {
        const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
        return tc::mbar_try_wait(&qt_full[s], ph) && tc::mbar_try_wait(&a0_full[s], ph);
      };
      auto issue_s = [&](uint32_t it) {   // S of local pair `it` into the accumulator of group it & 1
        const uint32_t s = it & 1u;
        tc::tc_fence_after();
        const uint32_t a0b = tc::smem_u32(a0s + s * kA0Stage), qtb = tc::smem_u32(qts + s * kQtStage);
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemS + s * kTmemSStride + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t da = tc::umma_desc_sw128(qtb + i * kQtClip + (ks >> 2) * kBox + (ks & 3) * 32, 16, 1024);
            const uint64_t db = tc::umma_desc_sw128(a0b + i * kA0Clip + (ks >> 2) * kA0Box + (ks & 3) * 32, 16, 1024);
            tc::umma_bf16(d, da, db, idesc_s, ks != 0);
          }
        }
        tc::umma_commit(&s_full[s]);
        tc::umma_commit(&qt_empty[s]);
      };
      auto u_ready = [&](uint32_t it) {   // P of the pair is in tensor memory and the previous pair's U has been drained
        if (!tc::mbar_try_wait(&p_full[it & 1u], (it >> 1) & 1u)) return false;
        return it == 0 || tc::mbar_try_wait(&u_empty[(it - 1) & 1u], ((it - 1) >> 1) & 1u);
      };
      auto issue_u = [&](uint32_t it) {
        const uint32_t s = it & 1u;
        tc::tc_fence_after();
        const uint32_t a0b = tc::smem_u32(a0s + s * kA0Stage);
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemU + ((uint32_t)(16 * i) << 16);
          const uint32_t pa = tmem + kTmemS + s * kTmemSStride + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < BTAP / 16; ++ks) {
            const uint64_t db = tc::umma_desc_sw128(a0b + i * kA0Clip + ks * 16 * 128, kA0Box, 1024);   // MN-major: K = token rows
            tc::umma_bf16_ts(d, pa + (uint32_t)(ks * 8), db, idesc_u, ks != 0);
          }
        }
        tc::umma_commit(&u_full[s]);
        tc::umma_commit(&a0_empty[s]);
      };
      // The tensor pipe runs ahead: S of pair it+1 is issued as soon as its operands have landed, before or after U of pair
      // it, whichever is ready first.  S(it+1) reuses the accumulator that held P(it-1); U(it-1) — its last reader — was
      // issued earlier and the pipe executes in order.
      if (n_local > 0) {
        while (!s_ready(0)) {
        }
        issue_s(0);
      }
      for (uint32_t it = 0; it < (uint32_t)n_local; ++it) {
        bool s_done = it + 1 >= (uint32_t)n_local, u_done = false;
        while (!s_done || !u_done) {
          if (!s_done && s_ready(it + 1)) {
            issue_s(it + 1);
            s_done = true;
          }
          if (!u_done && u_ready(it)) {
            issue_u(it);
            u_done = true;
          }
        }
      }
    }
  } else {
    // ---- softmax + epilogue warps.  Warp w may touch TMEM lanes [32*(w%4), +32): sub-partition == head. ----
    const int h = warp & 3, grp = (warp - 2) >> 2;
    const int ci = lane >> 4, q = lane & 15;          // this thread's row: clip ci of the pair, head h, query q
    const uint32_t tlane = (uint32_t)(32 * h) << 16;
    const uint32_t tS = tmem + kTmemS + grp * kTmemSStride + tlane;
    const float bsc = a.use_prior ? __ldg(a.bias_scale) : 0.f;
    const int g = lane >> 2;
    uint32_t it = grp;
    for (int p = blockIdx.x + grp * gridDim.x; p < npairs; p += 2 * gridDim.x, it += 2) {
      const uint32_t ph = (it >> 1) & 1u;
      // bias inputs of this row: qv_i = w_vq·v0_i + prior part (fusion.py:171); ka_j = w_ak·a0_j + prior part (:173)
      const int c = min(2 * p + ci, a.B - 1);
      float qv = 0.f, pk = 0.f;
      if (a.use_prior) {
        qv = __ldg(a.tqv + (int64_t)c * 16 + q) + __ldg(a.pc + (int64_t)c * 4);
        pk = __ldg(a.pc + (int64_t)c * 4 + 1);
      }
      const float* ka = a.tok + (int64_t)c * Ta;
      tc::mbar_wait(&s_full[grp], ph);
      tc::tc_fence_after();
      // pass 1: row maximum of the raw scores.  tanh is bounded, so m = max + |s| bounds the biased row from above and
      // the softmax is shift-invariant; the row never has to be held in registers.
      float m = -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < BTAP; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tS + (uint32_t)c0, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (c0 + i < Ta) m = fmaxf(m, __uint_as_float(r[i]));
      }
      const float nm = -(m + fabsf(bsc)) * kL2E;
      // pass 2: numerators -> bf16x2 -> back into tensor memory over the scores already consumed (16 P columns replace
      // S columns [16k, 16k+16) after S columns [32k, 32k+32) have been read); row sum for the epilogue
      float l = 0.f;
#pragma unroll 1
      for (int c0 = 0; c0 < BTAP; c0 += 32) {
        float kav[32];
        if (a.use_prior) {
#pragma unroll
          for (int i = 0; i < 32; ++i) kav[i] = __ldg(ka + min(c0 + i, Ta - 1));
        }
        uint32_t r[32];
        tc::tmem_ld_32x32(tS + (uint32_t)c0, r);
        tc::tmem_ld_wait();
        uint32_t pk2[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float e0, e1;
          {
            float sc = __uint_as_float(r[i]);
            if (a.use_prior) sc = fmaf(tc::tanh_fast(qv + kav[i] + pk), bsc, sc);
            e0 = c0 + i < Ta ? ex2_fast(fmaf(sc, kL2E, nm)) : 0.f;
          }
          {
            float sc = __uint_as_float(r[i + 1]);
            if (a.use_prior) sc = fmaf(tc::tanh_fast(qv + kav[i + 1] + pk), bsc, sc);
            e1 = c0 + i + 1 < Ta ? ex2_fast(fmaf(sc, kL2E, nm)) : 0.f;
          }
          l += e0 + e1;
          pk2[i >> 1] = tc::pack_bf16(e0, e1);
        }
        tc::tmem_st_32x32_x16(tS + (uint32_t)(c0 >> 1), pk2);
      }
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(&p_full[grp]);
      const float inv_l = 1.f / l;
      // ---- epilogue: U rows scaled by 1/l, bf16, mma A-fragment order ----
      tc::mbar_wait(&u_full[grp], ph);
      tc::tc_fence_after();
#pragma unroll 1
      for (int i = 0; i < 2; ++i) {
        uint32_t r[64];
        tc::tmem_ld_16x256_x16(tmem + kTmemU + tlane + ((uint32_t)(16 * i) << 16), r);
        tc::tmem_ld_wait();
        const float inv0 = __shfl_sync(0xffffffffu, inv_l, 16 * i + g), inv1 = __shfl_sync(0xffffffffu, inv_l, 16 * i + g + 8);
        const int cc = 2 * p + i;
        if (cc < a.B) {
          uint4* up = reinterpret_cast<uint4*>(a.u) + (((int64_t)cc * BH + h) * 32 + lane) * 8;
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
      tc::mbar_arrive(&u_empty[grp]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

constexpr int kV2ASmem = 2 * kA0Stage + 2 * kQtStage + 18 * 8 + 1024;

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
  static_assert(kV2ASmem <= 227 * 1024, "v2a_tc_kernel shared memory");
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
