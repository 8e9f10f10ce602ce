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
//   warp 0: TMA producer (a0 and Qt rings, one pair ahead)    warp 1: MMA issuer (S of pair i+1 right behind U of pair i)
//   warps 2-9: softmax + epilogue.  Warps w and w+4 share a TMEM sub-partition (= head) and split the 160 key columns of
//   every row in halves, so all eight warps work on the pair whose a0 tile is resident and the tile is released early.
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>

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
constexpr int kKaPitch = 164;              // floats; 656 B: the two clips' key scores start in different bank groups
constexpr int kKaStage = 2 * kKaPitch * 4;
// S [0,160) fp32 | P [160,240) bf16x2 | X [240,244): partial row max / sum exchanged between the two warps of a head | U [256,384)
constexpr uint32_t kTmemS = 0, kTmemP = 160, kTmemX = 240, kTmemU = 256;
constexpr float kL2E = 1.4426950408889634f;
// shared memory, no slack: the TMA rings fill all but 3 KB of the 227 KB a CTA can have
constexpr int kV2AOffKa = 2 * kA0Stage + 2 * kQtStage;       // [2 stages][2 clips][kKaPitch] f32: w_ak·a0_j, travels with the a0 stage
constexpr int kV2AOffBar = kV2AOffKa + 2 * kKaStage;         // 12 mbarriers + TMEM base
constexpr int kV2ASmem = kV2AOffBar + 12 * 8 + 8;
static_assert(kV2ASmem <= 227 * 1024, "v2a_tc_kernel shared memory");

struct V2ATcArgs {
  const float *tqv, *tok, *pc;   // [B,16] w_vq·v0_i ; [2,B,160] w_ak·a0_j | w_aq·a0_j ; [B,4]
  const float* bias_scale;
  __nv_bfloat16* u;              // [B,4,8,32] x uint4: mma A-fragment order, [clip][head][ks][lane]
  int B, Ta, use_prior;
  unsigned long long* trace;     // XATTN_TC_TRACE=1: [8 events][64 pairs] globaltimer stamps of CTA 0 (pipeline debugging)
};

__device__ __forceinline__ void trace_mark(unsigned long long* t, int ev, uint32_t it) {
  if (t != nullptr && blockIdx.x == 0 && it < 64) {
    unsigned long long now;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
    t[ev * 64 + it] = now;
  }
}

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// numerators of one strip of N key columns starting at column j0 of the row: e = 2^(S·log2e + tanh(qk + ka_j)·s·log2e - m')
template <int N>
__device__ __forceinline__ void v2a_strip(const uint32_t (&r)[N], uint32_t (&pk)[N / 2], const float* ka, int j0, int Ta, float qk,
                                          float bl, float nm, bool use_prior, float& l0, float& l1) {
  float e[N];
#pragma unroll
  for (int i = 0; i < N; i += 4) {
    float4 k4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (use_prior) k4 = *reinterpret_cast<const float4*>(ka + j0 + i);
    const float kk[4] = {k4.x, k4.y, k4.z, k4.w};
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      float x = fmaf(__uint_as_float(r[i + v]), kL2E, nm);
      if (use_prior) x = fmaf(tc::tanh_fast(qk + kk[v]), bl, x);
      e[i + v] = ex2_fast(x);
    }
  }
  if (j0 + N > Ta) {   // key padding (and whatever the padded score slots held) contributes nothing
#pragma unroll
    for (int i = 0; i < N; ++i)
      if (j0 + i >= Ta) e[i] = 0.f;
  }
#pragma unroll
  for (int i = 0; i < N; i += 2) {
    l0 += e[i];
    l1 += e[i + 1];
    pk[i >> 1] = tc::pack_bf16(e[i], e[i + 1]);
  }
}

__global__ void __launch_bounds__(kV2AThreads, 1) v2a_tc_kernel(const __grid_constant__ CUtensorMap tmA0,
                                                                 const __grid_constant__ CUtensorMap tmQt, V2ATcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a0s = smem;                              // [2 stages][2 clips][2 boxes]
  uint8_t* qts = smem + 2 * kA0Stage;               // [2 stages][2 clips][2 boxes]
  float* ka_s = reinterpret_cast<float*>(smem + kV2AOffKa);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kV2AOffBar);
  uint64_t *a0_full = bars, *a0_empty = bars + 2, *qt_full = bars + 4, *qt_empty = bars + 6, *s_full = bars + 8,
           *p_full = bars + 9, *u_full = bars + 10, *u_empty = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ta = a.Ta;
  const int npairs = (a.B + 1) / 2;
  const int n_local = (npairs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (threadIdx.x == 0) {
    if ((tc::smem_u32(smem) & 1023u) != 0) __trap();   // the swizzled TMA boxes need a 1024-byte aligned base
    tc::tma_prefetch_desc(&tmA0);
    tc::tma_prefetch_desc(&tmQt);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&a0_full[i], 1);
      tc::mbar_init(&a0_empty[i], 1);
      tc::mbar_init(&qt_full[i], 1);
      tc::mbar_init(&qt_empty[i], 1);
    }
    tc::mbar_init(s_full, 1);
    tc::mbar_init(p_full, 256);
    tc::mbar_init(u_full, 1);
    tc::mbar_init(u_empty, 256);
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
        while (!tc::mbar_try_wait(&qt_empty[s], ph ^ 1u)) __nanosleep(64);   // not latency critical: leave the issue slots to the softmax warps
        tc::mbar_expect_tx(&qt_full[s], kQtStage);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = qts + s * kQtStage + i * kQtClip;
          const int row = (2 * p + i) * 64;
          tc::tma_load_2d(dst, &tmQt, &qt_full[s], 0, row);
          tc::tma_load_2d(dst + kBox, &tmQt, &qt_full[s], 64, row);
        }
        while (!tc::mbar_try_wait(&a0_empty[s], ph ^ 1u)) __nanosleep(64);
        trace_mark(a.trace, 0, it);   // a0 stage free, loads issued
        tc::mbar_expect_tx(&a0_full[s], kA0Stage + (a.use_prior ? 2 * BTAP * 4 : 0));
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = a0s + s * kA0Stage + i * kA0Clip;
          const int row = (2 * p + i) * Ta;   // rows past the clip / the tensor are harmless: P is zero there, OOB reads as zero
          tc::tma_load_2d(dst, &tmA0, &a0_full[s], 0, row);
          tc::tma_load_2d(dst + kA0Box, &tmA0, &a0_full[s], 64, row);
          if (a.use_prior)   // the clip's key-side token scores ride along
            tc::bulk_load_1d(ka_s + (s * 2 + i) * kKaPitch, a.tok + (int64_t)min(2 * p + i, a.B - 1) * BTAP, BTAP * 4, &a0_full[s]);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = tc::umma_idesc_bf16_ex(64, BTAP, false, false);
      constexpr uint32_t idesc_u = tc::umma_idesc_bf16_ex(64, BE, false, true);
      auto issue_s = [&](uint32_t it) {
        const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
        tc::mbar_wait(&qt_full[s], ph);
        tc::mbar_wait(&a0_full[s], ph);
        trace_mark(a.trace, 1, it);   // operands landed
        tc::tc_fence_after();
        // descriptors differ from the stage's base descriptor only in the 16-byte-unit start address: add, don't rebuild
        const uint64_t da0 = tc::umma_desc_sw128(tc::smem_u32(qts + s * kQtStage), 16, 1024);
        const uint64_t db0 = tc::umma_desc_sw128(tc::smem_u32(a0s + s * kA0Stage), 16, 1024);
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemS + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t da = da0 + (uint64_t)((i * kQtClip + (ks >> 2) * kBox + (ks & 3) * 32) >> 4);
            const uint64_t db = db0 + (uint64_t)((i * kA0Clip + (ks >> 2) * kA0Box + (ks & 3) * 32) >> 4);
            tc::umma_bf16(d, da, db, idesc_s, ks != 0);
          }
        }
        tc::umma_commit(s_full);
        tc::umma_commit(&qt_empty[s]);
      };
      if (n_local > 0) issue_s(0);
      for (uint32_t it = 0; it < (uint32_t)n_local; ++it) {
        const uint32_t s = it & 1u;
        tc::mbar_wait(p_full, it & 1u);                          // all 128 rows of P are in tensor memory, S is dead
        if (it > 0) tc::mbar_wait(u_empty, (it - 1) & 1u);      // previous pair's U has left tensor memory
        trace_mark(a.trace, 2, it);   // P complete, U issued
        tc::tc_fence_after();
        const uint64_t db0 = tc::umma_desc_sw128(tc::smem_u32(a0s + s * kA0Stage), kA0Box, 1024);   // MN-major: K = token rows
        for (int i = 0; i < 2; ++i) {
          const uint32_t d = tmem + kTmemU + ((uint32_t)(16 * i) << 16);
          const uint32_t pa = tmem + kTmemP + ((uint32_t)(16 * i) << 16);
#pragma unroll
          for (int ks = 0; ks < BTAP / 16; ++ks)
            tc::umma_bf16_ts(d, pa + (uint32_t)(ks * 8), db0 + (uint64_t)((i * kA0Clip + ks * 16 * 128) >> 4), idesc_u, ks != 0);
        }
        tc::umma_commit(u_full);
        tc::umma_commit(&a0_empty[s]);
        if (it + 1 < (uint32_t)n_local) issue_s(it + 1);          // scores of the next pair while this pair's epilogue runs
      }
    }
  } else {
    // ---- softmax + epilogue warps.  Warp w may touch TMEM lanes [32*(w%4), +32): sub-partition == head. ----
    const int h = warp & 3, half = (warp - 2) >> 2;
    const int ci = lane >> 4, q = lane & 15;          // this thread's row: clip ci of the pair, head h, query q
    const uint32_t tlane = (uint32_t)(32 * h) << 16;
    const uint32_t tS = tmem + kTmemS + tlane + (uint32_t)(80 * half);
    const uint32_t tP = tmem + kTmemP + tlane + (uint32_t)(40 * half);
    const uint32_t tX = tmem + kTmemX + tlane;
    const float bsc = a.use_prior ? __ldg(a.bias_scale) : 0.f;
    const float bl = bsc * kL2E;
    const int g = lane >> 2;
    const int j0 = 80 * half;
    // qk = qv_i + the prior parts of both bias heads: w_vq·v0_i + pc[v_query] + pc[a_key]   (fusion.py:171, :173)
    auto qk_fetch = [&](int p, float& t0, float& t1, float& t2) {
      t0 = t1 = t2 = 0.f;
      if (a.use_prior && p < npairs) {
        const int c = min(2 * p + ci, a.B - 1);
        t0 = __ldg(a.tqv + (int64_t)c * 16 + q);
        t1 = __ldg(a.pc + (int64_t)c * 4);
        t2 = __ldg(a.pc + (int64_t)c * 4 + 1);
      }
    };
    float q0, q1, q2;
    qk_fetch(blockIdx.x, q0, q1, q2);
    uint32_t it = 0;
    for (int p = blockIdx.x; p < npairs; p += gridDim.x, ++it) {
      const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
      const float qk = q0 + q1 + q2;
      tc::mbar_wait(s_full, it & 1u);
      tc::mbar_wait(&a0_full[s], ph);                 // already complete (S used it); acquires the TMA-written key scores
      if (threadIdx.x == 64) trace_mark(a.trace, 3, it);   // worker: scores ready
      tc::tc_fence_after();
      uint32_t r0[32], r1[32], r2[16];                // this thread's 80 scores
      tc::tmem_ld_32x32(tS, r0);
      tc::tmem_ld_32x32(tS + 32, r1);
      tc::tmem_ld_32x32_x16(tS + 64, r2);
      tc::tmem_ld_wait();
      // row maximum of the raw scores over both halves.  tanh is bounded, so m + |s| bounds the biased row from above;
      // the softmax is shift-invariant.
      float m = -INFINITY;
      if (j0 + 80 <= Ta) {
#pragma unroll
        for (int i = 0; i < 32; ++i) m = fmaxf(m, fmaxf(__uint_as_float(r0[i]), __uint_as_float(r1[i])));
#pragma unroll
        for (int i = 0; i < 16; ++i) m = fmaxf(m, __uint_as_float(r2[i]));
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (j0 + i < Ta) m = fmaxf(m, __uint_as_float(r0[i]));
          if (j0 + 32 + i < Ta) m = fmaxf(m, __uint_as_float(r1[i]));
        }
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (j0 + 64 + i < Ta) m = fmaxf(m, __uint_as_float(r2[i]));
      }
      tc::tmem_st_32x32_x1(tX + (uint32_t)half, __float_as_uint(m));
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::named_bar_sync(2 + h, 64);
      tc::tc_fence_after();
      m = fmaxf(m, __uint_as_float(tc::tmem_ld_32x32_x1(tX + (uint32_t)(half ^ 1))));   // waits below, with the strip loads
      tc::tmem_ld_wait();
      if (threadIdx.x == 64) trace_mark(a.trace, 4, it);   // worker: row max known
      const float nm = -(m + fabsf(bsc)) * kL2E;
      // numerators -> bf16x2 -> tensor memory (the A operand of the U product); partial row sum
      float l0 = 0.f, l1 = 0.f;
      const float* ka = ka_s + (s * 2 + ci) * kKaPitch;
      {
        uint32_t pk[16];
        v2a_strip<32>(r0, pk, ka, j0, Ta, qk, bl, nm, a.use_prior, l0, l1);
        tc::tmem_st_32x32_x16(tP, pk);
        v2a_strip<32>(r1, pk, ka, j0 + 32, Ta, qk, bl, nm, a.use_prior, l0, l1);
        tc::tmem_st_32x32_x16(tP + 16, pk);
        uint32_t pk8[8];
        v2a_strip<16>(r2, pk8, ka, j0 + 64, Ta, qk, bl, nm, a.use_prior, l0, l1);
        tc::tmem_st_32x32_x8(tP + 32, pk8);
      }
      const float l = l0 + l1;
      tc::tc_fence_before();
      tc::named_bar_sync(2 + h, 64);                  // partner is past its epilogue of the previous pair: its sum slot is free
      tc::tc_fence_after();
      tc::tmem_st_32x32_x1(tX + 2u + (uint32_t)half, __float_as_uint(l));
      tc::tmem_st_wait();
      tc::tc_fence_before();
      tc::mbar_arrive(p_full);
      if (threadIdx.x == 64) trace_mark(a.trace, 5, it);   // worker: P written
      qk_fetch(p + gridDim.x, q0, q1, q2);            // next pair's row constants: in flight during the U product
      // ---- epilogue: this warp's head of clip `half`; U rows scaled by 1/l, bf16, mma A-fragment order ----
      tc::mbar_wait(u_full, it & 1u);                 // (completes after every p_full arrival: the partner's sums are in place)
      if (threadIdx.x == 64) trace_mark(a.trace, 6, it);   // worker: U ready
      tc::tc_fence_after();
      {
        uint32_t r[64];
        tc::tmem_ld_16x256_x16(tmem + kTmemU + tlane + ((uint32_t)(16 * half) << 16), r);
        const float lp = __uint_as_float(tc::tmem_ld_32x32_x1(tX + 2u + (uint32_t)(half ^ 1)));
        tc::tmem_ld_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(u_empty);
        const float lt = l + lp;                      // whole-row sum of this lane's row
        const float inv0 = 1.f / __shfl_sync(0xffffffffu, lt, 16 * half + g);
        const float inv1 = 1.f / __shfl_sync(0xffffffffu, lt, 16 * half + g + 8);
        const int cc = 2 * p + half;
        if (cc < a.B) {
          uint4* up = reinterpret_cast<uint4*>(a.u) + ((int64_t)cc * BH + h) * 256 + lane;   // [ks][lane]: coalesced
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            uint4 v;
            v.x = tc::pack_bf16(__uint_as_float(r[8 * ks + 0]) * inv0, __uint_as_float(r[8 * ks + 1]) * inv0);
            v.y = tc::pack_bf16(__uint_as_float(r[8 * ks + 2]) * inv1, __uint_as_float(r[8 * ks + 3]) * inv1);
            v.z = tc::pack_bf16(__uint_as_float(r[8 * ks + 4]) * inv0, __uint_as_float(r[8 * ks + 5]) * inv0);
            v.w = tc::pack_bf16(__uint_as_float(r[8 * ks + 6]) * inv1, __uint_as_float(r[8 * ks + 7]) * inv1);
            up[ks * 32] = v;
          }
        }
      }
      if (threadIdx.x == 64) trace_mark(a.trace, 7, it);   // worker: U stored
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

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
  a.trace = nullptr;
  static unsigned long long* trace_buf = nullptr;
  static const bool tracing = getenv("XATTN_TC_TRACE") != nullptr;
  if (tracing) {
    if (trace_buf == nullptr) XF_CUDA(cudaMalloc(&trace_buf, 8 * 64 * 8));
    XF_CUDA(cudaMemsetAsync(trace_buf, 0, 8 * 64 * 8, x.st));
    a.trace = trace_buf;
  }
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
  if (tracing) {   // debugging aid: timeline of CTA 0, microseconds from its first load
    static unsigned long long h[8 * 64];
    XF_CUDA(cudaStreamSynchronize(x.st));
    XF_CUDA(cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost));
    const char* names[8] = {"load_issue", "operands_in", "u_issue", "w_s_ready", "w_max", "w_p_done", "w_u_ready", "w_u_stored"};
    fprintf(stderr, "v2a_tc trace (us), B=%d\n%-12s", c.B, "pair");
    for (int e = 0; e < 8; ++e) fprintf(stderr, "%12s", names[e]);
    fprintf(stderr, "\n");
    for (int it = 0; it < 16 && h[it] != 0; ++it) {
      fprintf(stderr, "%-12d", it);
      for (int e = 0; e < 8; ++e) fprintf(stderr, "%12.2f", h[e * 64 + it] ? (double)(h[e * 64 + it] - h[0]) * 1e-3 : -1.0);
      fprintf(stderr, "\n");
    }
  }
  return XATTN_OK;
}

}  // namespace xf
