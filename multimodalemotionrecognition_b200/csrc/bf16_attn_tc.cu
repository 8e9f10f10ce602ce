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
//   U leaves row-major as [16·clip + i][128h + e] bf16, the A operand of the video_mix GEMM (gemm_tc.cuh).
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
  __nv_bfloat16* u;              // [16·B, 4·128]: row (clip, i), columns 128h + e
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
                                                                 const __grid_constant__ CUtensorMap tmQt,
                                                                 const __grid_constant__ CUtensorMap tmU, V2ATcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a0s = smem;                              // [2 stages][2 clips][2 boxes]
  uint8_t* qts = smem + 2 * kA0Stage;               // [2 stages][2 clips][2 boxes]
  float* ka_s = reinterpret_cast<float*>(smem + kV2AOffKa);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kV2AOffBar);
  uint64_t *a0_full = bars, *a0_empty = bars + 2, *qt_full = bars + 4, *u_staged = bars + 6, *s_full = bars + 8,
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
    tc::tma_prefetch_desc(&tmU);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&a0_full[i], 1);
      tc::mbar_init(&a0_empty[i], 1);
      tc::mbar_init(&qt_full[i], 1);
      tc::mbar_init(&u_staged[i], 256);   // the pair's U rows are staged in its (dead) Qt stage
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
  tc::pdl_wait();     // everything above overlapped the previous kernel's tail; its results are visible from here on
  tc::pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      // Producer: keeps the a0 ring full, and owns the Qt stages end to end — pair i's Qt stage is dead once its scores exist,
      // the softmax warps stage the pair's U rows in it, this thread sends them out with the copy engine (whole lines) and, when
      // the engine has read them, refills the stage with the Qt of pair i + 2.
      auto load_qt = [&](uint32_t it) {
        const uint32_t s = it & 1u;
        const int p = blockIdx.x + (int)it * (int)gridDim.x;
        tc::mbar_expect_tx(&qt_full[s], kQtStage);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = qts + s * kQtStage + i * kQtClip;
          const int row = (2 * p + i) * 64;
          tc::tma_load_2d(dst, &tmQt, &qt_full[s], 0, row);
          tc::tma_load_2d(dst + kBox, &tmQt, &qt_full[s], 64, row);
        }
      };
      auto load_a0 = [&](uint32_t it) {
        const uint32_t s = it & 1u;
        const int p = blockIdx.x + (int)it * (int)gridDim.x;
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
      };
      const uint32_t nl = (uint32_t)(n_local > 0 ? n_local : 0);
      uint32_t na = 0, nst = 0;   // next pair whose a0 is to be loaded / whose U is to be stored
      for (; na < 2 && na < nl; ++na) {
        load_qt(na);
        load_a0(na);
      }
      while (nst < nl) {
        bool progressed = false;
        if (na < nl && tc::mbar_test_wait(&a0_empty[na & 1u], ((na >> 1) & 1u) ^ 1u)) {
          load_a0(na++);
          progressed = true;
        }
        if (tc::mbar_test_wait(&u_staged[nst & 1u], (nst >> 1) & 1u)) {
          const uint32_t s = nst & 1u;
          const int p = blockIdx.x + (int)nst * (int)gridDim.x;
#pragma unroll
          for (int g8 = 0; g8 < 8; ++g8) tc::tma_store_2d(&tmU, qts + s * kQtStage + g8 * 4096, 64 * g8, 32 * p);
          tc::bulk_commit_group();
          tc::bulk_store_wait_read();
          if (nst + 2 < nl) load_qt(nst + 2);
          ++nst;
          progressed = true;
        }
        if (!progressed) __nanosleep(64);   // not latency critical: leave the issue slots to the softmax warps
      }
      tc::bulk_store_wait_all();
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
      // ---- epilogue: this warp's head, both clips (lane = (clip, row)), columns [64·half, +64); rows scaled by 1/l, bf16,
      //      row-major [16·clip + q][128h + e]: the A operand of the video_mix GEMM ----
      tc::mbar_wait(u_full, it & 1u);                 // (completes after every p_full arrival: the partner's sums are in place)
      if (threadIdx.x == 64) trace_mark(a.trace, 6, it);   // worker: U ready
      tc::tc_fence_after();
      {
        uint32_t r0u[32], r1u[32];
        tc::tmem_ld_32x32(tmem + kTmemU + tlane + (uint32_t)(64 * half), r0u);
        tc::tmem_ld_32x32(tmem + kTmemU + tlane + (uint32_t)(64 * half + 32), r1u);
        const float lp = __uint_as_float(tc::tmem_ld_32x32_x1(tX + 2u + (uint32_t)(half ^ 1)));
        tc::tmem_ld_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(u_empty);
        const float inv = 1.f / (l + lp);             // whole-row sum of this lane's row
        // 128 bytes of this lane's row -> the pair's Qt stage, as box (2h + half) of eight [32 rows][64 columns] swizzled boxes
        // (rows of clips past the batch are clipped by the tensor map)
        {
          uint8_t* stg = qts + s * kQtStage + (2 * h + half) * 4096 + lane * 128;
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            uint4 w;
            w.x = tc::pack_bf16(__uint_as_float(r0u[8 * v + 0]) * inv, __uint_as_float(r0u[8 * v + 1]) * inv);
            w.y = tc::pack_bf16(__uint_as_float(r0u[8 * v + 2]) * inv, __uint_as_float(r0u[8 * v + 3]) * inv);
            w.z = tc::pack_bf16(__uint_as_float(r0u[8 * v + 4]) * inv, __uint_as_float(r0u[8 * v + 5]) * inv);
            w.w = tc::pack_bf16(__uint_as_float(r0u[8 * v + 6]) * inv, __uint_as_float(r0u[8 * v + 7]) * inv);
            *reinterpret_cast<uint4*>(stg + (((v ^ lane) & 7) << 4)) = w;
          }
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            uint4 w;
            w.x = tc::pack_bf16(__uint_as_float(r1u[8 * v + 0]) * inv, __uint_as_float(r1u[8 * v + 1]) * inv);
            w.y = tc::pack_bf16(__uint_as_float(r1u[8 * v + 2]) * inv, __uint_as_float(r1u[8 * v + 3]) * inv);
            w.z = tc::pack_bf16(__uint_as_float(r1u[8 * v + 4]) * inv, __uint_as_float(r1u[8 * v + 5]) * inv);
            w.w = tc::pack_bf16(__uint_as_float(r1u[8 * v + 6]) * inv, __uint_as_float(r1u[8 * v + 7]) * inv);
            *reinterpret_cast<uint4*>(stg + ((((4 + v) ^ lane) & 7) << 4)) = w;
          }
          tc::fence_proxy_async();
          tc::mbar_arrive(&u_staged[s]);
        }
      }
      if (threadIdx.x == 64) trace_mark(a.trace, 7, it);   // worker: U stored
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem, 512);
}


// ---------------------------------------------------------------------------------------------------------------------------
// a2v_tc_kernel — audio queries the updated video (fusion.py:398), residual + LayerNorm (:399), temporal mean (:402).
// Two clips per pipeline step, their audio tokens stacked as 320 rows (clip 1 starts at row 160) and cut into three row
// tiles: rows [0,128) and [128,256) as M=128 products, rows [256,320) as an M=64 product.  160 = 5 x 32, so every warp's
// 32 rows belong to one clip.
//   S  = a0_rows[M,128] · [Kt_0 ; Kt_1]^T        N = 128: columns 0-63 are the (head, key) scores against clip 0's video,
//                                                64-127 against clip 1's; a row reads the half of its own clip
//   P  = per-head softmax over the 16 keys of (S + c_hi + tanh(qa_j + kv_i)·s)    one thread per token row   (fusion.py:176)
//        written back over S as bf16x2, the other clip's half as zeros
//   O  = P[M,128] · [Vt_0 ; Vt_1]                A = P from tensor memory, B = the two Vt tiles stacked along K (MN-major);
//                                                the zero half of P selects the clip, out_proj is already folded into Vt
//   y  = LayerNorm(a0 + O) without the affine (out_proj's bias bo2 rides in Vt), rows kept in registers; the 32 rows of a warp are summed with a
//        shuffle butterfly (4 columns per lane come out) and leave as one partial per (clip, tile, warp); the head kernel
//        adds a clip's partials in a fixed order and applies gamma/Ta, beta (mean_t LN(x) = gamma·mean xhat + beta).
//   warp 0: TMA producer   warp 1: MMA issuer, polls for whichever of {next S, next O} is ready
//   warps 2-9 / 10-17: two softmax+LayerNorm groups alternating row tiles, each with its own 256-column TMEM buffer.
//   Inside a group, warps w and w+4 share a TMEM sub-partition (the same 32 rows) and split the work in halves: two of
//   the four heads in the softmax, 64 of the 128 columns in the LayerNorm; the row statistics meet in two output columns
//   the thread has already loaded.
constexpr int kA2VThreads = 64 + 512;
constexpr int kA2VHalf = 2 * kA0Box;                         // a0 stage layout: [feature half][clip][160 rows][128 B]
constexpr int kKvHalf = 2 * kBox;                            // Kt / Vt layout:  [feature half][clip][64 rows][128 B]
constexpr int kKvTile = 2 * kKvHalf;                         // 32 KB: both clips
constexpr int kA2VOffKt = 2 * kA0Stage;
constexpr int kA2VOffVt = kA2VOffKt + kKvTile;
constexpr int kA2VOffKc = kA2VOffVt + kKvTile;               // [2 stages][2 clips][80] f32: kv_i (16) | c_hi (64), travels with the a0 stage
constexpr int kA2VOffBar = kA2VOffKc + 4 * 80 * 4;
constexpr int kA2VSmem = kA2VOffBar + 16 * 8 + 8;
static_assert(kA2VSmem <= 227 * 1024, "a2v_tc_kernel shared memory");
constexpr int kPoolSlots = 8;                                // partials per clip: even clip 5 (tile 0 x4, tile 1 warp 0), odd clip 7

struct A2VTcArgs {
  const float *tok, *tkv, *pc, *cv;    // [2,B,160], [B,16], [B,4], [B,64]
  const float* bias_scale;
  float* poolp;                        // [B][kPoolSlots][128] partial sums of xhat over a warp's rows
  __nv_bfloat16* xhat_out;             // [B][Ta][128] the normalised rows themselves (learned pooling, temporal.py:9-26) or nullptr
  int B, Ta, T, use_prior;
  unsigned long long* trace;
};

template <bool EMIT>   // EMIT: also write the normalised audio rows (A2VTcArgs::xhat_out) for the learned pooling
__global__ void __launch_bounds__(kA2VThreads, 1) a2v_tc_kernel(const __grid_constant__ CUtensorMap tmA0,
                                                                 const __grid_constant__ CUtensorMap tmKt,
                                                                 const __grid_constant__ CUtensorMap tmVt, A2VTcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a0s = smem;
  uint8_t* kts = smem + kA2VOffKt;
  uint8_t* vts = smem + kA2VOffVt;
  float* kc_s = reinterpret_cast<float*>(smem + kA2VOffKc);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kA2VOffBar);
  uint64_t *a0_full = bars, *a0_empty = bars + 2, *kt_full = bars + 4, *kt_empty = bars + 5, *vt_full = bars + 6,
           *vt_empty = bars + 7, *s_full = bars + 8, *p_full = bars + 10, *o_full = bars + 12, *o_empty = bars + 14;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ta = a.Ta;
  const int npairs = (a.B + 1) / 2;
  const int n_local = (npairs - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const uint32_t n_tiles = 3u * (uint32_t)(n_local > 0 ? n_local : 0);

  if (threadIdx.x == 0) {
    if ((tc::smem_u32(smem) & 1023u) != 0) __trap();
    tc::tma_prefetch_desc(&tmA0);
    tc::tma_prefetch_desc(&tmKt);
    tc::tma_prefetch_desc(&tmVt);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&a0_full[i], 1);
      tc::mbar_init(&a0_empty[i], 768);   // released by the rows themselves, after their residual read (3 tiles x 256 threads)
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&p_full[i], 256);
      tc::mbar_init(&o_full[i], 1);
      tc::mbar_init(&o_empty[i], 256);
    }
    tc::mbar_init(kt_full, 1);
    tc::mbar_init(kt_empty, 1);
    tc::mbar_init(vt_full, 1);
    tc::mbar_init(vt_empty, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  tc::pdl_wait();     // everything above overlapped the previous kernel's tail; its results are visible from here on
  tc::pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int p = blockIdx.x; p < npairs; p += gridDim.x, ++it) {
        const uint32_t s = it & 1u, ph = (it >> 1) & 1u;
        const int c0 = 2 * p, c1 = min(2 * p + 1, a.B - 1);
        while (!tc::mbar_try_wait(&a0_empty[s], ph ^ 1u)) __nanosleep(64);
        trace_mark(a.trace, 0, 3 * it);   // a0 stage free, loads issued
        tc::mbar_expect_tx(&a0_full[s], kA0Stage + 2 * 80 * 4);
        for (int i = 0; i < 2; ++i) {
          uint8_t* dst = a0s + s * kA0Stage + i * kA0Box;
          const int row = (2 * p + i) * Ta;   // rows past the clip / the tensor never reach the result (masked rows, OOB = 0)
          tc::tma_load_2d(dst, &tmA0, &a0_full[s], 0, row);
          tc::tma_load_2d(dst + kA2VHalf, &tmA0, &a0_full[s], 64, row);
          const int c = i == 0 ? c0 : c1;   // the clip's key-side constants live and die with its a0 tile
          tc::bulk_load_1d(kc_s + (s * 2 + i) * 80, a.tkv + (int64_t)c * 16, 64, &a0_full[s]);
          tc::bulk_load_1d(kc_s + (s * 2 + i) * 80 + 16, a.cv + (int64_t)c * 64, 256, &a0_full[s]);
        }
        // Kt (+ the per-clip key constants) — free once the three score products of the previous pair have retired
        while (!tc::mbar_try_wait(kt_empty, (it & 1u) ^ 1u)) __nanosleep(64);
        tc::mbar_expect_tx(kt_full, kKvTile);
        for (int i = 0; i < 2; ++i) {
          const int c = i == 0 ? c0 : c1;
          tc::tma_load_2d(kts + i * kBox, &tmKt, kt_full, 0, c * 64);
          tc::tma_load_2d(kts + kKvHalf + i * kBox, &tmKt, kt_full, 64, c * 64);
        }
        while (!tc::mbar_try_wait(vt_empty, (it & 1u) ^ 1u)) __nanosleep(64);
        tc::mbar_expect_tx(vt_full, kKvTile);
        for (int i = 0; i < 2; ++i) {
          const int c = i == 0 ? c0 : c1;
          tc::tma_load_2d(vts + i * kBox, &tmVt, vt_full, 0, c * 64);
          tc::tma_load_2d(vts + kKvHalf + i * kBox, &tmVt, vt_full, 64, c * 64);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc128 = tc::umma_idesc_bf16_ex(128, 128, false, false), idesc64 = tc::umma_idesc_bf16_ex(64, 128, false, false);
      constexpr uint32_t idesc128o = tc::umma_idesc_bf16_ex(128, 128, false, true), idesc64o = tc::umma_idesc_bf16_ex(64, 128, false, true);
      // tile n: pair n / 3, row tile n % 3, TMEM buffer n & 1
      auto s_ready = [&](uint32_t n) {
        const uint32_t it = n / 3, t = n % 3, b = n & 1u;
        if (t == 0 && !(tc::mbar_test_wait(kt_full, it & 1u) && tc::mbar_test_wait(&a0_full[it & 1u], (it >> 1) & 1u))) return false;
        return n < 2 || tc::mbar_test_wait(&o_full[b], ((n - 2) >> 1) & 1u);   // the buffer's previous P has been consumed
      };
      auto issue_s = [&](uint32_t n) {
        const uint32_t it = n / 3, t = n % 3, b = n & 1u, s = it & 1u;
        trace_mark(a.trace, 1, n);   // S issued
        tc::tc_fence_after();
        const uint64_t da0 = tc::umma_desc_sw128(tc::smem_u32(a0s + s * kA0Stage) + t * 128 * 128, 16, 1024);
        const uint64_t db0 = tc::umma_desc_sw128(tc::smem_u32(kts), 16, 1024);
        const uint32_t d = tmem + b * 256;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t da = da0 + (uint64_t)(((ks >> 2) * kA2VHalf + (ks & 3) * 32) >> 4);
          const uint64_t db = db0 + (uint64_t)(((ks >> 2) * kKvHalf + (ks & 3) * 32) >> 4);
          tc::umma_bf16(d, da, db, t == 2 ? idesc64 : idesc128, ks != 0);
        }
        tc::umma_commit(&s_full[b]);
        if (t == 2) tc::umma_commit(kt_empty);   // last score product of the pair
      };
      auto o_ready = [&](uint32_t n) {
        const uint32_t it = n / 3, t = n % 3, b = n & 1u;
        if (t == 0 && !tc::mbar_test_wait(vt_full, it & 1u)) return false;
        if (n >= 2 && !tc::mbar_test_wait(&o_empty[b], ((n - 2) >> 1) & 1u)) return false;   // the buffer's previous O is in registers
        return tc::mbar_test_wait(&p_full[b], (n >> 1) & 1u);
      };
      auto issue_o = [&](uint32_t n) {
        const uint32_t t = n % 3, b = n & 1u;
        trace_mark(a.trace, 2, n);   // O issued
        tc::tc_fence_after();
        const uint64_t db0 = tc::umma_desc_sw128(tc::smem_u32(vts), kKvHalf, 1024);   // MN-major: K = (clip, head, key) rows
        const uint32_t d = tmem + b * 256 + 128, pa = tmem + b * 256;
        // K order of P: step ks = 4·(head pair) + 2·clip + (head in pair), 16 keys each — so that a thread's probabilities
        // and zeros land inside the score columns it read itself (see the softmax below); Vt rows are (clip, head, key)
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const int vrow = 64 * ((ks >> 1) & 1) + 32 * (ks >> 2) + 16 * (ks & 1);
          tc::umma_bf16_ts(d, pa + (uint32_t)(ks * 8), db0 + (uint64_t)((vrow * 128) >> 4), t == 2 ? idesc64o : idesc128o, ks != 0);
        }
        tc::umma_commit(&o_full[b]);
        if (t == 2) tc::umma_commit(vt_empty);
      };
      uint32_t ns = 0, no = 0;   // next score product / next output product to issue
      while (no < n_tiles) {
        bool progressed = false;
        if (no < ns && o_ready(no)) {
          issue_o(no++);
          progressed = true;
        }
        if (ns < n_tiles && ns <= no + 2 && s_ready(ns)) {
          issue_s(ns++);
          progressed = true;
        }
        if (!progressed) __nanosleep(20);
      }
    }
  } else {
    // ---- softmax + LayerNorm warps: group g handles the tiles with n & 1 == g in TMEM buffer g ----
    const int ww = warp - 2, grp = ww >> 3, hf = (ww >> 2) & 1, sp = warp & 3;
    const uint32_t tlane = (uint32_t)(32 * sp) << 16;
    const uint32_t tbuf = tmem + grp * 256 + tlane;
    const uint32_t pair_bar = 1 + grp * 4 + sp;      // named barrier of the two warps that share these 32 rows
    const float bsc = a.use_prior ? __ldg(a.bias_scale) : 0.f;
    const float bl = bsc * kL2E;
    const bool tracer = hf == 0 && sp == 2 && lane == 0;
    // row of tile n owned by this thread: R in the 320-row pair (clip 1 starts at 160), clip c, token j
    auto row_of = [&](uint32_t n, int& R, int& ci, int& c, bool& valid) {
      const uint32_t it = n / 3, t = n % 3;
      const int p = blockIdx.x + (int)it * (int)gridDim.x;
      R = 128 * (int)t + (t == 2 ? 16 * sp + (lane & 15) : 32 * sp + lane);
      ci = R >= BTAP ? 1 : 0;                                                 // warp-uniform: 160 = 5 x 32
      c = 2 * p + ci;
      valid = n < n_tiles && R - BTAP * ci < Ta && c < a.B && !(t == 2 && lane >= 16);
    };
    // a_query token score + prior part (fusion.py:174) and the v_key prior part of the row's clip
    auto row_consts = [&](uint32_t n, float& qa, float& pk) {
      int R, ci, c;
      bool valid;
      row_of(n, R, ci, c, valid);
      qa = pk = 0.f;
      if (a.use_prior && valid) {
        qa = __ldg(a.tok + ((int64_t)a.B + c) * BTAP + (R - BTAP * ci)) + __ldg(a.pc + (int64_t)c * 4 + 2);
        pk = __ldg(a.pc + (int64_t)c * 4 + 3);
      }
    };
    float qa_n, pk_n;
    row_consts(grp, qa_n, pk_n);
    // Per tile: phase A (scores -> probabilities) and phase B (output -> LayerNorm -> pooled partial).  The kernel is bound by
    // instruction issue (67% of the issue slots), not by the tensor-core round trips: running B one group iteration late, behind
    // A of the group's next tile, was measured and is slower (113 vs 107 us at B=4096).
    auto phase_a = [&](uint32_t n) {
      const uint32_t it = n / 3, s = it & 1u;
      int R, ci, c;
      bool valid;
      row_of(n, R, ci, c, valid);
      const float qa = qa_n, pk = pk_n;
      tc::mbar_wait(&s_full[grp], (n >> 1) & 1u);
      tc::mbar_wait(&a0_full[s], (it >> 1) & 1u);   // already complete; acquires the TMA-written key constants
      if (tracer) trace_mark(a.trace, 3, n);   // worker: scores ready
      tc::tc_fence_after();
      {
        uint32_t r0[32];                      // heads 2hf, 2hf+1 x 16 keys of this row against its own clip
        tc::tmem_ld_32x32(tbuf + (uint32_t)(64 * ci + 32 * hf), r0);
        const float* kc = kc_s + (s * 2 + ci) * 80;
        // (element-wise math on fp32 pairs: FADD2 / FMUL2 / FFMA2 halve the issue slots of this issue-bound kernel)
        uint64_t bia2[8];   // the prior bias is the same for every head (fusion.py:351-354): once per key
        {
          const uint64_t qp2 = tc::f2_pack(qa + pk, qa + pk), bl2 = tc::f2_pack(bl, bl);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            bia2[i] = 0ull;
            if (a.use_prior) {
              const float2 k2 = *reinterpret_cast<const float2*>(kc + 2 * i);
              float t0_, t1_;
              tc::f2_unpack(tc::f2_add(qp2, tc::f2_pack(k2.x, k2.y)), t0_, t1_);
              bia2[i] = tc::f2_mul(tc::f2_pack(tc::tanh_fast(t0_), tc::tanh_fast(t1_)), bl2);
            }
          }
          if (a.T < 16) {   // padded video tokens never receive probability mass
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float b0_, b1_;
              tc::f2_unpack(bia2[i], b0_, b1_);
              if (2 * i >= a.T) b0_ = -INFINITY;
              if (2 * i + 1 >= a.T) b1_ = -INFINITY;
              bia2[i] = tc::f2_pack(b0_, b1_);
            }
          }
        }
        tc::tmem_ld_wait();
        uint32_t pk2[16];
        const uint64_t l2e2 = tc::f2_pack(kL2E, kL2E);
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          uint64_t xq[8];
          float m = -INFINITY;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float2 c2 = *reinterpret_cast<const float2*>(kc + 16 + 32 * hf + 16 * h2 + 2 * i);
            xq[i] = tc::f2_fma(tc::f2_add(tc::f2_pack_u(r0[16 * h2 + 2 * i], r0[16 * h2 + 2 * i + 1]), tc::f2_pack(c2.x, c2.y)), l2e2, bia2[i]);
            float x0_, x1_;
            tc::f2_unpack(xq[i], x0_, x1_);
            m = fmaxf(m, fmaxf(x0_, x1_));
          }
          const uint64_t nm2 = tc::f2_pack(-m, -m);
          uint64_t l2 = 0ull;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float x0_, x1_;
            tc::f2_unpack(tc::f2_add(xq[i], nm2), x0_, x1_);
            xq[i] = tc::f2_pack(ex2_fast(x0_), ex2_fast(x1_));
            l2 = tc::f2_add(l2, xq[i]);
          }
          float l0, l1;
          tc::f2_unpack(l2, l0, l1);
          const float inv = __frcp_rn(l0 + l1);
          const uint64_t inv2 = tc::f2_pack(inv, inv);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float x0_, x1_;
            tc::f2_unpack(tc::f2_mul(xq[i], inv2), x0_, x1_);
            pk2[8 * h2 + i] = tc::pack_bf16(x0_, x1_);
          }
        }
        uint32_t z[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) z[i] = 0u;
        // P columns [32 hf, +32) = (clip 0 | clip 1) x (2 heads x 16 keys as bf16x2): for a clip-0 row exactly the score columns
        // this thread just read, for a clip-1 row clip-0 score columns nobody reads — the partner warp's scores stay intact
        tc::tmem_st_32x32_x16(tbuf + (uint32_t)(32 * hf + 16 * ci), pk2);
        tc::tmem_st_32x32_x16(tbuf + (uint32_t)(32 * hf + 16 * (ci ^ 1)), z);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        tc::mbar_arrive(&p_full[grp]);
        if (tracer) trace_mark(a.trace, 4, n);   // worker: P written
      }
      row_consts(n + 2, qa_n, pk_n);           // next tile's row constants: in flight during the LayerNorm below
    };
    auto phase_b = [&](uint32_t n) {
      // ---- a1 = LayerNorm(a0 + out [+ bo2, inside out]) (fusion.py:399) accumulated into the temporal mean (:402);
      //      this thread: columns [64 hf, 64 hf + 64) of its row ----
      const uint32_t it = n / 3, t = n % 3, s = it & 1u;
      int R, ci, c;
      bool valid;
      row_of(n, R, ci, c, valid);
      constexpr int HC = BE / 2, HP = HC / 2;
      uint64_t xp[HP];   // the half row as fp32 pairs: this phase is bound by instruction issue, so its element-wise math is packed (FADD2 / FFMA2)
      const uint8_t* arow = a0s + s * kA0Stage + hf * kA2VHalf + R * 128;
      tc::mbar_wait(&o_full[grp], (n >> 1) & 1u);
      if (tracer) trace_mark(a.trace, 5, n);   // worker: O ready
      tc::tc_fence_after();
      {
        uint32_t t0[32], t1[32];
        tc::tmem_ld_32x32(tbuf + 128u + (uint32_t)(64 * hf), t0);
        tc::tmem_ld_32x32(tbuf + 160u + (uint32_t)(64 * hf), t1);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          xp[i] = tc::f2_pack_u(t0[2 * i], t0[2 * i + 1]);
          xp[16 + i] = tc::f2_pack_u(t1[2 * i], t1[2 * i + 1]);
        }
      }
      uint64_t sum2 = 0ull, sq2 = 0ull;
#pragma unroll
      for (int k = 0; k < 8; ++k) {   // + this row of a0 (the residual, 16-byte chunk k of the swizzled half row)
        const uint4 res = *reinterpret_cast<const uint4*>(arow + (((k ^ R) & 7) << 4));
        const uint32_t w[4] = {res.x, res.y, res.z, res.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = 4 * k + e;   // (bo2 arrives inside O: folded into Vt by video_update)
          xp[j] = tc::f2_add(xp[j], tc::f2_pack_u(w[e] << 16, w[e] & 0xFFFF0000u));
          sum2 = tc::f2_add(sum2, xp[j]);
          sq2 = tc::f2_fma(xp[j], xp[j], sq2);
        }
      }
      // the two halves of the row meet: (sum, sum of squares) through the first two of the output columns this thread has
      // just loaded (its own half: the partner never reads them, and the buffer is not refilled before o_empty below)
      float su, sq;
      {
        float a0_, a1_;
        tc::f2_unpack(sum2, a0_, a1_);
        su = a0_ + a1_;
        tc::f2_unpack(sq2, a0_, a1_);
        sq = a0_ + a1_;
      }
      {
        const uint32_t tx = tbuf + 128u;
        tc::tmem_st_32x32_x1(tx + (uint32_t)(64 * hf), __float_as_uint(su));
        tc::tmem_st_32x32_x1(tx + (uint32_t)(64 * hf + 1), __float_as_uint(sq));
        tc::tmem_st_wait();
        tc::tc_fence_before();
        tc::named_bar_sync(pair_bar, 64);
        tc::tc_fence_after();
        const uint32_t psu = tc::tmem_ld_32x32_x1(tx + (uint32_t)(64 * (hf ^ 1)));
        const uint32_t psq = tc::tmem_ld_32x32_x1(tx + (uint32_t)(64 * (hf ^ 1) + 1));
        tc::tmem_ld_wait();
        su += __uint_as_float(psu);
        sq += __uint_as_float(psq);
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&o_empty[grp]);          // O and the exchange columns are done with: the next output product may land
      tc::mbar_arrive(&a0_empty[s]);           // and so is this row's residual
      if (tracer) trace_mark(a.trace, 6, n);   // worker: O + residual in registers
      const float mu = su * (1.f / BE);
      // padded / absent rows contribute nothing: scale and shift of zero (their x is finite: real or zero-filled a0 rows, and
      // the unused lanes of the M=64 tile hold the previous tile's finite values)
      const float rs = valid ? rsqrtf(fmaxf(sq * (1.f / BE) - mu * mu, 0.f) + 1e-5f) : 0.f;
      const float nmr = -mu * rs;
      const uint64_t rs2 = tc::f2_pack(rs, rs), nmr2 = tc::f2_pack(nmr, nmr);
#pragma unroll
      for (int i = 0; i < HP; ++i) xp[i] = tc::f2_fma(xp[i], rs2, nmr2);
      if (EMIT && valid) {   // the learned pooler wants the tokens: this thread's 64 columns of its row, 128 bytes
        uint4* dst = reinterpret_cast<uint4*>(a.xhat_out + ((int64_t)c * Ta + (R - BTAP * ci)) * BE + HC * hf);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          uint32_t w4[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float lo, hi;
            tc::f2_unpack(xp[4 * k + e], lo, hi);
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(lo, hi);
            w4[e] = *reinterpret_cast<const uint32_t*>(&h2);
          }
          dst[k] = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
      }
      // sum over the warp's 32 rows: butterfly, the live columns halve at every step (pairs stay pairs down to the last step)
#pragma unroll
      for (int off = 16, cnt = HP / 2; off >= 1; off >>= 1, cnt >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < cnt; ++i) {
          float s0_, s1_;
          tc::f2_unpack(up ? xp[i] : xp[i + cnt], s0_, s1_);
          const uint64_t keep = up ? xp[i + cnt] : xp[i];
          xp[i] = tc::f2_add(keep, tc::f2_pack(__shfl_xor_sync(0xffffffffu, s0_, off), __shfl_xor_sync(0xffffffffu, s1_, off)));
        }
      }
      // after the five steps lane L holds two columns: bit k of L chose the upper half at step k
      {
        int base = 0;
#pragma unroll
        for (int off = 16, cnt = HC / 2; off >= 1; off >>= 1, cnt >>= 1)
          if (lane & off) base += cnt;
        // slot of this (clip, tile, warp): even clip: tile 0 warps 0-3 -> 0-3, tile 1 warp 0 -> 4; odd clip: tile 1 warps 1-3 -> 0-2, tile 2 -> 3-6
        const int slot = ci == 0 ? (t == 0 ? sp : 4) : (t == 1 ? sp - 1 : 3 + sp);
        float x0_, x1_;
        tc::f2_unpack(xp[0], x0_, x1_);
        if (c < a.B)
          *reinterpret_cast<float2*>(a.poolp + ((int64_t)c * kPoolSlots + slot) * BE + HC * hf + base) = make_float2(x0_, x1_);
      }
      if (tracer) trace_mark(a.trace, 7, n);   // worker: tile done
    };
    for (uint32_t n = grp; n < n_tiles; n += 2) {
      phase_a(n);
      phase_b(n);
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
  alignas(64) CUtensorMap tmA0, tmQt, tmU;
  XF_TRY(make_tmap_bf16_2d(&tmA0, x.ws + x.w.a0, (int64_t)c.B * c.Ta, BE, BTAP, 64, BOP));
  XF_TRY(make_tmap_bf16_2d(&tmQt, x.ws + x.w.qt, (int64_t)c.B * 64, BE, 64, 64, BOP));
  XF_TRY(make_tmap_bf16_2d(&tmU, x.ws + x.w.u, (int64_t)c.B * 16, 4 * BE, 32, 64));
  XF_TRY(ensure_dyn_smem(v2a_tc_kernel, kV2ASmem));
  const int npairs = (c.B + 1) / 2;
  const int grid = npairs < num_sms() ? npairs : num_sms();
  XF_CUDA(launch_chain(v2a_tc_kernel, dim3(grid), dim3(kV2AThreads), kV2ASmem, x.st, tmA0, tmQt, tmU, a));
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



int launch_a2v_tc(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  A2VTcArgs a;
  a.tok = reinterpret_cast<const float*>(x.ws + x.w.tok);
  a.tkv = reinterpret_cast<const float*>(x.ws + x.w.tkv);
  a.pc = reinterpret_cast<const float*>(x.ws + x.w.pc);
  a.cv = reinterpret_cast<const float*>(x.ws + x.w.cv);
  a.bias_scale = P(XP_BIAS_SCALE);
  a.poolp = reinterpret_cast<float*>(x.ws + x.w.poolp);
  a.xhat_out = c.pool_mode == XATTN_POOL_ATTN ? reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.xhat) : nullptr;
  a.B = c.B; a.Ta = c.Ta; a.T = c.T; a.use_prior = c.use_prior;
  a.trace = nullptr;
  static unsigned long long* trace_buf = nullptr;
  static const bool tracing = getenv("XATTN_TC_TRACE") != nullptr;
  if (tracing) {
    if (trace_buf == nullptr) XF_CUDA(cudaMalloc(&trace_buf, 8 * 64 * 8));
    XF_CUDA(cudaMemsetAsync(trace_buf, 0, 8 * 64 * 8, x.st));
    a.trace = trace_buf;
  }
  alignas(64) CUtensorMap tmA0, tmKt, tmVt;
  XF_TRY(make_tmap_bf16_2d(&tmA0, x.ws + x.w.a0, (int64_t)c.B * c.Ta, BE, BTAP, 64, BOP));
  XF_TRY(make_tmap_bf16_2d(&tmKt, x.ws + x.w.kt, (int64_t)c.B * 64, BE, 64, 64, BOP));
  XF_TRY(make_tmap_bf16_2d(&tmVt, x.ws + x.w.vt, (int64_t)c.B * 64, BE, 64, 64, BOP));
  XF_TRY(ensure_dyn_smem(a2v_tc_kernel<false>, kA2VSmem));
  XF_TRY(ensure_dyn_smem(a2v_tc_kernel<true>, kA2VSmem));
  const int npairs = (c.B + 1) / 2;
  const int grid = npairs < num_sms() ? npairs : num_sms();
  if (a.xhat_out != nullptr)
    XF_CUDA(launch_chain(a2v_tc_kernel<true>, dim3(grid), dim3(kA2VThreads), kA2VSmem, x.st, tmA0, tmKt, tmVt, a));
  else
    XF_CUDA(launch_chain(a2v_tc_kernel<false>, dim3(grid), dim3(kA2VThreads), kA2VSmem, x.st, tmA0, tmKt, tmVt, a));
  count_launch();
  XF_LAUNCH_CHECK();
  if (tracing) {   // debugging aid: timeline of CTA 0 by row tile, microseconds from its first load
    static unsigned long long h[8 * 64];
    XF_CUDA(cudaStreamSynchronize(x.st));
    XF_CUDA(cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost));
    const char* names[8] = {"load_issue", "s_issue", "o_issue", "w_s_ready", "w_p_done", "w_o_ready", "w_o_in_regs", "w_done"};
    fprintf(stderr, "a2v_tc trace (us), B=%d\n%-8s", c.B, "tile");
    for (int e = 0; e < 8; ++e) fprintf(stderr, "%12s", names[e]);
    fprintf(stderr, "\n");
    for (int n = 0; n < 24; ++n) {
      fprintf(stderr, "%-8d", n);
      for (int e = 0; e < 8; ++e) fprintf(stderr, "%12.2f", h[e * 64 + n] ? (double)(h[e * 64 + n] - h[0]) * 1e-3 : -1.0);
      fprintf(stderr, "\n");
    }
  }
  return XATTN_OK;
}

}  // namespace xf
