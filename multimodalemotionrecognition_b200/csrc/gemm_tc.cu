// Token-flat projection GEMM on the 5th-gen tensor cores:   Y[M,N] = X[M,K] · W[N,K]^T + bias[N]
//   X, W bf16 (K contiguous), fp32 accumulation in TMEM, Y bf16 (row pitch in elements) or fp32.
// Used for the two input projections of the xattn head (reference src/models/fusion.py):
//   v_in_proj :372                         X = video_emb [B*T, 512],  W = v_in_proj.weight
//   a_in_proj ∘ audio_seq_proj :378 + :386 X = audio_emb [B*Ta, 768], W = W_in · W_seq (folded at pack time)
//
// Persistent, warp-specialised kernel: one CTA per SM loops over 128(M) x 128(N) output tiles.
//   warp 0    : TMA producer — cp.async.bulk.tensor (SWIZZLE_128B boxes of 64 x 128 bf16) into a 5-stage smem ring that
//               keeps running across tile boundaries (the HBM stream never drains between tiles)
//   warp 1    : one elected lane issues tcgen05.mma (M=128, N=128, K=16; 4 per stage) into one of FOUR 128-column TMEM
//               accumulators; tcgen05.commit frees the smem stage / publishes the accumulator
//   warps 2-5 : epilogue — tcgen05.ld (32 lanes x 32 columns), + bias, convert; the bf16 tile is staged in smem as the
//               exact global image (rows at the output pitch are contiguous) and leaves with ONE cp.async.bulk store,
//               while the MMA warp already fills the other accumulator with the next tile.
// EXTRAS (the a0 projection): the epilogue also emits, from data it already holds,
//   tok[0][m], tok[1][m] = w_ak·y_m, w_aq·y_m   the token parts of a_key_bias / a_query_bias (fusion.py:173-174), from the fp32 row
//   colsum[clip][slot][n]           column sums of a warp's 32 output rows that belong to `clip` (rows are clips of seg_len
//                                   tokens); slot = 32-row group - first group of the clip: a reader adds <= 6 partials in order
//                                   (the temporal mean that feeds prior_net, fusion.py:180)
// Two further epilogues run the video-side update of the xattn head as plain GEMMs over folded weights (bf16_pack.cu):
//   EPI_LNRES    v1 = LayerNorm(v0 + [U_0|U_1|U_2|U_3]·G^T + b)  (fusion.py:394-395 with the value and output projections
//                folded per head), + the temporal mean of v1 (:401) and c_hi = v1_i·wc_h + cc_h; rows are clips of 16 slots
//   EPI_HEADROWS Y[m, 128h' .. +128) lands in rows 16h + i of clip m/16 of one of two [clips][64][128] operands
//                (Kt | Vt of a2v_attn: N = 1024, column block h' = 4·which + h)
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int kStagesDefault = 5;
constexpr int kAccs = 4;                               // TMEM accumulators (4 x 128 columns = all of TMEM): the MMA warp can run 3 tiles ahead of the epilogue
constexpr int kStageBytes = (BM * BK + BN * BK) * 2;   // 32 KB
constexpr int gemm_threads(int epg) { return 64 + 128 * epg; }   // TMA warp, MMA warp, EPG groups of 4 epilogue warps
constexpr int kStageBox = BM * 128;                      // one staged box: 128 rows x 64 bf16 (STAGED: N == 128, dense rows)
constexpr int kBarBytes = 256;
constexpr int kVecBytes = 6144;   // bias [<= 1024] f32 | 2 KB of per-mode vectors

enum { EPI_F32 = 0, EPI_BF16 = 1, EPI_STAGED = 2, EPI_EXTRAS = 3, EPI_LNRES = 4, EPI_HEADROWS = 5 };
//   EPI_F32 / EPI_BF16: rows go straight to global.  EPI_STAGED: bf16, N == 128, dense rows: the tile is staged in smem and
//   leaves in global order.  EPI_EXTRAS: EPI_STAGED + token scores + column sums.

__host__ __device__ constexpr int gemm_stage_region(int epi) {
  return (epi == EPI_STAGED || epi == EPI_EXTRAS) ? 2 * kStageBox   // one bf16 tile
         : epi == EPI_HEADROWS                        ? 4 * kStageBox   // one bf16 tile per epilogue group
         : epi == EPI_LNRES                           ? 8192            // statistics / c_hi exchange between the column halves
                                                      : 0;
}
__host__ __device__ constexpr int gemm_smem_bytes(int epi, int stages) {
  return stages * kStageBytes + gemm_stage_region(epi) + kVecBytes + kBarBytes + 1024;
}

struct GemmExtras {
  const float* sc_w;   // [2][128] fp32 score weights
  float* tok;          // [2][M / seg_len][kTokPitch]
  float* colsum;       // [M / seg_len][kColsumSlots][128]
  int seg_len;
  // EPI_HEADROWS: column blocks 4..7 go to out2
  void* out2;
  // EPI_LNRES
  const float* res;            // v0 [clips][T][128] fp32
  const float *gamma, *beta;   // LayerNorm affine [128]
  const float* wc;             // [4][128] + [4]: c_hi = y·wc_h + wc[512 + h]
  float *cv, *pooled;          // [clips][64], [clips][256] (first 128 = temporal mean)
  int T, clips;
  int relu;                    // EPI_F32 / EPI_BF16: y = max(y, 0) after the bias
};

// One step of the transpose-reduce over lanes: x[0..N) (+)= the partner's other half; N values remain.
template <int N, int O, int LEN>
__device__ __forceinline__ void lnres_fold(float (&x)[LEN], int lane) {
  const bool up = (lane & O) != 0;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    const float send = up ? x[k] : x[k + N];
    const float keep = up ? x[k + N] : x[k];
    x[k] = keep + __shfl_xor_sync(0xffffffffu, send, O);
  }
}

// EPG groups of epilogue warps take the tiles of a CTA in turn (tile l of the CTA -> group l % EPG): for the short-K products the
// epilogue, not the operand stream, is what a tile costs.
template <int EPI, int kStages, int EPG>
__global__ void __launch_bounds__(gemm_threads(EPG), 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                  const __grid_constant__ CUtensorMap tmB,
                                                                  const __grid_constant__ CUtensorMap tmO,
                                                                  const __grid_constant__ CUtensorMap tmO2,
                                                                  const float* __restrict__ bias, void* __restrict__ out,
                                                                  int64_t out_pitch, int M, int N, int K, GemmExtras ex) {
  constexpr bool OUT_BF16 = EPI != EPI_F32 && EPI != EPI_LNRES;
  constexpr bool STAGED = EPI == EPI_STAGED || EPI == EPI_EXTRAS || EPI == EPI_HEADROWS;   // the tile is written to smem first
  constexpr bool EXTRAS = EPI == EPI_EXTRAS;
  constexpr bool HEADROWS = EPI == EPI_HEADROWS;
  constexpr bool LNRES = EPI == EPI_LNRES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* stage_out = smem + kStages * kStageBytes;   // STAGED: 2 boxes [128 rows][64 bf16], swizzled
  float* bias_s = reinterpret_cast<float*>(stage_out + gemm_stage_region(EPI));   // [N] (HEADROWS) or the current 128-column slice
  float2* scw_s = reinterpret_cast<float2*>(bias_s + 1024);   // EXTRAS: [128] (w_ak[n], w_aq[n])
  float4* wc_s = reinterpret_cast<float4*>(bias_s + 1024);    // LNRES: [128] (wc_0[k], wc_1[k], wc_2[k], wc_3[k]); gamma, beta, cc behind the bias
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(bias_s) + kVecBytes);
  uint64_t* empty = full + kStages;
  uint64_t* acc_full = empty + kStages;        // [kAccs]
  uint64_t* acc_empty = acc_full + kAccs;      // [kAccs]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kAccs);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KB = K / BK;
  const int tiles_n = N / BN;
  const int tiles_m = (M + BM - 1) / BM;
  const int num_tiles = tiles_m * tiles_n;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < kAccs; ++b) {
      tc::mbar_init(&acc_full[b], 1);
      tc::mbar_init(&acc_empty[b], LNRES ? 4 * EPG : 4);   // one arrival per epilogue warp that reads it
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, kAccs * BN);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  tc::pdl_wait();     // no-op unless launched as part of the forward chain (launch_gemm_tc, chain = true)
  tc::pdl_launch();

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;   // global k-iteration counter: the ring does not restart at tile boundaries
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          tc::mbar_wait(&empty[s], ph ^ 1u);
          tc::mbar_expect_tx(&full[s], kStageBytes);
          uint8_t* a = smem + s * kStageBytes;
          tc::tma_load_2d(a, &tmA, &full[s], kb * BK, m0);
          tc::tma_load_2d(a + BM * BK * 2, &tmB, &full[s], kb * BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = tc::umma_idesc_bf16(BM, BN);
      uint32_t it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
        const uint32_t ab = lt % kAccs, aph = (lt / kAccs) & 1u;
        tc::mbar_wait(&acc_empty[ab], aph ^ 1u);   // epilogue has drained this accumulator (passes on first use)
        tc::tc_fence_after();
        const uint32_t tmem_acc = tmem_base + ab * BN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          tc::mbar_wait(&full[s], ph);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + s * kStageBytes);
          const uint64_t da = tc::umma_desc_k_sw128(a_addr);
          const uint64_t db = tc::umma_desc_k_sw128(a_addr + BM * BK * 2);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)   // +32 bytes (= 2 in the >>4 address field) per K=16 step inside the swizzle atom
            tc::umma_bf16(tmem_acc, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          tc::umma_commit(&empty[s]);
        }
        tc::umma_commit(&acc_full[ab]);
      }
    }
  } else {
    // ---- epilogue warps: warp w may only touch TMEM lanes [32*(w%4), +32) ----
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int eg = (warp - 2) >> 2;            // epilogue group
    const int et = (threadIdx.x - 64) & 127;   // thread within the group
    if (EXTRAS) scw_s[et] = make_float2(__ldg(ex.sc_w + et), __ldg(ex.sc_w + BN + et));
    if (HEADROWS) {   // every column block's bias stays resident
      for (int i = et; i < N; i += 128) bias_s[i] = __ldg(bias + i);   // (every group writes the same values)
      tc::named_bar_sync(2 + eg, 128);
    }
    if (LNRES) {   // (both column halves write the same values; each waits for its own copy)
      bias_s[et] = __ldg(bias + et);
      bias_s[BN + et] = __ldg(ex.gamma + et);
      bias_s[2 * BN + et] = __ldg(ex.beta + et);
      if (et < 4) bias_s[3 * BN + et] = __ldg(ex.wc + 4 * BN + et);
      wc_s[et] = make_float4(__ldg(ex.wc + et), __ldg(ex.wc + BN + et), __ldg(ex.wc + 2 * BN + et), __ldg(ex.wc + 3 * BN + et));
      tc::named_bar_sync(2 + eg, 128);
    }
    if (HEADROWS) stage_out += eg * 2 * kStageBox;
    uint32_t lt = 0;
    int cur_n0 = -1;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
      const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
      const uint32_t ab = lt % kAccs, aph = (lt / kAccs) & 1u;
      if (!LNRES && EPG > 1 && (int)(lt % EPG) != eg) continue;   // (LNRES: the groups split the columns of every tile)
      if (!HEADROWS && !LNRES && n0 != cur_n0) {   // (re)load the bias slice; all epilogue threads pass the barrier below before reading it
        static_assert(EPG == 1 || HEADROWS || LNRES, "the bias slice is shared by the groups");
        tc::named_bar_sync(2, 128);
        bias_s[et] = __ldg(bias + n0 + et);
        if (EPI == EPI_F32 || EPI == EPI_BF16) bias_s[1024 + et] = 0.f;
        cur_n0 = n0;
        tc::named_bar_sync(2, 128);   // (also publishes scw_s / the LayerNorm vectors on the first pass)
      }
      const float* bs = bias_s + (HEADROWS ? n0 : 0);
      if (STAGED && !HEADROWS) __syncwarp();   // this warp's copy-out of the previous tile has finished reading its staged rows
      if (HEADROWS) {   // the group's previous tile has left the staging buffer
        if (et == 0) tc::bulk_store_wait_read();
        tc::named_bar_sync(2 + eg, 128);
      }
      if (LNRES) {
        // ---- v1 = LayerNorm(v0 + acc + b): a row (clip, slot) = (m / 16, m % 16) belongs to the two threads with this lane in
        //      the warps of its TMEM sub-partition, group eg owning columns [64·eg, +64); v0 has T <= 16 rows per clip.
        //      All global traffic is 32 bytes per lane (whole sectors); only the statistics and c_hi cross between the halves. ----
        static_assert(!LNRES || EPG == 2, "LNRES: two column halves");
        constexpr int HC = BN / 2;
        const int T = ex.T;
        const int mm = m0 + row, clip = mm >> 4, slot = mm & 15, c0 = HC * eg;
        const bool live = clip < ex.clips;
        float2* xs = reinterpret_cast<float2*>(stage_out) + (lt & 1u) * 2 * BM;          // [parity][half][row] (sum, sum of squares)
        float4* xc = reinterpret_cast<float4*>(stage_out + 4096) + (lt & 1u) * BM;       // [parity][row] c_hi partials of half 1
        float x[HC];
        {   // the residual half row, in flight while the accumulator is still being produced
          const float* rp = ex.res + ((int64_t)clip * T + slot) * BN + c0;
          if (live && slot < T) {
#pragma unroll
            for (int k = 0; k < HC; k += 8)
              tc::ldg256f(rp + k, x[k], x[k + 1], x[k + 2], x[k + 3], x[k + 4], x[k + 5], x[k + 6], x[k + 7]);
          } else {
#pragma unroll
            for (int k = 0; k < HC; ++k) x[k] = 0.f;
          }
        }
        tc::mbar_wait(&acc_full[ab], aph);
        tc::tc_fence_after();
#pragma unroll
        for (int c = 0; c < HC; c += 32) {
          uint32_t r[32];
          tc::tmem_ld_32x32(tmem_base + ab * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)(c0 + c), r);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) x[c + i] += __uint_as_float(r[i]) + bs[c0 + c + i];
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&acc_empty[ab]);
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
        for (int k = 0; k < HC; k += 2) {
          s0 += x[k]; s1 += x[k + 1];
          q0 = fmaf(x[k], x[k], q0); q1 = fmaf(x[k + 1], x[k + 1], q1);
        }
        xs[eg * BM + row] = make_float2(s0 + s1, q0 + q1);
        tc::named_bar_sync(4 + q, 64);   // the two warps of this sub-partition
        const float2 oth = xs[(eg ^ 1) * BM + row];
        const float mu = ((s0 + s1) + oth.x) * (1.f / BN);
        const float rstd = rsqrtf(fmaxf(((q0 + q1) + oth.y) * (1.f / BN) - mu * mu, 0.f) + 1e-5f);
        float cvh0 = 0.f, cvh1 = 0.f, cvh2 = 0.f, cvh3 = 0.f;
        __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(out) + (int64_t)mm * BN + c0;
#pragma unroll
        for (int k16 = 0; k16 < HC; k16 += 16) {
          uint32_t pk[8];
#pragma unroll
          for (int e = 0; e < 16; e += 4) {
            const int k = k16 + e;
            const float4 gm = *reinterpret_cast<const float4*>(bias_s + BN + c0 + k);
            const float4 bt = *reinterpret_cast<const float4*>(bias_s + 2 * BN + c0 + k);
            const float y0 = (x[k + 0] - mu) * rstd * gm.x + bt.x;
            const float y1 = (x[k + 1] - mu) * rstd * gm.y + bt.y;
            const float y2 = (x[k + 2] - mu) * rstd * gm.z + bt.z;
            const float y3 = (x[k + 3] - mu) * rstd * gm.w + bt.w;
            const float4 w0 = wc_s[c0 + k + 0], w1 = wc_s[c0 + k + 1], w2 = wc_s[c0 + k + 2], w3 = wc_s[c0 + k + 3];
            cvh0 = fmaf(y0, w0.x, fmaf(y1, w1.x, fmaf(y2, w2.x, fmaf(y3, w3.x, cvh0))));
            cvh1 = fmaf(y0, w0.y, fmaf(y1, w1.y, fmaf(y2, w2.y, fmaf(y3, w3.y, cvh1))));
            cvh2 = fmaf(y0, w0.z, fmaf(y1, w1.z, fmaf(y2, w2.z, fmaf(y3, w3.z, cvh2))));
            cvh3 = fmaf(y0, w0.w, fmaf(y1, w1.w, fmaf(y2, w2.w, fmaf(y3, w3.w, cvh3))));
            pk[e >> 1] = tc::pack_bf16(y0, y1);
            pk[(e >> 1) + 1] = tc::pack_bf16(y2, y3);
            // rows past T do not enter the temporal mean (fusion.py:401 averages the T real tokens)
            x[k + 0] = slot < T ? y0 : 0.f; x[k + 1] = slot < T ? y1 : 0.f;
            x[k + 2] = slot < T ? y2 : 0.f; x[k + 3] = slot < T ? y3 : 0.f;
          }
          if (live) tc::stg256(orow + k16, pk[0], pk[1], pk[2], pk[3], pk[4], pk[5], pk[6], pk[7]);
        }
        if (eg == 1) xc[row] = make_float4(cvh0, cvh1, cvh2, cvh3);
        tc::named_bar_sync(4 + q, 64);
        if (eg == 0 && live) {
          const float4 o = xc[row];
          float* cvp = ex.cv + (int64_t)clip * 64 + slot;
          cvp[0] = cvh0 + o.x + bias_s[3 * BN + 0]; cvp[16] = cvh1 + o.y + bias_s[3 * BN + 1];
          cvp[32] = cvh2 + o.z + bias_s[3 * BN + 2]; cvp[48] = cvh3 + o.w + bias_s[3 * BN + 3];
        }
        // temporal mean: the 16 rows of a clip are the 16 lanes of a half-warp.  Transpose-reduce: after the step with lane
        // offset o a lane keeps the half of the columns whose index bit matches its lane bit, so lane l ends with the
        // complete sums of columns 64·eg + 4·(l % 16) .. +4
        lnres_fold<32, 8>(x, lane);
        lnres_fold<16, 4>(x, lane);
        lnres_fold<8, 2>(x, lane);
        lnres_fold<4, 1>(x, lane);
        if (live) {
          const float invT = 1.f / (float)T;
          *reinterpret_cast<float4*>(ex.pooled + (int64_t)clip * 2 * BN + c0 + 4 * (lane & 15)) =
              make_float4(x[0] * invT, x[1] * invT, x[2] * invT, x[3] * invT);
        }
        continue;
      }
      tc::mbar_wait(&acc_full[ab], aph);
      tc::tc_fence_after();
      const int m = m0 + row;
      float sc0 = 0.f, sc1 = 0.f;
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem_base + ab * BN + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
        tc::tmem_ld_wait();
        if (EXTRAS) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float y = __uint_as_float(r[i]) + bs[c + i];
            const float2 w = scw_s[c + i];
            sc0 = fmaf(y, w.x, sc0);
            sc1 = fmaf(y, w.y, sc1);
          }
        }
        if (EPI == EPI_F32 || EPI == EPI_BF16) {   // bias (+ ReLU) in place; the stores below then add nothing
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float y = __uint_as_float(r[i]) + bs[c + i];
            if (ex.relu) y = fmaxf(y, 0.f);
            r[i] = __float_as_uint(y);
          }
        }
        const float* bq = (EPI == EPI_F32 || EPI == EPI_BF16) ? bias_s + 1024 : bs;   // (zeros: the bias is already in)
        if (OUT_BF16) {
          // direct rows: 32 columns = 64 bytes = two whole sectors per lane
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out) + (int64_t)m * out_pitch + n0 + c;
          if (STAGED || m < M) {
            uint32_t pk[16];
#pragma unroll
            for (int v = 0; v < 16; ++v)
              pk[v] = tc::pack_bf16(__uint_as_float(r[2 * v]) + bq[c + 2 * v], __uint_as_float(r[2 * v + 1]) + bq[c + 2 * v + 1]);
            if (STAGED) {   // 16-byte chunk q of the row -> box q / 8, slot (q % 8) ^ (row % 8)
#pragma unroll
              for (int v = 0; v < 4; ++v) {
                const int q8 = (c >> 3) + v;
                *reinterpret_cast<uint4*>(stage_out + (q8 >> 3) * kStageBox + row * 128 + (((q8 ^ row) & 7) << 4)) =
                    make_uint4(pk[4 * v], pk[4 * v + 1], pk[4 * v + 2], pk[4 * v + 3]);
              }
            } else if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
              tc::stg256(o, pk[0], pk[1], pk[2], pk[3], pk[4], pk[5], pk[6], pk[7]);
              tc::stg256(o + 16, pk[8], pk[9], pk[10], pk[11], pk[12], pk[13], pk[14], pk[15]);
            } else {
#pragma unroll
              for (int v = 0; v < 4; ++v)
                *reinterpret_cast<uint4*>(o + 8 * v) = make_uint4(pk[4 * v], pk[4 * v + 1], pk[4 * v + 2], pk[4 * v + 3]);
            }
          }
        } else if (m < M) {
          float* o = reinterpret_cast<float*>(out) + (int64_t)m * out_pitch + n0 + c;
          if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
#pragma unroll
            for (int v = 0; v < 4; ++v)
              tc::stg256f(o + 8 * v, __uint_as_float(r[8 * v + 0]) + bq[c + 8 * v + 0], __uint_as_float(r[8 * v + 1]) + bq[c + 8 * v + 1],
                          __uint_as_float(r[8 * v + 2]) + bq[c + 8 * v + 2], __uint_as_float(r[8 * v + 3]) + bq[c + 8 * v + 3],
                          __uint_as_float(r[8 * v + 4]) + bq[c + 8 * v + 4], __uint_as_float(r[8 * v + 5]) + bq[c + 8 * v + 5],
                          __uint_as_float(r[8 * v + 6]) + bq[c + 8 * v + 6], __uint_as_float(r[8 * v + 7]) + bq[c + 8 * v + 7]);
          } else {
#pragma unroll
            for (int v = 0; v < 8; ++v) {
              float4 f;
              f.x = __uint_as_float(r[4 * v + 0]) + bq[c + 4 * v + 0];
              f.y = __uint_as_float(r[4 * v + 1]) + bq[c + 4 * v + 1];
              f.z = __uint_as_float(r[4 * v + 2]) + bq[c + 4 * v + 2];
              f.w = __uint_as_float(r[4 * v + 3]) + bq[c + 4 * v + 3];
              *reinterpret_cast<float4*>(o + 4 * v) = f;
            }
          }
        }
      }
      // accumulator fully read: hand it back to the MMA warp
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&acc_empty[ab]);
      if (EXTRAS && m < M) {   // two planes [2][clips][kTokPitch]: 16-byte aligned rows per clip for the attention kernels
        const int clip = m / ex.seg_len, j = m - clip * ex.seg_len;
        ex.tok[(int64_t)clip * kTokPitch + j] = sc0;
        ex.tok[((int64_t)(M / ex.seg_len) + clip) * kTokPitch + j] = sc1;
      }
      if (HEADROWS) {
        // column block n0/128 = 4·which + h: rows (clip, i) of the tile are rows 16h + i of 8 consecutive clips of the [clips][64][128]
        // operand `which` (out | out2) — one 3D box per 64-column half, written by the copy engine as whole lines
        tc::fence_proxy_async();
        tc::named_bar_sync(2 + eg, 128);
        if (et == 0) {
          const int nt = n0 >> 7;
          const void* tm = nt < 4 ? reinterpret_cast<const void*>(&tmO) : reinterpret_cast<const void*>(&tmO2);
          tc::tma_store_3d(tm, stage_out, 0, 16 * (nt & 3), m0 >> 4);
          tc::tma_store_3d(tm, stage_out + kStageBox, 64, 16 * (nt & 3), m0 >> 4);
          tc::bulk_commit_group();
        }
      } else if (STAGED) {
        // Each warp owns 32 rows of the tile end to end: no CTA-level barrier, no asynchronous store to wait for.  The rows
        // were staged swizzled (conflict-free 16-byte writes); they leave in global order, 512 contiguous bytes per access.
        __syncwarp();
        const int wrow0 = q * 32, mrow0 = m0 + wrow0;
        {
          uint4* gout = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(out) + (int64_t)mrow0 * BN);
#pragma unroll 4
          for (int i = 0; i < 16; ++i) {
            const int lin = i * 32 + lane, rr = lin >> 4, q8 = lin & 15;   // row in the warp's block, 16-byte chunk of the row
            if (mrow0 + rr < M)
              gout[lin] = *reinterpret_cast<const uint4*>(stage_out + (q8 >> 3) * kStageBox + (wrow0 + rr) * 128 + (((q8 ^ rr) & 7) << 4));
          }
        }
        if (EXTRAS) {   // lane owns columns 4*lane .. 4*lane+3: sum them over the warp's rows of each clip present
          const int L = ex.seg_len;
          const int cb = 4 * lane, chunk = (cb & 63) >> 3;
          const uint8_t* colp = stage_out + (cb >> 6) * kStageBox + wrow0 * 128 + (cb & 7) * 2;
          const int nrows = min(32, M - mrow0);
          const int clip_last = (mrow0 + nrows - 1) / L;
          for (int clip = mrow0 / L; clip <= clip_last; ++clip) {
            const int r0 = max(clip * L, mrow0) - mrow0, r1 = min((clip + 1) * L, mrow0 + nrows) - mrow0;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            for (int rr = r0; rr < r1; ++rr) {
              const uint2 v = *reinterpret_cast<const uint2*>(colp + rr * 128 + (((chunk ^ rr) & 7) << 4));
              a0 += __uint_as_float(v.x << 16); a1 += __uint_as_float(v.x & 0xFFFF0000u);
              a2 += __uint_as_float(v.y << 16); a3 += __uint_as_float(v.y & 0xFFFF0000u);
            }
            const int slot = mrow0 / 32 - (clip * L) / 32;   // 32-row group of the output relative to the clip's first group
            *reinterpret_cast<float4*>(ex.colsum + ((int64_t)clip * kColsumSlots + slot) * BN + cb) = make_float4(a0, a1, a2, a3);
          }
        }
      }
    }
  }
  if (HEADROWS && warp >= 2 && ((threadIdx.x - 64) & 127) == 0) tc::bulk_store_wait_all();
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, kAccs * BN);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

template <int EPI, int STAGES = kStagesDefault, int EPG = 1>
int launch_gemm_tc(cudaStream_t st, const CUtensorMap& tmA, const CUtensorMap& tmB, const float* bias, void* out, int64_t pitch,
                   int M, int N, int K, const GemmExtras& ex, const CUtensorMap* tmO = nullptr, const CUtensorMap* tmO2 = nullptr,
                   bool chain = false) {
  if (tmO == nullptr) tmO = &tmA;   // (only EPI_HEADROWS stores through tensor maps)
  if (tmO2 == nullptr) tmO2 = &tmA;
  constexpr int smem = gemm_smem_bytes(EPI, STAGES);
  static_assert(smem <= 227 * 1024, "projection GEMM: shared memory");
  XF_TRY(ensure_dyn_smem(gemm_tc_kernel<EPI, STAGES, EPG>, smem));
  const int tiles = ((M + BM - 1) / BM) * (N / BN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  if (chain)
    XF_CUDA(launch_chain(gemm_tc_kernel<EPI, STAGES, EPG>, dim3(grid), dim3(gemm_threads(EPG)), smem, st, tmA, tmB, *tmO, *tmO2, bias, out,
                         pitch, M, N, K, ex));
  else
    gemm_tc_kernel<EPI, STAGES, EPG><<<grid, gemm_threads(EPG), smem, st>>>(tmA, tmB, *tmO, *tmO2, bias, out, pitch, M, N, K, ex);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace

int make_tmap_bf16_2d(void* tmap_out, const void* base, int64_t rows, int64_t cols, int box_rows, int box_cols, int64_t pitch) {
  if (pitch <= 0) pitch = cols;
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)pitch * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(reinterpret_cast<CUtensorMap*>(tmap_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                  gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld)", (int)r,
                                 (long long)rows, (long long)cols);
  return XATTN_OK;
}

// dense [d2][d1][d0] bf16 tensor (d0 contiguous), boxes b2 x b1 x b0 with the 128-byte swizzle
int make_tmap_bf16_3d(void* tmap_out, const void* base, int64_t d0, int64_t d1, int64_t d2, int b0, int b1, int b2) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t gstride[2] = {(cuuint64_t)d0 * 2, (cuuint64_t)d0 * d1 * 2};
  cuuint32_t box[3] = {(cuuint32_t)b0, (cuuint32_t)b1, (cuuint32_t)b2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(reinterpret_cast<CUtensorMap*>(tmap_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), gdim,
                  gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled (3D) failed with CUresult %d", (int)r);
  return XATTN_OK;
}

int gemm_tc_bf16(cudaStream_t st, const void* X, const void* W, const float* bias, void* out, bool out_bf16, int64_t out_pitch,
                 int M, int N, int K, const float* sc_w, float* tok, float* colsum, int seg_len, bool relu) {
  if (M <= 0) return XATTN_OK;
  const bool extras = sc_w != nullptr;
  const bool staged0 = out_bf16 && N == BN && out_pitch == BN;
  if (extras && !(staged0 && !relu && seg_len > 0 && M % seg_len == 0 &&
                  (seg_len + 30) / 32 + 1 <= kColsumSlots && seg_len <= kTokPitch))
    XF_FAIL(XATTN_ERR_SHAPE, "projection GEMM extras need dense bf16 output with N == 128 and rows in clips of <= 160 tokens");
  GemmExtras ex{};
  ex.sc_w = sc_w; ex.tok = tok; ex.colsum = colsum; ex.seg_len = seg_len; ex.relu = relu ? 1 : 0;
  const bool staged = !relu && staged0;
  if (N % BN != 0 || K % BK != 0)
    XF_FAIL(XATTN_ERR_SHAPE, "tcgen05 projection GEMM needs N %% 128 == 0 and K %% 64 == 0 (N=%d K=%d)", N, K);
  if (out_pitch < N || (out_bf16 && out_pitch % 8 != 0) || (!out_bf16 && out_pitch % 4 != 0))
    XF_FAIL(XATTN_ERR_SHAPE, "output pitch %lld must be >= N and keep rows 16-byte aligned", (long long)out_pitch);
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap_bf16_2d(&tmA, X, M, K, BM, BK));
  XF_TRY(make_tmap_bf16_2d(&tmB, W, N, K, BN, BK));
  if (extras) return launch_gemm_tc<EPI_EXTRAS>(st, tmA, tmB, bias, out, out_pitch, M, N, K, ex);
  if (staged) return launch_gemm_tc<EPI_STAGED>(st, tmA, tmB, bias, out, out_pitch, M, N, K, ex);
  if (out_bf16) return launch_gemm_tc<EPI_BF16>(st, tmA, tmB, bias, out, out_pitch, M, N, K, ex);
  return launch_gemm_tc<EPI_F32>(st, tmA, tmB, bias, out, out_pitch, M, N, K, ex);
}

int gemm_tc_video_mix(cudaStream_t st, const void* U, const void* G, const float* bias, const float* v0, const float* gamma,
                      const float* beta, const float* wc, void* v1, float* cv, float* pooled, int clips, int T) {
  if (clips <= 0) return XATTN_OK;
  const int M = clips * 16;
  GemmExtras ex{};
  ex.res = v0; ex.gamma = gamma; ex.beta = beta; ex.wc = wc; ex.cv = cv; ex.pooled = pooled; ex.T = T; ex.clips = clips;
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap_bf16_2d(&tmA, U, M, 4 * BN, BM, BK));
  XF_TRY(make_tmap_bf16_2d(&tmB, G, BN, 4 * BN, BN, BK));
  return launch_gemm_tc<EPI_LNRES, 6, 2>(st, tmA, tmB, bias, v1, BN, M, BN, 4 * BN, ex, nullptr, nullptr, true);
}

int gemm_tc_video_kv(cudaStream_t st, const void* v1, const void* F, const float* bias, void* kt, void* vt, int clips) {
  if (clips <= 0) return XATTN_OK;
  const int M = clips * 16;
  GemmExtras ex{};
  ex.out2 = vt;
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap_bf16_2d(&tmA, v1, M, BN, BM, BK));
  XF_TRY(make_tmap_bf16_2d(&tmB, F, 8 * BN, BN, BN, BK));
  alignas(64) CUtensorMap tmK, tmV;   // [clips][64][128], boxes of 8 clips x 16 rows x 64 columns
  XF_TRY(make_tmap_bf16_3d(&tmK, kt, BN, 64, clips, 64, 16, 8));
  XF_TRY(make_tmap_bf16_3d(&tmV, vt, BN, 64, clips, 64, 16, 8));
  return launch_gemm_tc<EPI_HEADROWS, 4, 2>(st, tmA, tmB, bias, kt, BN, M, 8 * BN, BN, ex, &tmK, &tmV, true);
}

}  // namespace xf
