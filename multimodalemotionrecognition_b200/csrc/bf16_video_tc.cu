// tcgen05 version of the video-side preparation of the bf16 path (reference src/models/fusion.py:372, :394, :171-181):
// per clip, from v0 = v_in_proj(video_emb) [T,128] (fp32, written by the projection GEMM)
//   Q~   = (v0·Wq1^T + bq1) / sqrt(dh)                    the query projection of v2a_attn (in_proj rows 0..E)
//   Qt_h = Q~_h · Wk1_h                                    [16,128] per head: the key projection folded onto the query side,
//                                                          so that  Q_h·K_h^T = Qt_h · a0^T  (bf16_layout.cuh)
//   mean_t v0, w_vq·v0_i, w_vk·v0_i                        inputs of the emotion prior / its bias heads
// Eight clips x 16 token slots form one 128-row tile, so the two products are M=128 tcgen05 MMAs with the 128x128 weights
// resident in shared memory:
//   load   : each warp reads its 32 rows with fully coalesced 512-byte accesses (a lane owns 4 columns of every row, which
//            makes mean_t v0 a private sum), converts to bf16 and writes the swizzled K-major A tile
//            the two token scores of its 16 rows are fp32 partial dot products, summed over the lanes by a 31-shuffle
//            transpose-reduce that leaves lane j with score j (tqv | tkv)
//   MMA 1  : D[0,128) = A·Wq1^T (K-major B)
//   pass 1 : thread = row: (D + bq1)/sqrt(dh) -> bf16 -> overwrites the A tile
//   MMA 2  : per head h, D_h = A[:, 32h:32h+32]·Wk1[32h:32h+32, :]  (K = 32; B = Wk1 read as an MN-major operand: rows = K)
//            heads 0,1 then 2,3 through two 128-column accumulators
//   pass 2 : thread = row: bf16, 64-byte strips staged per warp and copied out as full half-lines of Qt[clip][16h+i][:]
// warps 0-3: rows (TMEM sub-partition = warp), warp 4: barrier init / TMEM alloc / weight TMA / MMA issue.
// Two CTAs per SM (108 KB, 256 TMEM columns each) cover each other's serial chain.
#include <cuda.h>
#include <stdlib.h>

#include "bf16_layout.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kVBox = 128 * 128;                 // 16 KB: 128 rows x 64 bf16, 128-byte swizzle
constexpr int kVOffWq = 0;                       // Wq1: 2 boxes (input features 0-63 | 64-127), rows = output feature
constexpr int kVOffWk = 2 * kVBox;               // Wk1: same image; read as MN-major (rows = K = d)
constexpr int kVOffA = 4 * kVBox;                // A tile: v0 (bf16), then Q~
constexpr int kVOffStage = 6 * kVBox;            // 4 warps x [32 rows][64 B]
constexpr int kVOffVec = kVOffStage + 4 * 32 * 64;   // bq1 [128] f32
constexpr int kVOffBar = kVOffVec + BE * 4;
constexpr int kVSmem = kVOffBar + 8 * 8 + 8;
static_assert(kVSmem <= 112 * 1024, "video_pre_tc: two CTAs per SM");
constexpr int kVThreads = 160;

struct VideoPreTcArgs {
  const float* v0;                 // [B,T,E]
  const float *bq1, *wvq, *wvk;    // wvq / wvk: first E entries of v_query_bias / v_key_bias weights (or nullptr)
  __nv_bfloat16* qt;               // [B,64,BOP]
  float *meanv, *tqv, *tkv;        // [B,E], [B,16], [B,16]
  int B, T;
  float scale;
};

__global__ void __launch_bounds__(kVThreads, 2) video_pre_tc_kernel(const __grid_constant__ CUtensorMap tmWq,
                                                                     const __grid_constant__ CUtensorMap tmWk, VideoPreTcArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* a_tile = smem + kVOffA;
  float* bq_s = reinterpret_cast<float*>(smem + kVOffVec);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kVOffBar);
  uint64_t *w_full = bars, *a_ready = bars + 1, *q_full = bars + 2, *q_ready = bars + 3, *qt_full = bars + 4, *d_free = bars + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = a.T;
  const int ntiles = (a.B + 7) / 8;

  if (threadIdx.x == 128) {
    if ((tc::smem_u32(smem) & 1023u) != 0) __trap();
    tc::tma_prefetch_desc(&tmWq);
    tc::tma_prefetch_desc(&tmWk);
    tc::mbar_init(w_full, 1);
    tc::mbar_init(a_ready, 128);
    tc::mbar_init(q_full, 1);
    tc::mbar_init(q_ready, 128);
    tc::mbar_init(&qt_full[0], 1);
    tc::mbar_init(&qt_full[1], 1);
    tc::mbar_init(d_free, 128);
    tc::fence_barrier_init();
    tc::mbar_expect_tx(w_full, 4 * kVBox);
    tc::tma_load_2d(smem + kVOffWq, &tmWq, w_full, 0, 0);
    tc::tma_load_2d(smem + kVOffWq + kVBox, &tmWq, w_full, 64, 0);
    tc::tma_load_2d(smem + kVOffWk, &tmWk, w_full, 0, 0);
    tc::tma_load_2d(smem + kVOffWk + kVBox, &tmWk, w_full, 64, 0);
  }
  if (warp == 4) tc::tmem_alloc(tmem_slot, 256);
  if (threadIdx.x < BE) bq_s[threadIdx.x] = __ldg(a.bq1 + threadIdx.x);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 4) {
    if (lane == 0) {
      constexpr uint32_t idesc_q = tc::umma_idesc_bf16_ex(128, 128, false, false);
      constexpr uint32_t idesc_t = tc::umma_idesc_bf16_ex(128, 128, false, true);
      const uint64_t dA = tc::umma_desc_sw128(tc::smem_u32(a_tile), 16, 1024);
      const uint64_t dWq = tc::umma_desc_sw128(tc::smem_u32(smem + kVOffWq), 16, 1024);
      const uint64_t dWk = tc::umma_desc_sw128(tc::smem_u32(smem + kVOffWk), kVBox, 1024);   // MN-major: LBO between the 64-column halves
      tc::mbar_wait(w_full, 0);
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const uint32_t ph = it & 1u;
        tc::mbar_wait(a_ready, ph);
        tc::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t off = (uint64_t)(((ks >> 2) * kVBox + (ks & 3) * 32) >> 4);
          tc::umma_bf16(tmem, dA + off, dWq + off, idesc_q, ks != 0);
        }
        tc::umma_commit(q_full);
        tc::mbar_wait(q_ready, ph);   // Q~ is in the A tile (and the first accumulators have been read)
        tc::tc_fence_after();
        for (int hp = 0; hp < 2; ++hp) {
          if (hp == 1) {
            tc::mbar_wait(d_free, ph);   // heads 0,1 have left the accumulators
            tc::tc_fence_after();
          }
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int h = 2 * hp + hh;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const uint64_t offa = (uint64_t)(((h >> 1) * kVBox + (h & 1) * 64 + kk * 32) >> 4);   // A columns 32h + 16kk ..
              const uint64_t offb = (uint64_t)(((32 * h + 16 * kk) * 128) >> 4);                  // Wk1 rows 32h + 16kk ..
              tc::umma_bf16(tmem + hh * 128, dA + offa, dWk + offb, idesc_t, kk != 0);
            }
          }
          tc::umma_commit(&qt_full[hp]);
        }
      }
    }
  } else {
    // ---- row warps: warp w owns rows 32w .. 32w+31 of the tile = clips 2w, 2w+1 of its eight ----
    const uint32_t tlane = (uint32_t)(32 * warp) << 16;
    const int row = 32 * warp + lane;
    uint8_t* stage = smem + kVOffStage + warp * (32 * 64);
    const float invT = 1.f / (float)T;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 wq4 = a.wvq != nullptr ? __ldg(reinterpret_cast<const float4*>(a.wvq) + lane) : zero4;
    const float4 wk4 = a.wvk != nullptr ? __ldg(reinterpret_cast<const float4*>(a.wvk) + lane) : zero4;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const uint32_t ph = it & 1u;
      // ---- load: 32 rows x 512 B, one row per access; lane owns columns 4*lane .. +3 ----
      {
        const int box = lane >> 4, chunk = (lane & 15) >> 1, sub = (lane & 1) * 8;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int c = tile * 8 + 2 * warp + half;
          float4 x[16];
#pragma unroll
          for (int i = 0; i < 16; ++i)
            x[i] = (c < a.B && i < T) ? __ldg(reinterpret_cast<const float4*>(a.v0 + ((int64_t)c * T + i) * BE) + lane)
                                      : zero4;
          float4 s = zero4;
          float sc[32];   // partial token scores: [0,16) = w_vq·v0_i, [16,32) = w_vk·v0_i over this lane's 4 columns
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int r = 32 * warp + 16 * half + i;
            uint2 pk;
            pk.x = tc::pack_bf16(x[i].x, x[i].y);
            pk.y = tc::pack_bf16(x[i].z, x[i].w);
            *reinterpret_cast<uint2*>(a_tile + box * kVBox + r * 128 + (((chunk ^ r) & 7) << 4) + sub) = pk;
            s.x += x[i].x; s.y += x[i].y; s.z += x[i].z; s.w += x[i].w;
            sc[i] = fmaf(wq4.x, x[i].x, fmaf(wq4.y, x[i].y, fmaf(wq4.z, x[i].z, wq4.w * x[i].w)));
            sc[16 + i] = fmaf(wk4.x, x[i].x, fmaf(wk4.y, x[i].y, fmaf(wk4.z, x[i].z, wk4.w * x[i].w)));
          }
          // transpose-reduce over the lanes: after the step with offset o, a lane keeps the half of the values whose
          // index bit o matches its own lane bit; lane j ends with the full sum of value j
#pragma unroll
          for (int o = 16; o >= 1; o >>= 1) {
            const bool up = (lane & o) != 0;
#pragma unroll
            for (int k = 0; k < o; ++k) {
              const float send = up ? sc[k] : sc[k + o];
              const float keep = up ? sc[k + o] : sc[k];
              sc[k] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
          }
          if (c < a.B) {
            reinterpret_cast<float4*>(a.meanv + (int64_t)c * BE)[lane] = make_float4(s.x * invT, s.y * invT, s.z * invT, s.w * invT);
            (lane < 16 ? a.tqv + (int64_t)c * 16 + lane : a.tkv + (int64_t)c * 16 + (lane - 16))[0] = sc[0];
          }
        }
      }
      tc::fence_proxy_async();
      tc::tc_fence_before();   // this thread's reads of the previous tile's accumulators are complete
      tc::mbar_arrive(a_ready);
      // ---- pass 1: Q~ = (v0 Wq1^T + bq1) / sqrt(dh) -> bf16 -> A tile ----
      tc::mbar_wait(q_full, ph);
      tc::tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < BE; c0 += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem + tlane + (uint32_t)c0, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          uint4 pk;
          const int cb = c0 + 8 * v;
          pk.x = tc::pack_bf16((__uint_as_float(r[8 * v + 0]) + bq_s[cb + 0]) * a.scale, (__uint_as_float(r[8 * v + 1]) + bq_s[cb + 1]) * a.scale);
          pk.y = tc::pack_bf16((__uint_as_float(r[8 * v + 2]) + bq_s[cb + 2]) * a.scale, (__uint_as_float(r[8 * v + 3]) + bq_s[cb + 3]) * a.scale);
          pk.z = tc::pack_bf16((__uint_as_float(r[8 * v + 4]) + bq_s[cb + 4]) * a.scale, (__uint_as_float(r[8 * v + 5]) + bq_s[cb + 5]) * a.scale);
          pk.w = tc::pack_bf16((__uint_as_float(r[8 * v + 6]) + bq_s[cb + 6]) * a.scale, (__uint_as_float(r[8 * v + 7]) + bq_s[cb + 7]) * a.scale);
          const int q8 = cb >> 3;
          *reinterpret_cast<uint4*>(a_tile + (q8 >> 3) * kVBox + row * 128 + (((q8 ^ row) & 7) << 4)) = pk;
        }
      }
      tc::fence_proxy_async();
      tc::tc_fence_before();
      tc::mbar_arrive(q_ready);
      // ---- pass 2: Qt_h rows -> bf16 -> Qt[clip][16h + i][:], 64-byte strips staged per warp, copied out as half lines ----
#pragma unroll 1
      for (int hp = 0; hp < 2; ++hp) {
        tc::mbar_wait(&qt_full[hp], ph);
        tc::tc_fence_after();
#pragma unroll 1
        for (int hh = 0; hh < 2; ++hh) {
          const int h = 2 * hp + hh;
#pragma unroll 1
          for (int c0 = 0; c0 < BE; c0 += 32) {
            uint32_t r[32];
            tc::tmem_ld_32x32(tmem + tlane + (uint32_t)(hh * 128 + c0), r);
            tc::tmem_ld_wait();
            __syncwarp();   // the previous strip has been copied out
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              uint4 pk;
              pk.x = tc::pack_bf16(__uint_as_float(r[8 * v + 0]), __uint_as_float(r[8 * v + 1]));
              pk.y = tc::pack_bf16(__uint_as_float(r[8 * v + 2]), __uint_as_float(r[8 * v + 3]));
              pk.z = tc::pack_bf16(__uint_as_float(r[8 * v + 4]), __uint_as_float(r[8 * v + 5]));
              pk.w = tc::pack_bf16(__uint_as_float(r[8 * v + 6]), __uint_as_float(r[8 * v + 7]));
              *reinterpret_cast<uint4*>(stage + lane * 64 + ((v ^ (lane >> 1)) & 3) * 16) = pk;   // 4 chunks per row, xor-swizzled
            }
            __syncwarp();
            // copy-out: 32 rows x 64 B = 128 x 16 B; lane handles chunks lane, lane+32, ..: row = idx / 4
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int idx = k * 32 + lane, rr = idx >> 2, v = idx & 3;
              const int cc = tile * 8 + 2 * warp + (rr >> 4);
              if (cc < a.B) {
                const uint4 val = *reinterpret_cast<const uint4*>(stage + rr * 64 + ((v ^ (rr >> 1)) & 3) * 16);
                *reinterpret_cast<uint4*>(a.qt + ((int64_t)cc * 64 + h * 16 + (rr & 15)) * BOP + c0 + 8 * v) = val;
              }
            }
          }
        }
        if (hp == 0) {
          tc::tc_fence_before();
          tc::mbar_arrive(d_free);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 4) tc::tmem_dealloc(tmem, 256);
}

}  // namespace

int launch_video_pre(const Bf16Ctx& x) {
  const xattn_config& c = *x.cfg;
  auto P = [&](int id) { return x.params + x.pl.off[id]; };
  VideoPreTcArgs a;
  a.v0 = reinterpret_cast<const float*>(x.ws + x.w.v0);
  a.bq1 = P(XP_V2A_IN_B);
  a.wvq = c.use_prior ? P(XP_VQ_W) : nullptr;
  a.wvk = c.use_prior ? P(XP_VK_W) : nullptr;
  a.qt = reinterpret_cast<__nv_bfloat16*>(x.ws + x.w.qt);
  a.meanv = reinterpret_cast<float*>(x.ws + x.w.meanv);
  a.tqv = reinterpret_cast<float*>(x.ws + x.w.tqv);
  a.tkv = reinterpret_cast<float*>(x.ws + x.w.tkv);
  a.B = c.B; a.T = c.T;
  a.scale = 1.0f / sqrtf((float)BDH);
  alignas(64) CUtensorMap tmWq, tmWk;
  XF_TRY(make_tmap_bf16_2d(&tmWq, x.packed + x.pk.Wq1, BE, BE, 128, 64, BEP));
  XF_TRY(make_tmap_bf16_2d(&tmWk, x.packed + x.pk.Wk1, BE, BE, 128, 64, BEP));
  XF_TRY(ensure_dyn_smem(video_pre_tc_kernel, kVSmem));
  const int ntiles = (c.B + 7) / 8;
  const int cap = 2 * num_sms();
  video_pre_tc_kernel<<<ntiles < cap ? ntiles : cap, kVThreads, kVSmem, x.st>>>(tmWq, tmWk, a);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
