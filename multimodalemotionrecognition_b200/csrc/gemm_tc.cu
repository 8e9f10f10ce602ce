// Token-flat projection GEMM on the 5th-gen tensor cores:   Y[M,N] = X[M,K] · W[N,K]^T + bias[N]
//   X, W bf16 (K contiguous), fp32 accumulation in TMEM, Y bf16 (row pitch in elements) or fp32.
// Used for the two input projections of the xattn head (reference src/models/fusion.py):
//   v_in_proj :372                         X = video_emb [B*T, 512],  W = v_in_proj.weight
//   a_in_proj ∘ audio_seq_proj :378 + :386 X = audio_emb [B*Ta, 768], W = W_in · W_seq (folded at pack time)
//
// Kernel shape: one 128(M) x 128(N) output tile per CTA, K in 64-element (128-byte) chunks.
//   warp 0    : TMA producer — cp.async.bulk.tensor (SWIZZLE_128B boxes of 64 x 128) into a kStages-deep smem ring
//   warp 1    : allocates 128 TMEM columns, one elected lane issues tcgen05.mma (M=128, N=128, K=16) x4 per stage and
//               tcgen05.commit's the stage back to the producer; a final commit signals the epilogue
//   warps 2-5 : epilogue — tcgen05.ld (32 lanes x 32 columns), + bias, convert, 16-byte global stores
// 96 KB of smem per CTA => two CTAs per SM, so one CTA's epilogue overlaps the other's main loop.
#include <cuda.h>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int kStages = 3;
constexpr int kStageBytes = (BM * BK + BN * BK) * 2;   // 32 KB
constexpr int kGemmThreads = 192;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;

template <bool OUT_BF16>
__global__ void __launch_bounds__(kGemmThreads) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                               const __grid_constant__ CUtensorMap tmB,
                                                               const float* __restrict__ bias, void* __restrict__ out,
                                                               int64_t out_pitch, int M, int N, int K) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* acc_full = empty + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int KB = K / BK;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmA);
    tc::tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(&full[s], 1);
      tc::mbar_init(&empty[s], 1);
    }
    tc::mbar_init(acc_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(tmem_slot, BN);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < KB; ++kb) {
        const int s = kb % kStages;
        const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
        tc::mbar_wait(&empty[s], ph ^ 1u);
        tc::mbar_expect_tx(&full[s], kStageBytes);
        uint8_t* a = smem + s * kStageBytes;
        tc::tma_load_2d(a, &tmA, &full[s], kb * BK, m0);
        tc::tma_load_2d(a + BM * BK * 2, &tmB, &full[s], kb * BK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = tc::umma_idesc_bf16(BM, BN);
      for (int kb = 0; kb < KB; ++kb) {
        const int s = kb % kStages;
        const uint32_t ph = (uint32_t)(kb / kStages) & 1u;
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
      tc::umma_commit(acc_full);
    }
  } else {
    // epilogue: warp w may only touch TMEM lanes [32*(w%4), +32)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int m = m0 + row;
    tc::mbar_wait(acc_full, 0);
    tc::tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < BN; c += 32) {
      uint32_t r[32];
      tc::tmem_ld_32x32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
      tc::tmem_ld_wait();
      if (m < M) {
        const float* bp = bias + n0 + c;
        if (OUT_BF16) {
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out) + (int64_t)m * out_pitch + n0 + c;
#pragma unroll
          for (int v = 0; v < 4; ++v) {
            uint4 pk;
            pk.x = tc::pack_bf16(__uint_as_float(r[8 * v + 0]) + __ldg(bp + 8 * v + 0), __uint_as_float(r[8 * v + 1]) + __ldg(bp + 8 * v + 1));
            pk.y = tc::pack_bf16(__uint_as_float(r[8 * v + 2]) + __ldg(bp + 8 * v + 2), __uint_as_float(r[8 * v + 3]) + __ldg(bp + 8 * v + 3));
            pk.z = tc::pack_bf16(__uint_as_float(r[8 * v + 4]) + __ldg(bp + 8 * v + 4), __uint_as_float(r[8 * v + 5]) + __ldg(bp + 8 * v + 5));
            pk.w = tc::pack_bf16(__uint_as_float(r[8 * v + 6]) + __ldg(bp + 8 * v + 6), __uint_as_float(r[8 * v + 7]) + __ldg(bp + 8 * v + 7));
            *reinterpret_cast<uint4*>(o + 8 * v) = pk;
          }
        } else {
          float* o = reinterpret_cast<float*>(out) + (int64_t)m * out_pitch + n0 + c;
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            float4 f;
            f.x = __uint_as_float(r[4 * v + 0]) + __ldg(bp + 4 * v + 0);
            f.y = __uint_as_float(r[4 * v + 1]) + __ldg(bp + 4 * v + 1);
            f.z = __uint_as_float(r[4 * v + 2]) + __ldg(bp + 4 * v + 2);
            f.w = __uint_as_float(r[4 * v + 3]) + __ldg(bp + 4 * v + 3);
            *reinterpret_cast<float4*>(o + 4 * v) = f;
          }
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_acc, BN);
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

}  // namespace

int make_tmap_bf16_2d(void* tmap_out, const void* base, int64_t rows, int64_t cols, int box_rows, int box_cols) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (fn == nullptr) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(reinterpret_cast<CUtensorMap*>(tmap_out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim,
                  gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld)", (int)r,
                                 (long long)rows, (long long)cols);
  return XATTN_OK;
}

int gemm_tc_bf16(cudaStream_t st, const void* X, const void* W, const float* bias, void* out, bool out_bf16, int64_t out_pitch,
                 int M, int N, int K) {
  if (M <= 0) return XATTN_OK;
  if (N % BN != 0 || K % BK != 0)
    XF_FAIL(XATTN_ERR_SHAPE, "tcgen05 projection GEMM needs N %% 128 == 0 and K %% 64 == 0 (N=%d K=%d)", N, K);
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap_bf16_2d(&tmA, X, M, K, BM, BK));
  XF_TRY(make_tmap_bf16_2d(&tmB, W, N, K, BN, BK));
  static bool configured = false;
  if (!configured) {
    XF_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    XF_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    configured = true;
  }
  dim3 grid((M + BM - 1) / BM, N / BN);
  if (out_bf16)
    gemm_tc_kernel<true><<<grid, kGemmThreads, kSmemBytes, st>>>(tmA, tmB, bias, out, out_pitch, M, N, K);
  else
    gemm_tc_kernel<false><<<grid, kGemmThreads, kSmemBytes, st>>>(tmA, tmB, bias, out, out_pitch, M, N, K);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
