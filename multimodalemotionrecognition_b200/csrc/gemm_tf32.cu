// Tensor-core GEMMs of the training path (precision = XATTN_PREC_TF32): every nn.Linear forward / input gradient / weight
// gradient of the Stage-1 step (reference src/train.py:204-228 over src/models/fusion.py:372-411) whose dimensions allow it
// runs here instead of on the SIMT kernel of gemm_simt.cuh.  fp32 tensors are read as they are — TMA boxes of 32 floats
// (128 B rows, SWIZZLE_128B) feed tcgen05.mma.kind::tf32 (the tensor core reads the top 19 bits of each operand), the
// accumulators are fp32 in TMEM.  The bf16 embeddings of the first layer use kind::f16 through the same kernels.
//
//   gemm_tc_rows<B_MN = false>  C[M,N] (+)= A[M,K] · W[N,K]^T (+ bias)(ReLU)      nn.Linear forward       (both K-major)
//   gemm_tc_rows<B_MN = true >  C[M,N] (+)= A[M,K] · W[K,N]                       dX = dY · W             (B MN-major: the
//                               weight is read where it lies, rows = the reduction index, no transposed copy)
//   gemm_tc_cols<T>             C[N1,N2] += A[M,N1]^T · B[M,N2]                   dW = dY^T · X           (A and B MN-major:
//                               the reduction runs down the rows of both activations; split over M, partial tiles meet in C
//                               with red.global.add.f32 — as the SIMT kernel's split-K did)
// All three: persistent warp-specialised CTAs (warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue), a 6-stage ring of
// 32 KB stages, two 128-column TMEM accumulators.
#include <cuda.h>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int TM = 128, TN = 128;
constexpr int kRowBytes = 128;                    // one swizzle row: 32 fp32 or 64 bf16
constexpr int kOpBytes = TM * kRowBytes;          // one operand tile of a stage: 16 KB
constexpr int kStage = 2 * kOpBytes;              // 32 KB
constexpr int kStages = 5;   // 160 KB of operands in flight per SM; leaves ~40 KB of shared memory for a co-resident small kernel (SideFork)
constexpr int kAccs = 2;
constexpr int kThreads = 64 + 128;
constexpr int kEpiPitch = 36;                     // floats per staged row: 16-byte aligned, quarter-warp float4 accesses conflict-free
constexpr int kEpiBytes = 4 * 32 * kEpiPitch * 4; // one 32 x 32 transpose buffer per epilogue warp
constexpr int kSmem = kStages * kStage + 1024 /*bias*/ + 256 /*barriers*/ + kEpiBytes + 1024 /*alignment*/;
constexpr int kRedRows = 32;                      // reduction rows per stage of the MN-major operands

// Instruction descriptors: D = f32 (bits 4-5 = 1); A/B format at bits 7-9 / 10-12: bf16 = 1, tf32 = 2; bit 15 / 16 = A / B is
// MN-major; N >> 3 at 17, M >> 4 at 24.
__host__ __device__ constexpr uint32_t idesc_of(int fmt, int M, int N, bool a_mn, bool b_mn) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <bool TF32>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  if (TF32) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    tc::umma_bf16(tmem_d, desc_a, desc_b, idesc, accumulate);
  }
}

// MN-major operand of 32-bit elements: the 128-byte swizzle with 32-byte atoms (UMMA layout type 1, SWIZZLE_128B_BASE32B; TMA mode
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B).  Rows (reduction index) are 128 B apart, the swizzle pattern repeats every 4 rows: in
// 16-byte units the canonical layout is ((8,n),(4,k)):((1,LBO),(8,SBO)) with SBO = 512 B (cute/atom/mma_traits_sm100.hpp).
__device__ __forceinline__ uint64_t desc_mn_32b(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;
  return d;
}
// MN-major operand tile of kRedRows rows: 32-bit elements as above, 16-bit elements with the plain 128-byte swizzle.
template <bool TF32>
__device__ __forceinline__ uint64_t desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  return TF32 ? desc_mn_32b(smem_addr, lbo_bytes) : tc::umma_desc_sw128(smem_addr, lbo_bytes, 1024);
}

struct Bars {
  uint64_t full[kStages], empty[kStages], acc_full[kAccs], acc_empty[kAccs];
  uint32_t tmem_slot;
};

__device__ __forceinline__ uint8_t* align1k(uint8_t* p) {
  return reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~(uintptr_t)1023);
}

__device__ __forceinline__ uint32_t prologue(Bars* bars, int warp, int lane, const void* tm0, const void* tm1) {
  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(tm0);
    tc::tma_prefetch_desc(tm1);
    for (int s = 0; s < kStages; ++s) {
      tc::mbar_init(&bars->full[s], 1);
      tc::mbar_init(&bars->empty[s], 1);
    }
    for (int b = 0; b < kAccs; ++b) {
      tc::mbar_init(&bars->acc_full[b], 1);
      tc::mbar_init(&bars->acc_empty[b], 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 1) tc::tmem_alloc(&bars->tmem_slot, kAccs * TN);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  return bars->tmem_slot;
}

// ---------------------------------------------------------------------------------------------------------------------------
// C[M,N] (+)= A[M,K] · op(W) (+ bias)(ReLU); fp32 everywhere, tf32 products.  K % 32 == 0, N % 128 == 0.
// ---------------------------------------------------------------------------------------------------------------------------
template <bool B_MN>
__global__ void __launch_bounds__(kThreads, 1) gemm_tf32_rows_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                     const __grid_constant__ CUtensorMap tmB,
                                                                     const float* __restrict__ bias, float* __restrict__ C, int64_t ldc,
                                                                     int M, int N, int K, int relu, int accumulate) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1k(smem_raw);
  float* bias_s = reinterpret_cast<float*>(smem + kStages * kStage);
  Bars* bars = reinterpret_cast<Bars*>(smem + kStages * kStage + 1024);
  float* epi_s = reinterpret_cast<float*>(smem + kStages * kStage + 1024 + 256);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int KB = K / 32, tiles_n = N / TN, tiles_m = (M + TM - 1) / TM, num_tiles = tiles_m * tiles_n;
  const uint32_t tmem_base = prologue(bars, warp, lane, &tmA, &tmB);

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * TM, n0 = (tile % tiles_n) * TN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kStages;
          tc::mbar_wait(&bars->empty[s], ((it / kStages) & 1u) ^ 1u);
          tc::mbar_expect_tx(&bars->full[s], kStage);
          uint8_t* a = smem + s * kStage;
          tc::tma_load_2d(a, &tmA, &bars->full[s], kb * 32, m0);
          if (!B_MN) {
            tc::tma_load_2d(a + kOpBytes, &tmB, &bars->full[s], kb * 32, n0);          // W[N,K]: 128 rows n, 32 k
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)                                                 // W[K,N]: 32 rows k, 4 blocks of 32 n
              tc::tma_load_2d(a + kOpBytes + j * (kRedRows * kRowBytes), &tmB, &bars->full[s], n0 + 32 * j, kb * 32);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_of(2, TM, TN, false, B_MN);
      uint32_t it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
        const uint32_t ab = lt % kAccs;
        tc::mbar_wait(&bars->acc_empty[ab], ((lt / kAccs) & 1u) ^ 1u);
        tc::tc_fence_after();
        const uint32_t acc = tmem_base + ab * TN;
        for (int kb = 0; kb < KB; ++kb, ++it) {
          const int s = it % kStages;
          tc::mbar_wait(&bars->full[s], (it / kStages) & 1u);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + s * kStage);
          const uint64_t da = tc::umma_desc_sw128(a_addr, 16, 1024);
          const uint64_t db = B_MN ? desc_mn_32b(a_addr + kOpBytes, kRedRows * kRowBytes) : tc::umma_desc_sw128(a_addr + kOpBytes, 16, 1024);
#pragma unroll
          for (int k = 0; k < 4; ++k)   // K = 8 tf32 per instruction: +32 B inside the swizzle row (K-major) / +1024 B = 8 rows (MN-major)
            umma<true>(acc, da + (uint64_t)(2 * k), db + (uint64_t)(B_MN ? 64 * k : 2 * k), idesc, (kb | k) != 0);
          tc::umma_commit(&bars->empty[s]);
        }
        tc::umma_commit(&bars->acc_full[ab]);
      }
    }
  } else {
    const int q = warp & 3, et = threadIdx.x - 64;
    uint32_t lt = 0;
    int cur_n0 = -1;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++lt) {
      const int m0 = (tile / tiles_n) * TM, n0 = (tile % tiles_n) * TN;
      const uint32_t ab = lt % kAccs;
      if (n0 != cur_n0) {
        tc::named_bar_sync(2, 128);
        bias_s[et] = bias != nullptr ? __ldg(bias + n0 + et) : 0.f;
        cur_n0 = n0;
        tc::named_bar_sync(2, 128);
      }
      tc::mbar_wait(&bars->acc_full[ab], (lt / kAccs) & 1u);
      tc::tc_fence_after();
      // TMEM hands a lane one ROW of the tile (32 columns at a time).  Storing that directly makes every store instruction touch 32
      // different 128-byte lines with 16 bytes each; the strip goes through a per-warp shared-memory transpose instead, so that a
      // store instruction writes four complete lines (rows rr = 4i + lane/8, float4 column lane%8) — and the `accumulate` loads
      // are coalesced the same way.
      float* buf = epi_s + q * (32 * kEpiPitch);
      const int rsub = lane >> 3, c4 = lane & 7;
#pragma unroll 1
      for (int c = 0; c < TN; c += 32) {
        float* o = C + (int64_t)(m0 + q * 32 + rsub) * ldc + n0 + c + 4 * c4;   // row rsub of the strip; + 4·i rows per step
        float4 old[8];
        if (accumulate) {   // the previous values: eight independent loads in flight while the accumulator strip arrives
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (m0 + q * 32 + 4 * i + rsub < M) old[i] = __ldcs(reinterpret_cast<const float4*>(o + (int64_t)(4 * i) * ldc));
        }
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem_base + ab * TN + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int v = 0; v < 8; ++v) {
          float4 f;
          f.x = __uint_as_float(r[4 * v + 0]) + bias_s[c + 4 * v + 0];
          f.y = __uint_as_float(r[4 * v + 1]) + bias_s[c + 4 * v + 1];
          f.z = __uint_as_float(r[4 * v + 2]) + bias_s[c + 4 * v + 2];
          f.w = __uint_as_float(r[4 * v + 3]) + bias_s[c + 4 * v + 3];
          if (relu) { f.x = fmaxf(f.x, 0.f); f.y = fmaxf(f.y, 0.f); f.z = fmaxf(f.z, 0.f); f.w = fmaxf(f.w, 0.f); }
          *reinterpret_cast<float4*>(buf + lane * kEpiPitch + 4 * v) = f;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = 4 * i + rsub;
          if (m0 + q * 32 + rr < M) {
            float4 f = *reinterpret_cast<const float4*>(buf + rr * kEpiPitch + 4 * c4);
            if (accumulate) { f.x += old[i].x; f.y += old[i].y; f.z += old[i].z; f.w += old[i].w; }
            *reinterpret_cast<float4*>(o + (int64_t)(4 * i) * ldc) = f;
          }
        }
        __syncwarp();
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&bars->acc_empty[ab]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, kAccs * TN);
}

// ---------------------------------------------------------------------------------------------------------------------------
// C[N1,N2] += A[M,N1]^T · B[M,N2]; A, B both TF32 (fp32 in memory) or both bf16; N1, N2 % 128 == 0.  Work items are
// (n1 tile, n2 tile, slice of M); a slice's partial tile is added into C with fp32 reductions.
// ---------------------------------------------------------------------------------------------------------------------------
template <bool TF32>
__global__ void __launch_bounds__(kThreads, 1) gemm_cols_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB, float* __restrict__ C,
                                                                int64_t ldc, int M, int N1, int N2, int splits, int rows_per_split) {
  constexpr int EW = TF32 ? 32 : 64;              // elements per 128-byte swizzle row
  constexpr int NBLK = TM / EW;                   // column blocks per operand tile: 4 (tf32) / 2 (bf16)
  constexpr int KI = TF32 ? 8 : 16;               // reduction rows per instruction
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align1k(smem_raw);
  Bars* bars = reinterpret_cast<Bars*>(smem + kStages * kStage + 1024);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t1 = N1 / TM, t2 = N2 / TN, num_items = t1 * t2 * splits;
  const uint32_t tmem_base = prologue(bars, warp, lane, &tmA, &tmB);
  auto item_rows = [&](int item, int& r0, int& r1) {
    const int sp = item / (t1 * t2);
    r0 = sp * rows_per_split;
    r1 = min(M, r0 + rows_per_split);
  };

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int tt = item % (t1 * t2), a0c = (tt / t2) * TM, b0c = (tt % t2) * TN;
        int r0, r1;
        item_rows(item, r0, r1);
        for (int r = r0; r < r1; r += kRedRows, ++it) {   // (rows past M read as zeros: they add nothing)
          const int s = it % kStages;
          tc::mbar_wait(&bars->empty[s], ((it / kStages) & 1u) ^ 1u);
          tc::mbar_expect_tx(&bars->full[s], 2 * NBLK * kRedRows * kRowBytes);   // (bf16: half of the stage is used)
          uint8_t* a = smem + s * kStage;
#pragma unroll
          for (int j = 0; j < NBLK; ++j) {
            tc::tma_load_2d(a + j * (kRedRows * kRowBytes), &tmA, &bars->full[s], a0c + EW * j, r);
            tc::tma_load_2d(a + kOpBytes + j * (kRedRows * kRowBytes), &tmB, &bars->full[s], b0c + EW * j, r);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_of(TF32 ? 2 : 1, TM, TN, true, true);
      uint32_t it = 0, lt = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++lt) {
        const uint32_t ab = lt % kAccs;
        int r0, r1;
        item_rows(item, r0, r1);
        tc::mbar_wait(&bars->acc_empty[ab], ((lt / kAccs) & 1u) ^ 1u);
        tc::tc_fence_after();
        const uint32_t acc = tmem_base + ab * TN;
        for (int r = r0; r < r1; r += kRedRows, ++it) {
          const int s = it % kStages;
          tc::mbar_wait(&bars->full[s], (it / kStages) & 1u);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + s * kStage);
          const uint64_t da = desc_mn<TF32>(a_addr, kRedRows * kRowBytes);
          const uint64_t db = desc_mn<TF32>(a_addr + kOpBytes, kRedRows * kRowBytes);
#pragma unroll
          for (int k = 0; k < kRedRows / KI; ++k) {
            const uint64_t off = (uint64_t)((k * KI * kRowBytes) >> 4);
            umma<TF32>(acc, da + off, db + off, idesc, (r != r0 || k != 0) ? 1u : 0u);
          }
          tc::umma_commit(&bars->empty[s]);
        }
        tc::umma_commit(&bars->acc_full[ab]);
      }
    }
  } else {
    const int q = warp & 3;
    uint32_t lt = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++lt) {
      const int tt = item % (t1 * t2), a0c = (tt / t2) * TM, b0c = (tt % t2) * TN;
      const uint32_t ab = lt % kAccs;
      tc::mbar_wait(&bars->acc_full[ab], (lt / kAccs) & 1u);
      tc::tc_fence_after();
      float* o = C + (int64_t)(a0c + q * 32 + lane) * ldc + b0c;
#pragma unroll 1
      for (int c = 0; c < TN; c += 32) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem_base + ab * TN + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
        tc::tmem_ld_wait();
        if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
#pragma unroll
          for (int v = 0; v < 8; ++v)   // red.global.add.v4.f32
            atomicAdd(reinterpret_cast<float4*>(o + c) + v, make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                                                                         __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3])));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) atomicAdd(o + c + i, __uint_as_float(r[i]));
        }
      }
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&bars->acc_empty[ab]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, kAccs * TN);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2D map over a row-major [rows, cols] tensor (pitch in elements), boxes of box_rows x (128 bytes of columns), 128-byte swizzle.
int make_tmap(CUtensorMap* tm, const void* base, bool f32, int64_t rows, int64_t cols, int64_t pitch, int box_rows, bool mn_major = false) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const int es = f32 ? 4 : 2;
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0 || (pitch * es) % 16 != 0)
    XF_FAIL(XATTN_ERR_ALIGN, "tensor-core GEMM operand: base and row pitch must be 16-byte aligned");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)pitch * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, (mn_major && f32) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) XF_FAIL(XATTN_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld cols=%lld)", (int)r,
                                 (long long)rows, (long long)cols);
  return XATTN_OK;
}

template <typename K>
int set_smem(K kernel) {
  return ensure_dyn_smem(kernel, kSmem);
}

}  // namespace

bool gemm_tf32_rows_ok(int M, int N, int K, int64_t lda, int64_t ldc) {
  return M > 0 && N % TN == 0 && K % 32 == 0 && lda % 4 == 0 && ldc % 4 == 0;
}
bool gemm_cols_ok(int M, int N1, int N2, int64_t lda, int64_t ldb, bool b_bf16) {
  return M > 0 && N1 % TM == 0 && N2 % TN == 0 && lda % (b_bf16 ? 8 : 4) == 0 && ldb % (b_bf16 ? 8 : 4) == 0;
}

int gemm_tf32_rows(cudaStream_t st, const float* A, int64_t lda, const float* W, bool w_is_kn, const float* bias, float* C, int64_t ldc,
                   int M, int N, int K, bool relu, bool accumulate) {
  if (!gemm_tf32_rows_ok(M, N, K, lda, ldc)) XF_FAIL(XATTN_ERR_SHAPE, "tf32 GEMM: M=%d N=%d K=%d lda=%lld ldc=%lld", M, N, K, (long long)lda, (long long)ldc);
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap(&tmA, A, true, M, K, lda, TM));
  if (!w_is_kn) XF_TRY(make_tmap(&tmB, W, true, N, K, K, TN));          // W[N,K]
  else XF_TRY(make_tmap(&tmB, W, true, K, N, N, kRedRows, true));       // W[K,N]
  XF_TRY(set_smem(gemm_tf32_rows_kernel<false>));
  XF_TRY(set_smem(gemm_tf32_rows_kernel<true>));
  const int tiles = ((M + TM - 1) / TM) * (N / TN);
  const int grid = tiles < num_sms() ? tiles : num_sms();
  if (!w_is_kn)
    gemm_tf32_rows_kernel<false><<<grid, kThreads, kSmem, st>>>(tmA, tmB, bias, C, ldc, M, N, K, relu ? 1 : 0, accumulate ? 1 : 0);
  else
    gemm_tf32_rows_kernel<true><<<grid, kThreads, kSmem, st>>>(tmA, tmB, bias, C, ldc, M, N, K, relu ? 1 : 0, accumulate ? 1 : 0);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int gemm_cols(cudaStream_t st, const void* A, int64_t lda, const void* B, int64_t ldb, bool bf16, float* C, int64_t ldc, int M, int N1,
              int N2) {
  if (!gemm_cols_ok(M, N1, N2, lda, ldb, bf16)) XF_FAIL(XATTN_ERR_SHAPE, "column GEMM: M=%d N1=%d N2=%d", M, N1, N2);
  alignas(64) CUtensorMap tmA, tmB;
  XF_TRY(make_tmap(&tmA, A, !bf16, M, N1, lda, kRedRows, true));
  XF_TRY(make_tmap(&tmB, B, !bf16, M, N2, ldb, kRedRows, true));
  XF_TRY(set_smem(gemm_cols_kernel<false>));
  XF_TRY(set_smem(gemm_cols_kernel<true>));
  const int tiles = (N1 / TM) * (N2 / TN);
  int splits = (2 * num_sms() + tiles - 1) / tiles;            // ~2 work items per CTA
  const int max_splits = (M + 4 * kRedRows - 1) / (4 * kRedRows);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int rows_per_split = ((M + splits - 1) / splits + kRedRows - 1) / kRedRows * kRedRows;
  splits = (M + rows_per_split - 1) / rows_per_split;
  const int items = tiles * splits;
  const int grid = items < num_sms() ? items : num_sms();
  if (bf16) gemm_cols_kernel<false><<<grid, kThreads, kSmem, st>>>(tmA, tmB, C, ldc, M, N1, N2, splits, rows_per_split);
  else gemm_cols_kernel<true><<<grid, kThreads, kSmem, st>>>(tmA, tmB, C, ldc, M, N1, N2, splits, rows_per_split);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
