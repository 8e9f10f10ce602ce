// Hardware probe for the tcgen05 features the attention kernels rely on beyond what the projection GEMM already uses:
//   * M = 64 accumulators placed in the lower (lanes 0-15) and upper (16-31) half of every TMEM sub-partition,
//   * an MN-major B operand (rows = K) read from a 128-byte-swizzled tile,
//   * hand-written (generic-proxy) swizzled operand tiles, and the 16x256b TMEM load shape.
// Two independent products run in one CTA:  D_lo = A[0:64]·op(B0)  -> lanes 0-15 of each sub-partition,
//                                           D_hi = A[64:128]·op(B1) -> lanes 16-31.
// mode 0: B_i given as [N][K] (K contiguous, K-major operand); mode 1: B_i given as [K][N] (N contiguous, MN-major operand).
// mode 2 / 3: as 0 / 1 with the A operands in TENSOR MEMORY (tcgen05.st of bf16x2-packed rows into the lanes of the matching D).
// Exported through xattn_debug_umma_probe for tests/test_gpu_bf16.py; not on the product path.
#include "common.cuh"
#include "tc.cuh"

namespace xf {

namespace {

constexpr int kProbeMaxBytes = 200 * 1024;

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                                                            float* __restrict__ out32, float* __restrict__ out16, int N, int K,
                                                            int mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int KA = (K + 63) / 64;                       // 64-element K blocks of the A tiles
  const int a_bytes = KA * 64 * 128;                  // one [64 rows] A operand: KA boxes of 64 x 128 B
  // B operand of product i: mode 0 -> KA boxes of [N rows x 128 B]; mode 1 -> N/64 boxes of [Kp rows x 128 B]
  const int Kp = (K + 15) / 16 * 16;
  const int b_box = (mode & 1) == 0 ? N * 128 : Kp * 128;
  const int b_bytes = ((mode & 1) == 0 ? KA : N / 64) * b_box;
  uint8_t* sA[2] = {smem, smem + a_bytes};
  uint8_t* sB[2] = {smem + 2 * a_bytes, smem + 2 * a_bytes + ((b_bytes + 1023) & ~1023)};
  for (int i = tid; i < (2 * a_bytes + 2 * ((b_bytes + 1023) & ~1023)) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  for (int p = 0; p < 2; ++p) {
    for (int i = tid; i < 64 * K; i += 128) {
      const int r = i / K, k = i % K;
      *reinterpret_cast<__nv_bfloat16*>(sA[p] + (k / 64) * 64 * 128 + tc::sw128_off(r, k % 64)) = A[(int64_t)(p * 64 + r) * K + k];
    }
    if ((mode & 1) == 0) {
      for (int i = tid; i < N * K; i += 128) {
        const int n = i / K, k = i % K;
        *reinterpret_cast<__nv_bfloat16*>(sB[p] + (k / 64) * b_box + tc::sw128_off(n, k % 64)) = B[(int64_t)p * N * K + i];
      }
    } else {
      for (int i = tid; i < K * N; i += 128) {
        const int k = i / N, n = i % N;
        *reinterpret_cast<__nv_bfloat16*>(sB[p] + (n / 64) * b_box + tc::sw128_off(k, n % 64)) = B[(int64_t)p * N * K + i];
      }
    }
  }
  if (tid == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_proxy_async();
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const bool a_tmem = mode >= 2;
  mode &= 1;
  constexpr uint32_t kTmemA = 256;   // A rows, packed: columns [256, 256 + K/2)
  if (a_tmem) {   // lane l of warp w holds row 16w + (l & 15) of product (l >> 4)
    const int p = lane >> 4, r = 16 * warp + (lane & 15);
    for (int k0 = 0; k0 < Kp; k0 += 32) {
      uint32_t v[16];
      for (int j = 0; j < 16; ++j) {
        const int k = k0 + 2 * j;
        const __nv_bfloat16 lo = k < K ? A[(int64_t)(p * 64 + r) * K + k] : __float2bfloat16(0.f);
        const __nv_bfloat16 hi = k + 1 < K ? A[(int64_t)(p * 64 + r) * K + k + 1] : __float2bfloat16(0.f);
        v[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
      }
      tc::tmem_st_32x32_x16(tmem + kTmemA + ((uint32_t)(32 * warp) << 16) + (uint32_t)(k0 / 2), v);
    }
    tc::tmem_st_wait();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
  }
  if (tid == 0) {
    const uint32_t idesc = tc::umma_idesc_bf16_ex(64, N, false, mode == 1);
    for (int p = 0; p < 2; ++p) {
      const uint32_t d = tmem + ((uint32_t)(16 * p) << 16);
      for (int ks = 0; ks < Kp / 16; ++ks) {
        const uint64_t da = tc::umma_desc_sw128(tc::smem_u32(sA[p]) + (ks / 4) * 64 * 128 + (ks % 4) * 32, 16, 1024);
        uint64_t db;
        if (mode == 0) db = tc::umma_desc_sw128(tc::smem_u32(sB[p]) + (ks / 4) * b_box + (ks % 4) * 32, 16, 1024);
        else db = tc::umma_desc_sw128(tc::smem_u32(sB[p]) + ks * 16 * 128, (uint32_t)b_box, 1024);
        if (a_tmem) tc::umma_bf16_ts(d, tmem + kTmemA + ((uint32_t)(16 * p) << 16) + (uint32_t)(ks * 8), db, idesc, ks != 0);
        else tc::umma_bf16(d, da, db, idesc, ks != 0);
      }
    }
    tc::umma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::tc_fence_after();
  for (int c = 0; c < N; c += 32) {
    uint32_t r[32];
    tc::tmem_ld_32x32(tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)c, r);
    tc::tmem_ld_wait();
    for (int i = 0; i < 32 && c + i < N; ++i) out32[(int64_t)(32 * warp + lane) * N + c + i] = __uint_as_float(r[i]);
  }
  if (N >= 128) {   // 16x256b shape on both halves of the sub-partition: [half][warp][lane][64]
    for (int p = 0; p < 2; ++p) {
      uint32_t r[64];
      tc::tmem_ld_16x256_x16(tmem + ((uint32_t)(32 * warp + 16 * p) << 16), r);
      tc::tmem_ld_wait();
      for (int i = 0; i < 64; ++i) out16[(((int64_t)p * 4 + warp) * 32 + lane) * 64 + i] = __uint_as_float(r[i]);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

}  // namespace

int umma_probe(cudaStream_t st, const void* A, const void* B, float* out32, float* out16, int N, int K, int mode) {
  if (N % 8 != 0 || N < 8 || N > 256 || K < 16 || K % 16 != 0 || K > 256 || mode < 0 || mode > 3 ||
      ((mode & 1) == 1 && N % 64 != 0) || ((mode & 1) == 0 && K % 64 != 0))
    XF_FAIL(XATTN_ERR_SHAPE, "umma probe: unsupported N=%d K=%d mode=%d", N, K, mode);
  const int KA = (K + 63) / 64, Kp = (K + 15) / 16 * 16;
  const int b_bytes = (mode & 1) == 0 ? KA * N * 128 : (N / 64) * Kp * 128;
  const int smem = 2 * KA * 64 * 128 + 2 * ((b_bytes + 1023) & ~1023) + 1024;
  if (smem > kProbeMaxBytes) XF_FAIL(XATTN_ERR_SHAPE, "umma probe: operands need %d bytes of smem", smem);
  XF_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  umma_probe_kernel<<<1, 128, smem, st>>>(reinterpret_cast<const __nv_bfloat16*>(A), reinterpret_cast<const __nv_bfloat16*>(B), out32,
                                         out16, N, K, mode);
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
