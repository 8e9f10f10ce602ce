// ldmatrix addressing of the m16n8k16 bf16 fragments over 136-element-pitch (272 B) smem tiles, shared by the per-clip
// kernels.  Fragment layout (PTX mma.m16n8k16): g = lane>>2, t = lane&3
//   A: a0=(g, 2t..), a1=(g+8, 2t..), a2=(g, 2t+8..), a3=(g+8, 2t+8..)      B: b0=(k=2t.., n=g), b1=(k=2t+8.., n=g)
//   C: c0=(g,2t) c1=(g,2t+1) c2=(g+8,2t) c3=(g+8,2t+1)
#pragma once
#include "bf16_layout.cuh"
#include "tc.cuh"

namespace xf {
namespace fr {

// A tile rows [row0, row0+16), k columns [16ks, 16ks+16) of a row-major [m][k] smem matrix.
__device__ __forceinline__ void ldA(uint32_t base, int row0, int ks, int lane, uint32_t (&a)[4]) {
  const int row = row0 + (lane & 15), col = ks * 16 + (lane >> 4) * 8;
  tc::ldsm_x4(base + (uint32_t)(row * BEP + col) * 2u, a[0], a[1], a[2], a[3]);
}
// B from a matrix stored [n][k] (k contiguous): n-tile nt, k-steps ks (b[0],b[1]) and ks+1 (b[2],b[3]); col_off shifts k.
__device__ __forceinline__ void ldB_nk(uint32_t base, int nt, int ks, int lane, uint32_t (&b)[4], int col_off = 0) {
  const int row = nt * 8 + (lane & 7), col = col_off + ks * 16 + (lane >> 3) * 8;
  tc::ldsm_x4(base + (uint32_t)(row * BEP + col) * 2u, b[0], b[1], b[2], b[3]);
}
// B from a matrix stored [k][n] (n contiguous): k rows [k0, k0+16), n-tiles 2*nt2 (b[0],b[1]) and 2*nt2+1 (b[2],b[3]).
__device__ __forceinline__ void ldB_kn(uint32_t base, int k0, int nt2, int lane, uint32_t (&b)[4]) {
  const int row = k0 + (lane & 15), col = nt2 * 16 + (lane >> 4) * 8;
  tc::ldsm_x4_t(base + (uint32_t)(row * BEP + col) * 2u, b[0], b[1], b[2], b[3]);
}
// C tiles (n-tiles 2ks, 2ks+1 of a 16 x N accumulator) -> A fragment of k-step ks for the next product.
__device__ __forceinline__ void c2a(const float (&c0)[4], const float (&c1)[4], uint32_t (&a)[4]) {
  a[0] = tc::pack_bf16(c0[0], c0[1]);
  a[1] = tc::pack_bf16(c0[2], c0[3]);
  a[2] = tc::pack_bf16(c1[0], c1[1]);
  a[3] = tc::pack_bf16(c1[2], c1[3]);
}
// cooperative copy of `bytes` (multiple of 16) global -> smem by the whole CTA
__device__ __forceinline__ void copy_g2s(void* dst, const void* src, int bytes, int tid, int nthreads) {
  const uint4* s = reinterpret_cast<const uint4*>(src);
  uint4* d = reinterpret_cast<uint4*>(dst);
  for (int i = tid; i < bytes / 16; i += nthreads) d[i] = __ldg(s + i);
}

}  // namespace fr
}  // namespace xf
