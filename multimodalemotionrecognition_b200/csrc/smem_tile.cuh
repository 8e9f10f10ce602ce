// Register-tiled products over operands that sit in shared memory — the per-clip, per-head matrix products of the fp32
// attention kernels (fwd_f32.cu / bwd_f32.cu) when the head width is one warp (dh == 32).  The plain loops they replace spent
// two shared-memory loads on every FMA; here a warp shares TM broadcast rows and every lane reuses them.
#pragma once
#include "common.cuh"

namespace xf {

// out[e] = mean over the L rows of x[L][E] (global, rows 16-byte aligned when E % 4 == 0), for every thread of the CTA; out and
// scratch ((nthreads/32)·E floats) in shared memory.  Warps stride over the rows, a lane owns four columns (float4 loads, eight
// in flight), partial sums meet in scratch.  The loop it replaces walked the rows with E of the CTA's threads, one scalar
// load per row.  Ends with __syncthreads().
__device__ __forceinline__ void clip_mean(const float* __restrict__ x, int L, int E, float* __restrict__ out, float* __restrict__ scratch,
                                          int tid, int nthreads) {
  const int lane = tid & 31, warp = tid >> 5, nw = nthreads >> 5;
  if ((E & 3) != 0) {
    for (int e = tid; e < E; e += nthreads) {
      float s = 0.f;
      for (int t = 0; t < L; ++t) s += x[(int64_t)t * E + e];
      out[e] = s / (float)L;
    }
    __syncthreads();
    return;
  }
  for (int c4 = lane; c4 < (E >> 2); c4 += 32) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int t = warp; t < L; t += nw) {
      const float4 u = __ldg(reinterpret_cast<const float4*>(x + (int64_t)t * E) + c4);
      s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
    }
    float* d = scratch + warp * E + 4 * c4;
    d[0] = s.x; d[1] = s.y; d[2] = s.z; d[3] = s.w;
  }
  __syncthreads();
  for (int e = tid; e < E; e += nthreads) {
    float s = 0.f;
    for (int w = 0; w < nw; ++w) s += scratch[w * E + e];
    out[e] = s / (float)L;
  }
  __syncthreads();
}

// Stage `rows` rows of 32 floats (one head's slice; global row stride gs floats, 16-byte aligned) into dst[rows][33], times scale.
// float4 loads, four in flight per thread before the first store: the scalar load->store loop this replaces held 40 % of the
// attention kernels' stall samples on its long-scoreboard waits.  Eight lanes cover one row; with the odd pitch the four
// scalar stores of a quarter-warp's rows fall in distinct banks.
__device__ __forceinline__ void stage_rows32(float* __restrict__ dst, const float* __restrict__ src, int rows, int64_t gs, float scale,
                                             int tid, int nthreads) {
  const int n4 = rows * 8;
  for (int base = 0; base < n4; base += 4 * nthreads) {
    float4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = base + tid + j * nthreads;
      if (i < n4) v[j] = __ldg(reinterpret_cast<const float4*>(src + (int64_t)(i >> 3) * gs + (i & 7) * 4));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = base + tid + j * nthreads;
      if (i < n4) {
        float* d = dst + (i >> 3) * 33 + (i & 7) * 4;
        d[0] = v[j].x * scale; d[1] = v[j].y * scale; d[2] = v[j].z * scale; d[3] = v[j].w * scale;
      }
    }
  }
}
// The same for two slices of one row at once (K and V of a head: src and src + second), so both sets of loads are in flight together.
__device__ __forceinline__ void stage_rows32x2(float* __restrict__ dst0, float* __restrict__ dst1, const float* __restrict__ src,
                                               int64_t second, int rows, int64_t gs, int tid, int nthreads) {
  const int n4 = rows * 8;
  for (int base = 0; base < n4; base += 2 * nthreads) {
    float4 v[2], w[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int i = base + tid + j * nthreads;
      if (i < n4) {
        const float* g = src + (int64_t)(i >> 3) * gs + (i & 7) * 4;
        v[j] = __ldg(reinterpret_cast<const float4*>(g));
        w[j] = __ldg(reinterpret_cast<const float4*>(g + second));
      }
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int i = base + tid + j * nthreads;
      if (i < n4) {
        const int o = (i >> 3) * 33 + (i & 7) * 4;
        dst0[o] = v[j].x; dst0[o + 1] = v[j].y; dst0[o + 2] = v[j].z; dst0[o + 3] = v[j].w;
        dst1[o] = w[j].x; dst1[o + 1] = w[j].y; dst1[o + 2] = w[j].z; dst1[o + 3] = w[j].w;
      }
    }
  }
}

// acc[i][j] = sum_k A[(a0+i)*pa + k] * B[(b0 + lane + 32j)*pb + k],  k < K.  Rows past the operands are clamped (the caller drops
// those results).  pa, pb odd (dh + 1): the lane-strided rows of B fall in distinct banks, the rows of A are broadcasts.
template <int TM, int TN>
__device__ __forceinline__ void tile_nt(const float* __restrict__ A, int a0, int a_rows, int pa, const float* __restrict__ B, int b0,
                                        int b_rows, int pb, int K, int lane, float (&acc)[TM][TN]) {
  const float* ap[TM];
  const float* bp[TN];
#pragma unroll
  for (int i = 0; i < TM; ++i) ap[i] = A + min(a0 + i, a_rows - 1) * pa;
#pragma unroll
  for (int j = 0; j < TN; ++j) bp[j] = B + min(b0 + lane + 32 * j, b_rows - 1) * pb;
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    float a[TM], b[TN];
#pragma unroll
    for (int i = 0; i < TM; ++i) a[i] = ap[i][k];
#pragma unroll
    for (int j = 0; j < TN; ++j) b[j] = bp[j][k];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
      for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
  }
}

// acc[i] = sum_k A[(r0+i)*pa + k] * B[k*pb + lane],  k < K.  pa % 4 == 0 and A 16-byte aligned: four k at a time come in as one
// broadcast float4 per row; entries A[r][k >= K] up to the next multiple of 4 must be readable (they are multiplied by clamped B
// rows and must be ZERO).  Rows past `rows` are clamped.
template <int TM>
__device__ __forceinline__ void tile_ak(const float* __restrict__ A, int r0, int rows, int pa, const float* __restrict__ B, int pb, int K,
                                        int lane, float (&acc)[TM]) {
  const float4* ap[TM];
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    ap[i] = reinterpret_cast<const float4*>(A + min(r0 + i, rows - 1) * pa);
    acc[i] = 0.f;
  }
  const float* bp = B + lane;
  const int K4 = (K + 3) >> 2;
#pragma unroll 2
  for (int k4 = 0; k4 < K4; ++k4) {
    const int k = 4 * k4;
    const float b0 = bp[min(k, K - 1) * pb], b1 = bp[min(k + 1, K - 1) * pb], b2 = bp[min(k + 2, K - 1) * pb], b3 = bp[min(k + 3, K - 1) * pb];
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const float4 a = ap[i][k4];
      acc[i] = fmaf(a.x, b0, fmaf(a.y, b1, fmaf(a.z, b2, fmaf(a.w, b3, acc[i]))));
    }
  }
}

// acc[i] = sum_k A[k*pa + m0 + i] * B[k*pb + lane],  k < K, TM in {2, 4}.  pa % 4 == 0, m0 % TM == 0: the TM consecutive columns of a
// row of A come in as one broadcast vector load (columns past the operand are padding the caller ignores).
template <int TM>
__device__ __forceinline__ void tile_am(const float* __restrict__ A, int m0, int pa, const float* __restrict__ B, int pb, int K, int lane,
                                        float (&acc)[TM]) {
  static_assert(TM == 2 || TM == 4, "tile_am: TM is 2 or 4");
#pragma unroll
  for (int i = 0; i < TM; ++i) acc[i] = 0.f;
  const float* ap = A + m0;
  const float* bp = B + lane;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float b = bp[k * pb];
    if (TM == 4) {
      const float4 a = *reinterpret_cast<const float4*>(ap + k * pa);
      acc[0] = fmaf(a.x, b, acc[0]); acc[1] = fmaf(a.y, b, acc[1]); acc[2] = fmaf(a.z, b, acc[2]); acc[3] = fmaf(a.w, b, acc[3]);
    } else {
      const float2 a = *reinterpret_cast<const float2*>(ap + k * pa);
      acc[0] = fmaf(a.x, b, acc[0]); acc[1] = fmaf(a.y, b, acc[1]);
    }
  }
}

}  // namespace xf
