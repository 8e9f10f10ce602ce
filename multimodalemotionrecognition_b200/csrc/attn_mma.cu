// Tensor-core attention core of the TRAINING path (precision = XATTN_PREC_TF32), forward and backward:
//   S_h = (Q_h / sqrt(dh)) K_h^T + tanh(qs_l + ks_s)·bias_scale ;  P_h = softmax(S_h) ;  O_h = dropout(P_h) V_h
// (nn.MultiheadAttention's need_weights branch, reference src/models/fusion.py:394 / :398 under loss.backward() at
// src/train.py:227; the bias is fusion.py:175-176, head-independent).  Same arguments, saved tensors (P before dropout), Philox
// masks and side outputs as the fp32 SIMT kernels attn_fwd32_kernel / attn_bwd32_kernel (fwd_f32.cu / bwd_f32.cu), which stay the
// 1e-4 parity path; here the six contractions per head run as mma.sync m16n8k8 tf32 tiles with fp32 accumulation:
//   forward   S = Q K^T                    O = Pd V
//   backward  dPd = dO V^T   dV = Pd^T dO   dK = dS^T Q   dQ = dS K
// The operands are the fp32 tensors themselves: the tensor core reads the upper 19 bits of each 32-bit word (tf32 by truncation),
// which is what oracle/torch_port.py (xattn_forward_tf32, attn_tf32) applies.
//
// One CTA per clip.  A work item is (head, tile of 16 queries); a warp owns an item end to end: the 16 x S score tile lives in the
// warp's mma accumulators (S <= 160: 20 key tiles of 8), goes through a warp-private shared-memory tile once — for the row-major
// dropout pass (one Philox evaluation per four consecutive elements, as the SIMT kernels draw them) and to come back as an A
// operand — and every global operand is read exactly once, straight into fragments: a permutation of the reduction index (the
// same on both operands) or of the output columns lets each thread fetch whole 16-byte pieces of a row.
//   v2a (L = T <= 16, S = Ta <= 160): 4 items per clip, one per warp.      a2v (L = Ta <= 160, S = T <= 16): 40 items, 8 warps.
#include "kernels_f32.cuh"

namespace xf {

namespace {

__device__ __forceinline__ void mma_tf32(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t fu(float x) { return __float_as_uint(x); }
// global -> shared without a register round trip: the copy completes while the warp issues its first products
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// pull `bytes` of a clip's operand towards L2 ahead of the fragment loads that will want it (one line per thread and step)
__device__ __forceinline__ void prefetch_l2(const float* base, int64_t bytes, int tid, int nthreads) {
  const char* p = reinterpret_cast<const char*>(base);
  for (int64_t o = (int64_t)tid * 128; o < bytes; o += (int64_t)nthreads * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(p + o));
}
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
// tanh for the prior bias tile: 1 - 2 / (e^{2x} + 1) on ex2.approx / rcp.approx — absolute error < 1e-6 (the bias is an additive
// score term next to tf32 products), exact limits ±1, six instructions instead of tanhf's ~35 with a divergent branch.  Forward and
// backward use the same function, so (1 - t²) in the backward is the derivative of what the forward added.  (The fp32 SIMT kernels
// keep tanhf.)
__device__ __forceinline__ float tanh_fast(float x) { return 1.f - __fdividef(2.f, __expf(2.f * x) + 1.f); }

// Fragment layouts of mma.m16n8k8 (lane = 4 g + t):
//   A[16x8]: a0 (g, t)  a1 (g+8, t)  a2 (g, t+4)  a3 (g+8, t+4)      B[8x8]: b0 (k = t, n = g)  b1 (k = t+4, n = g)
//   C[16x8]: c0 (g, 2t)  c1 (g, 2t+1)  c2 (g+8, 2t)  c3 (g+8, 2t+1)
// "K-permuted" operand: a 32-wide reduction index d is cut into four k-steps; logical position (ks, t + 4 j) of step ks is
// physical d = 8 t + 2 ks + j, so a thread's eight values of a row are the 32 contiguous bytes [8t, 8t+8).
struct RowFrag {   // two rows (g, g+8) of a 32-wide K-permuted A operand: a[ks] = {row0[2ks], row1[2ks], row0[2ks+1], row1[2ks+1]}
  uint32_t a[4][4];
};
__device__ __forceinline__ void load_rowfrag(RowFrag& f, const float* row0, const float* row1, bool ok0, bool ok1, float scale, int t) {
  const float4 x0 = ldg4(row0 + 8 * t), x1 = ldg4(row0 + 8 * t + 4), y0 = ldg4(row1 + 8 * t), y1 = ldg4(row1 + 8 * t + 4);
  const float s0 = ok0 ? scale : 0.f, s1 = ok1 ? scale : 0.f;   // rows past the tile's last query contribute nothing
  const float u[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w}, v[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    f.a[ks][0] = fu(u[2 * ks] * s0);
    f.a[ks][1] = fu(v[2 * ks] * s1);
    f.a[ks][2] = fu(u[2 * ks + 1] * s0);
    f.a[ks][3] = fu(v[2 * ks + 1] * s1);
  }
}
// acc[16 x 8] += A(rows g, g+8; K-permuted) · X[row = n0 + g][32]^T  — X row read as the matching K-permuted B operand
__device__ __forceinline__ void mma_rows_nt(float (&acc)[4], const RowFrag& f, const float* xrow, int t) {
  const float4 k0 = ldg4(xrow + 8 * t), k1 = ldg4(xrow + 8 * t + 4);
  const float kk[8] = {k0.x, k0.y, k0.z, k0.w, k1.x, k1.y, k1.z, k1.w};
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) mma_tf32(acc, f.a[ks][0], f.a[ks][1], f.a[ks][2], f.a[ks][3], fu(kk[2 * ks]), fu(kk[2 * ks + 1]));
}
__device__ __forceinline__ void mma_rows_nt_frag(float (&acc)[4], const RowFrag& f, const float (&kk)[8]) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) mma_tf32(acc, f.a[ks][0], f.a[ks][1], f.a[ks][2], f.a[ks][3], fu(kk[2 * ks]), fu(kk[2 * ks + 1]));
}
__device__ __forceinline__ void load_row8(float (&kk)[8], const float* xrow, int t) {
  const float4 k0 = ldg4(xrow + 8 * t), k1 = ldg4(xrow + 8 * t + 4);
  kk[0] = k0.x; kk[1] = k0.y; kk[2] = k0.z; kk[3] = k0.w; kk[4] = k1.x; kk[5] = k1.y; kk[6] = k1.z; kk[7] = k1.w;
}
// out[16 x 32] += A[16 x 8 ksteps·8] (canonical, from the shared-memory tile `ps`, pitch P) · X[row = key][32]: "N-permuted"
// output — logical column n of output tile i is physical column 4 n + i, so thread g fetches X[key][4g .. 4g+3] in one piece
// and a thread ends up with the physical columns 8t .. 8t+7 of its two rows.
template <int PITCH>
__device__ __forceinline__ void mma_tile_x(float (&o)[4][4], const float* ps, int kstep, const float* xrow0, const float* xrow1, int g,
                                           int t) {
  const uint32_t a0 = fu(ps[g * PITCH + 8 * kstep + t]), a1 = fu(ps[(g + 8) * PITCH + 8 * kstep + t]);
  const uint32_t a2 = fu(ps[g * PITCH + 8 * kstep + t + 4]), a3 = fu(ps[(g + 8) * PITCH + 8 * kstep + t + 4]);
  const float4 v0 = ldg4(xrow0 + 4 * g), v1 = ldg4(xrow1 + 4 * g);
  mma_tf32(o[0], a0, a1, a2, a3, fu(v0.x), fu(v1.x));
  mma_tf32(o[1], a0, a1, a2, a3, fu(v0.y), fu(v1.y));
  mma_tf32(o[2], a0, a1, a2, a3, fu(v0.z), fu(v1.z));
  mma_tf32(o[3], a0, a1, a2, a3, fu(v0.w), fu(v1.w));
}
template <int PITCH>
__device__ __forceinline__ void mma_tile_x_frag(float (&o)[4][4], const float* ps, int kstep, const float4& v0, const float4& v1, int g, int t) {
  const uint32_t a0 = fu(ps[g * PITCH + 8 * kstep + t]), a1 = fu(ps[(g + 8) * PITCH + 8 * kstep + t]);
  const uint32_t a2 = fu(ps[g * PITCH + 8 * kstep + t + 4]), a3 = fu(ps[(g + 8) * PITCH + 8 * kstep + t + 4]);
  mma_tf32(o[0], a0, a1, a2, a3, fu(v0.x), fu(v1.x));
  mma_tf32(o[1], a0, a1, a2, a3, fu(v0.y), fu(v1.y));
  mma_tf32(o[2], a0, a1, a2, a3, fu(v0.z), fu(v1.z));
  mma_tf32(o[3], a0, a1, a2, a3, fu(v0.w), fu(v1.w));
}
// the two rows (g, g+8) of an N-permuted 16 x 32 result -> global rows (physical columns 8t .. 8t+7 each)
__device__ __forceinline__ void store_rows_x(const float (&o)[4][4], float* row0, float* row1, bool ok0, bool ok1, float scale, int t) {
  if (ok0) {
    *reinterpret_cast<float4*>(row0 + 8 * t) = make_float4(o[0][0] * scale, o[1][0] * scale, o[2][0] * scale, o[3][0] * scale);
    *reinterpret_cast<float4*>(row0 + 8 * t + 4) = make_float4(o[0][1] * scale, o[1][1] * scale, o[2][1] * scale, o[3][1] * scale);
  }
  if (ok1) {
    *reinterpret_cast<float4*>(row1 + 8 * t) = make_float4(o[0][2] * scale, o[1][2] * scale, o[2][2] * scale, o[3][2] * scale);
    *reinterpret_cast<float4*>(row1 + 8 * t + 4) = make_float4(o[0][3] * scale, o[1][3] * scale, o[2][3] * scale, o[3][3] * scale);
  }
}

// Row-major pass of a warp over its 16 x S tile: element (r, s) has the linear index base_r + s in the [B,H,L,S] tensor; the four
// dropout factors of indices 4c .. 4c+3 come from ONE Philox evaluation (drop_factor4), so a lane takes one aligned group of four
// at a time — (row, group) pairs are dealt to the lanes flat, so short rows (S = 16: four groups) keep the whole warp busy.
// fn4(r, s, f[4]) gets a whole group inside the row (s % 4 == 0 in index space), fn(r, s, f) single elements at the row ends.
template <typename F4, typename F1>
__device__ __forceinline__ void for_each_with_dropout(int rows_valid, int S, int64_t base0, int64_t row_stride, float p, uint64_t seed,
                                                      uint64_t offset, uint32_t stream, int lane, F4&& fn4, F1&& fn) {
  const int G = (S + 6) >> 2;   // upper bound of the aligned groups a row can touch
  // flat index lane, lane + 32, ... over (row, group) without a division per step: (r, grp) advance by 32 = dr·G + dg
  const int dr = 32 / G, dg = 32 - dr * G;
  int r = lane / G, grp = lane - r * G;
  while (r < rows_valid) {
    const int64_t base = base0 + (int64_t)r * row_stride;
    const int rem = (int)(base & 3);
    if (grp < ((S + rem + 3) >> 2)) {
      float f[4] = {1.f, 1.f, 1.f, 1.f};
      if (p > 0.f) drop_factor4(p, seed, offset, stream, (uint64_t)(base - rem) + 4u * (uint64_t)grp, f);
      const int s0 = 4 * grp - rem;
      if (s0 >= 0 && s0 + 3 < S) {
        fn4(r, s0, f);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (s0 + j >= 0 && s0 + j < S) fn(r, s0 + j, f[j]);
      }
    }
    r += dr;
    grp += dg;
    if (grp >= G) {
      grp -= G;
      ++r;
    }
  }
}

// --------------------------------------------------------------------------------------------------------------------------
template <int NT, int NW>
__global__ void __launch_bounds__(NW * 32, NW == 4 ? 4 : 2) attn_fwd_mma_kernel(AttnArgs a) {
  constexpr int PITCH = NT * 8 + 4;   // (≡ 4 mod 32: the canonical A-fragment reads of a tile are conflict-free)
  extern __shared__ __align__(16) float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int L = a.L, S = a.S, E = a.E, H = a.H;
  const int n_mt = (L + 15) >> 4, Lp = n_mt * 16;
  float* bt = sm;                                           // [Lp][PITCH]: tanh(qs_l + ks_s)·bias_scale, shared by the heads
  float* ps = sm + (a.qs ? Lp * PITCH : 0) + warp * 16 * PITCH;
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  if (NT <= 2) prefetch_l2(q, (int64_t)L * E * 4, tid, NW * 32);   // a2v: ten query tiles per head, fetched item by item
  if (a.qs) {
    const float bscale = a.bias_scale[0];
    const float* qs = a.qs + (int64_t)b * L;
    const float* ks = a.ks + (int64_t)b * S;
    for (int i = tid; i < Lp * NT * 8; i += NW * 32) {
      const int l = i / (NT * 8), s = i - l * (NT * 8);
      bt[l * PITCH + s] = (l < L && s < S) ? tanh_fast(qs[l] + ks[s]) * bscale : 0.f;
    }
    __syncthreads();
  }
  // Short key side (a2v: 16 video tokens): the K / V fragments of a head are the same for all of its query tiles, and a warp stays on
  // one head (NW % H == 0) — they are fetched once per warp instead of once per item.
  constexpr bool HOIST = NT <= 2;
  constexpr int NH = HOIST ? NT : 1;
  float kfr[NH][8];
  float4 vfr[NH][2];
  int cur_h = -1;
  for (int item = warp; item < H * n_mt; item += NW) {
    const int h = item % H, m0 = (item / H) * 16;
    if (HOIST && h != cur_h) {
      cur_h = h;
#pragma unroll
      for (int nt = 0; nt < NH; ++nt) {
        load_row8(kfr[nt], kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + h * 32, t);
        vfr[nt][0] = ldg4(kv + (int64_t)min(nt * 8 + t, S - 1) * 2 * E + E + h * 32 + 4 * g);
        vfr[nt][1] = ldg4(kv + (int64_t)min(nt * 8 + t + 4, S - 1) * 2 * E + E + h * 32 + 4 * g);
      }
    }
    const int r0 = m0 + g, r1 = r0 + 8;
    const bool ok0 = r0 < L, ok1 = r1 < L;
    RowFrag fq;   // q scaled before the product, as torch does
    load_rowfrag(fq, q + (int64_t)min(r0, L - 1) * E + h * 32, q + (int64_t)min(r1, L - 1) * E + h * 32, ok0, ok1, a.scale, t);
    float acc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
      if (nt * 8 < S) {
        if (HOIST) mma_rows_nt_frag(acc[nt], fq, kfr[nt < NH ? nt : 0]);
        else mma_rows_nt(acc[nt], fq, kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + h * 32, t);
      }
    }
    // bias, key padding, row maxima (rows g and g+8; a row's scores sit in the four lanes of a quad)
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      if (a.qs) {
        const float2 b0 = *reinterpret_cast<const float2*>(bt + (m0 + g) * PITCH + c);
        const float2 b1 = *reinterpret_cast<const float2*>(bt + (m0 + g + 8) * PITCH + c);
        acc[nt][0] += b0.x; acc[nt][1] += b0.y; acc[nt][2] += b1.x; acc[nt][3] += b1.y;
      }
      if (c >= S) acc[nt][0] = acc[nt][2] = -INFINITY;
      if (c + 1 >= S) acc[nt][1] = acc[nt][3] = -INFINITY;
      mx0 = fmaxf(mx0, fmaxf(acc[nt][0], acc[nt][1]));
      mx1 = fmaxf(mx1, fmaxf(acc[nt][2], acc[nt][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      acc[nt][0] = expf(acc[nt][0] - mx0); acc[nt][1] = expf(acc[nt][1] - mx0);
      acc[nt][2] = expf(acc[nt][2] - mx1); acc[nt][3] = expf(acc[nt][3] - mx1);
      s0 += acc[nt][0] + acc[nt][1];
      s1 += acc[nt][2] + acc[nt][3];
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
    const float i0 = 1.f / s0, i1 = 1.f / s1;
    __syncwarp();   // the previous item's output product has finished reading the tile
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      *reinterpret_cast<float2*>(ps + g * PITCH + c) = make_float2(acc[nt][0] * i0, acc[nt][1] * i0);
      *reinterpret_cast<float2*>(ps + (g + 8) * PITCH + c) = make_float2(acc[nt][2] * i1, acc[nt][3] * i1);
    }
    __syncwarp();
    // save P (before dropout) for the backward — with the dropout decision in its sign: +P kept, -P dropped (P >= 0, -0 counts
    // as dropped), so the backward regenerates no Philox mask — and apply the mask to the tile in place
    const int rows_valid = min(16, L - m0);
    const int ldp = a.ldp > 0 ? a.ldp : S;   // row stride of P = the index space of the Philox draws (a multiple of 4 under tf32)
    const int64_t base0 = (((int64_t)b * H + h) * L + m0) * ldp;
    if (a.P != nullptr || a.p_drop > 0.f) {
      const bool vec = (ldp & 3) == 0 && (reinterpret_cast<uintptr_t>(a.P) & 15) == 0;   // rows start on 16-byte boundaries
      for_each_with_dropout(
          rows_valid, S, base0, ldp, a.p_drop, a.seed, a.offset, a.rng_stream, lane,
          [&](int r, int s, const float(&f)[4]) {
            float* pp = ps + r * PITCH + s;
            if (vec) {   // (s % 4 == 0 then, and the tile pitch keeps rows 16-byte aligned)
              const float4 pv = *reinterpret_cast<const float4*>(pp);
              if (a.P != nullptr)
                *reinterpret_cast<float4*>(a.P + base0 + (int64_t)r * ldp + s) =
                    make_float4(f[0] > 0.f ? pv.x : -pv.x, f[1] > 0.f ? pv.y : -pv.y, f[2] > 0.f ? pv.z : -pv.z, f[3] > 0.f ? pv.w : -pv.w);
              *reinterpret_cast<float4*>(pp) = make_float4(pv.x * f[0], pv.y * f[1], pv.z * f[2], pv.w * f[3]);
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float pj = pp[j];
                if (a.P != nullptr) a.P[base0 + (int64_t)r * ldp + s + j] = f[j] > 0.f ? pj : -pj;
                pp[j] = pj * f[j];
              }
            }
          },
          [&](int r, int s, float f) {
            const float pj = ps[r * PITCH + s];
            if (a.P != nullptr) a.P[base0 + (int64_t)r * ldp + s] = f > 0.f ? pj : -pj;
            ps[r * PITCH + s] = pj * f;
          });
      __syncwarp();
    }
    // O = Pd V
    float o[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    if (HOIST) {
#pragma unroll
      for (int kstep = 0; kstep < NH; ++kstep)
        if (kstep * 8 < S) mma_tile_x_frag<PITCH>(o, ps, kstep, vfr[kstep][0], vfr[kstep][1], g, t);
    } else {
#pragma unroll 4
      for (int kstep = 0; kstep < NT; ++kstep) {
        if (kstep * 8 >= S) break;
        const float* v0 = kv + (int64_t)min(kstep * 8 + t, S - 1) * 2 * E + E + h * 32;       // (P is zero past the last key)
        const float* v1 = kv + (int64_t)min(kstep * 8 + t + 4, S - 1) * 2 * E + E + h * 32;
        mma_tile_x<PITCH>(o, ps, kstep, v0, v1, g, t);
      }
    }
    float* orow = a.O + (int64_t)b * L * E + h * 32;
    store_rows_x(o, orow + (int64_t)min(r0, L - 1) * E, orow + (int64_t)min(r1, L - 1) * E, ok0, ok1, 1.f, t);
  }
}

// --------------------------------------------------------------------------------------------------------------------------
// Backward.  Per item: dPd = dO V^T (accumulators); P and the dropout decisions arrive in ONE value per element, written by the
// forward: x = +P where the element was kept, -P where it was dropped (P >= 0, so the sign is free; -0 counts as dropped) — no
// Philox evaluation here.
// dS = P ⊙ (dP − rowsum(dP ⊙ P)) stays in the accumulator registers; the bias gradients are taken from it in place.  The tile then
// holds Pd for dV = Pd^T dO, is overwritten with dS for dK = dS^T Qs (A read transposed from the tile) and dQ = scale · dS K (A
// read canonically).  A warp always works on the same head (NW % H == 0), so dK / dV of a head with several query tiles (a2v: 10
// tiles of 16 queries, 16 keys) and the column sums for the in_proj bias gradients accumulate in registers across the warp's items
// and meet the other warps of the head in shared memory once, after the last item.
template <int NT, int NW, bool ACC>
__global__ void __launch_bounds__(NW * 32, NW == 4 ? 3 : 2) attn_bwd_mma_kernel(AttnBwdArgs a) {
  constexpr int PITCH = NT * 8 + 4;
  extern __shared__ __align__(16) float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int L = a.L, S = a.S, E = a.E, H = a.H;
  const int n_mt = (L + 15) >> 4, Lp = n_mt * 16;
  const bool prior = a.qs != nullptr;
  float* bt = sm;                                           // [Lp][PITCH] tanh(qs_l + ks_s) (NOT scaled here)
  float* tiles = sm + (prior ? Lp * PITCH : 0);
  float* ps = tiles + warp * 16 * PITCH;                    // ±P, then Pd, then dS
  float* accs = tiles + NW * 16 * PITCH;
  float* dqs_acc = accs;                                    // [Lp]
  float* dks_acc = dqs_acc + Lp;                            // [NT*8]
  float* col_acc = dks_acc + NT * 8;                        // [3][E]: column sums of dq | dk | dv
  float* dscale_acc = col_acc + 3 * E;                      // [1] (+3 pad)
  float* dkv_acc = dscale_acc + 4;                          // ACC: [NW / H or 1][H][NT*8][64] (dK | dV per key)
  // NW % H == 0: a warp stays on head warp % H, and the NW / H warps of a head each own one slot of dkv_acc (plain stores, summed
  // at the end); otherwise one slot, met by shared-memory atomics
  const bool own_slot = ACC && (NW % H) == 0;
  const int n_slot = own_slot ? NW / H : 1;
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float* dO = a.dO + (int64_t)b * L * E;
  const float bscale = prior ? a.bias_scale[0] : 0.f;
  const float inv_keep = a.p_drop > 0.f ? 1.f / (1.f - a.p_drop) : 1.f;
  prefetch_l2(kv, (int64_t)S * 2 * E * 4, tid, NW * 32);
  prefetch_l2(dO, (int64_t)L * E * 4, tid, NW * 32);
  prefetch_l2(q, (int64_t)L * E * 4, tid, NW * 32);
  prefetch_l2(a.P + (int64_t)b * H * L * (a.ldp > 0 ? a.ldp : S), (int64_t)H * L * (a.ldp > 0 ? a.ldp : S) * 4, tid, NW * 32);
  {
    const int nacc = Lp + NT * 8 + 3 * E + 4 + (ACC ? n_slot * H * NT * 8 * 64 : 0);
    for (int i = tid; i < nacc; i += NW * 32) accs[i] = 0.f;
    if (prior) {
      const float* qs = a.qs + (int64_t)b * L;
      const float* ks = a.ks + (int64_t)b * S;
      for (int i = tid; i < Lp * NT * 8; i += NW * 32) {
        const int l = i / (NT * 8), s = i - l * (NT * 8);
        bt[l * PITCH + s] = (l < L && s < S) ? tanh_fast(qs[l] + ks[s]) : 0.f;
      }
    }
    __syncthreads();
  }
  // per-warp accumulators over the warp's items (one head): column sums for the bias gradients, dK / dV of the (single) key tile
  float bk[8], bv[8], bq[8], dka[4][4], dva[4][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) bk[i] = bv[i] = bq[i] = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) dka[i][0] = dka[i][1] = dka[i][2] = dka[i][3] = dva[i][0] = dva[i][1] = dva[i][2] = dva[i][3] = 0.f;
  constexpr bool HOIST = false;   // (hoisting the head's V / K fragments as the forward does costs spills here: 217 -> 229 us measured)
  constexpr int NH = HOIST ? NT : 1;
  float vfr[NH][8];
  float4 kfr[NH][2];
  int cur_h = -1;
  auto flush = [&]() {   // this warp's sums of head cur_h -> the CTA's accumulators
    if (cur_h < 0) return;
    if (ACC) {
      float* d0 = dkv_acc + (((int64_t)(own_slot ? warp / H : 0) * H + cur_h) * NT * 8 + g) * 64 + 8 * t;
      float* d1 = d0 + 8 * 64;
      if (own_slot) {   // (flushed once: the head never changes)
        *reinterpret_cast<float4*>(d0) = make_float4(dka[0][0], dka[1][0], dka[2][0], dka[3][0]);
        *reinterpret_cast<float4*>(d0 + 4) = make_float4(dka[0][1], dka[1][1], dka[2][1], dka[3][1]);
        *reinterpret_cast<float4*>(d0 + 32) = make_float4(dva[0][0], dva[1][0], dva[2][0], dva[3][0]);
        *reinterpret_cast<float4*>(d0 + 36) = make_float4(dva[0][1], dva[1][1], dva[2][1], dva[3][1]);
        *reinterpret_cast<float4*>(d1) = make_float4(dka[0][2], dka[1][2], dka[2][2], dka[3][2]);
        *reinterpret_cast<float4*>(d1 + 4) = make_float4(dka[0][3], dka[1][3], dka[2][3], dka[3][3]);
        *reinterpret_cast<float4*>(d1 + 32) = make_float4(dva[0][2], dva[1][2], dva[2][2], dva[3][2]);
        *reinterpret_cast<float4*>(d1 + 36) = make_float4(dva[0][3], dva[1][3], dva[2][3], dva[3][3]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (!own_slot) {
          atomicAdd(d0 + i, dka[i][0]); atomicAdd(d0 + 4 + i, dka[i][1]); atomicAdd(d0 + 32 + i, dva[i][0]); atomicAdd(d0 + 36 + i, dva[i][1]);
          atomicAdd(d1 + i, dka[i][2]); atomicAdd(d1 + 4 + i, dka[i][3]); atomicAdd(d1 + 32 + i, dva[i][2]); atomicAdd(d1 + 36 + i, dva[i][3]);
        }
        dka[i][0] = dka[i][1] = dka[i][2] = dka[i][3] = dva[i][0] = dva[i][1] = dva[i][2] = dva[i][3] = 0.f;
      }
    }
    if (a.dbq != nullptr) {   // in_proj bias gradients: sum the rows over g, one shared-memory add per column
#pragma unroll
      for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int o2 = 4; o2 <= 16; o2 <<= 1) {
          bq[i] += __shfl_xor_sync(0xffffffffu, bq[i], o2);
          bk[i] += __shfl_xor_sync(0xffffffffu, bk[i], o2);
          bv[i] += __shfl_xor_sync(0xffffffffu, bv[i], o2);
        }
      }
      if (g == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          atomicAdd(col_acc + cur_h * 32 + 8 * t + i, bq[i]);
          atomicAdd(col_acc + E + cur_h * 32 + 8 * t + i, bk[i]);
          atomicAdd(col_acc + 2 * E + cur_h * 32 + 8 * t + i, bv[i]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) bk[i] = bv[i] = bq[i] = 0.f;
  };
  for (int item = warp; item < H * n_mt; item += NW) {
    const int h = item % H, m0 = (item / H) * 16;
    if (h != cur_h) {
      flush();
      cur_h = h;
      if (HOIST) {
#pragma unroll
        for (int nt = 0; nt < NH; ++nt) {
          load_row8(vfr[nt], kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + E + h * 32, t);
          kfr[nt][0] = ldg4(kv + (int64_t)min(nt * 8 + t, S - 1) * 2 * E + h * 32 + 4 * g);
          kfr[nt][1] = ldg4(kv + (int64_t)min(nt * 8 + t + 4, S - 1) * 2 * E + h * 32 + 4 * g);
        }
      }
    }
    const int r0 = m0 + g, r1 = r0 + 8;
    const bool ok0 = r0 < L, ok1 = r1 < L;
    const int rows_valid = min(16, L - m0);
    const int ldp = a.ldp > 0 ? a.ldp : S;
    const int64_t base0 = (((int64_t)b * H + h) * L + m0) * ldp;
    // ---- ±P (sign = dropout decision, written by the forward) starts its way into the tile: asynchronous copies that land while
    //      the dPd products below are issued; rows / keys past the end are zero ----
    __syncwarp();   // the previous item's products have finished reading the tile
    {
      const float* pg = a.P + base0;
      if ((ldp & 3) == 0 && (reinterpret_cast<uintptr_t>(a.P) & 15) == 0) {
        for (int i = lane; i < 16 * NT * 2; i += 32) {   // 16-byte pieces; the last piece of a row may straddle S (padding is not P)
          const int r = i / (NT * 2), s4 = (i - r * (NT * 2)) * 4;
          float* dst = ps + r * PITCH + s4;
          const float* src = pg + (int64_t)r * ldp + s4;
          if (r < rows_valid && s4 + 3 < S) {
            cp_async16(dst, src);
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (r < rows_valid && s4 + j < S) cp_async4(dst + j, src + j);
              else dst[j] = 0.f;
            }
          }
        }
      } else {
        for (int i = lane; i < 16 * NT * 8; i += 32) {
          const int r = i / (NT * 8), s1 = i - r * (NT * 8);
          if (r < rows_valid && s1 < S) cp_async4(ps + r * PITCH + s1, pg + (int64_t)r * ldp + s1);
          else ps[r * PITCH + s1] = 0.f;
        }
      }
    }
    // ---- dPd = dO V^T ----
    float acc[NT][4];
    {
      RowFrag fo;
      load_rowfrag(fo, dO + (int64_t)min(r0, L - 1) * E + h * 32, dO + (int64_t)min(r1, L - 1) * E + h * 32, ok0, ok1, 1.f, t);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
        if (nt * 8 < S) {
          if (HOIST) mma_rows_nt_frag(acc[nt], fo, vfr[nt < NH ? nt : 0]);
          else mma_rows_nt(acc[nt], fo, kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + E + h * 32, t);
        }
      }
    }
    cp_async_wait_all();
    __syncwarp();
    // ---- dS at the accumulator positions (kept in acc); bias gradients; Pd into the tile ----
    float rd0 = 0.f, rd1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      const float2 x0 = *reinterpret_cast<const float2*>(ps + g * PITCH + c), x1 = *reinterpret_cast<const float2*>(ps + (g + 8) * PITCH + c);
      // dP = dPd ⊙ mask (0 where dropped, 1/keep where kept)
      acc[nt][0] *= signbit(x0.x) ? 0.f : inv_keep; acc[nt][1] *= signbit(x0.y) ? 0.f : inv_keep;
      acc[nt][2] *= signbit(x1.x) ? 0.f : inv_keep; acc[nt][3] *= signbit(x1.y) ? 0.f : inv_keep;
      rd0 = fmaf(acc[nt][0], fabsf(x0.x), fmaf(acc[nt][1], fabsf(x0.y), rd0));
      rd1 = fmaf(acc[nt][2], fabsf(x1.x), fmaf(acc[nt][3], fabsf(x1.y), rd1));
    }
    rd0 += __shfl_xor_sync(0xffffffffu, rd0, 1); rd0 += __shfl_xor_sync(0xffffffffu, rd0, 2);
    rd1 += __shfl_xor_sync(0xffffffffu, rd1, 1); rd1 += __shfl_xor_sync(0xffffffffu, rd1, 2);
    float dq0 = 0.f, dq1 = 0.f, dsc = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      const float2 x0 = *reinterpret_cast<const float2*>(ps + g * PITCH + c), x1 = *reinterpret_cast<const float2*>(ps + (g + 8) * PITCH + c);
      const float p00 = fabsf(x0.x), p01 = fabsf(x0.y), p10 = fabsf(x1.x), p11 = fabsf(x1.y);
      acc[nt][0] = p00 * (acc[nt][0] - rd0); acc[nt][1] = p01 * (acc[nt][1] - rd0);   // dS = P ⊙ (dP − rowdot)
      acc[nt][2] = p10 * (acc[nt][2] - rd1); acc[nt][3] = p11 * (acc[nt][3] - rd1);
      if (prior) {
        const float2 t0 = *reinterpret_cast<const float2*>(bt + (m0 + g) * PITCH + c), t1 = *reinterpret_cast<const float2*>(bt + (m0 + g + 8) * PITCH + c);
        dsc = fmaf(acc[nt][0], t0.x, fmaf(acc[nt][1], t0.y, fmaf(acc[nt][2], t1.x, fmaf(acc[nt][3], t1.y, dsc))));
        const float u00 = acc[nt][0] * (1.f - t0.x * t0.x), u01 = acc[nt][1] * (1.f - t0.y * t0.y);
        const float u10 = acc[nt][2] * (1.f - t1.x * t1.x), u11 = acc[nt][3] * (1.f - t1.y * t1.y);
        dq0 += u00 + u01;
        dq1 += u10 + u11;
        float k0 = u00 + u10, k1 = u01 + u11;                                          // over this lane's two rows; then over g
        k0 += __shfl_xor_sync(0xffffffffu, k0, 4); k0 += __shfl_xor_sync(0xffffffffu, k0, 8); k0 += __shfl_xor_sync(0xffffffffu, k0, 16);
        k1 += __shfl_xor_sync(0xffffffffu, k1, 4); k1 += __shfl_xor_sync(0xffffffffu, k1, 8); k1 += __shfl_xor_sync(0xffffffffu, k1, 16);
        if (g == 0 && nt * 8 < S) {
          atomicAdd(dks_acc + c, k0 * bscale);
          atomicAdd(dks_acc + c + 1, k1 * bscale);
        }
      }
      // Pd replaces ±P (same positions, same lane)
      *reinterpret_cast<float2*>(ps + g * PITCH + c) = make_float2(signbit(x0.x) ? 0.f : p00 * inv_keep, signbit(x0.y) ? 0.f : p01 * inv_keep);
      *reinterpret_cast<float2*>(ps + (g + 8) * PITCH + c) = make_float2(signbit(x1.x) ? 0.f : p10 * inv_keep, signbit(x1.y) ? 0.f : p11 * inv_keep);
    }
    if (prior) {
      dq0 += __shfl_xor_sync(0xffffffffu, dq0, 1); dq0 += __shfl_xor_sync(0xffffffffu, dq0, 2);
      dq1 += __shfl_xor_sync(0xffffffffu, dq1, 1); dq1 += __shfl_xor_sync(0xffffffffu, dq1, 2);
      if (t == 0) {
        atomicAdd(dqs_acc + m0 + g, dq0 * bscale);
        atomicAdd(dqs_acc + m0 + g + 8, dq1 * bscale);
      }
#pragma unroll
      for (int o2 = 16; o2 > 0; o2 >>= 1) dsc += __shfl_xor_sync(0xffffffffu, dsc, o2);
      if (lane == 0) atomicAdd(dscale_acc, dsc);
    }
    __syncwarp();
    const int n_km = (S + 15) >> 4;
    // ---- dV = Pd^T dO: rows = keys, reduction over the tile's 16 queries (two k-steps), N-permuted columns ----
    {
      float4 on[2][2];   // B operand: dO rows of queries 8 k2 + t (+4), physical columns 4g .. 4g+3
#pragma unroll
      for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = m0 + 8 * k2 + t + 4 * j;
          const float4 x = ldg4(dO + (int64_t)min(r, L - 1) * E + h * 32 + 4 * g);
          const float s1 = r < L ? 1.f : 0.f;
          on[k2][j] = make_float4(x.x * s1, x.y * s1, x.z * s1, x.w * s1);
        }
      for (int km = 0; km < n_km; ++km) {
        float dv[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f;
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          // A[key][query] read transposed from the tile: a0 (key g, query t)  a1 (key g+8, t)  a2 (g, t+4)  a3 (g+8, t+4)
          const int qa = 8 * k2 + t, ka = 16 * km + g;
          const uint32_t p0 = fu(ps[qa * PITCH + ka]), p1 = fu(ps[qa * PITCH + ka + 8]), p2 = fu(ps[(qa + 4) * PITCH + ka]),
                         p3 = fu(ps[(qa + 4) * PITCH + ka + 8]);
          const float4 o0 = on[k2][0], o1 = on[k2][1];
          mma_tf32(dv[0], p0, p1, p2, p3, fu(o0.x), fu(o1.x)); mma_tf32(dv[1], p0, p1, p2, p3, fu(o0.y), fu(o1.y));
          mma_tf32(dv[2], p0, p1, p2, p3, fu(o0.z), fu(o1.z)); mma_tf32(dv[3], p0, p1, p2, p3, fu(o0.w), fu(o1.w));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // (rows past the last key hold exact zeros: Pd is zero there)
          bv[i] += dv[i][0] + dv[i][2];
          bv[4 + i] += dv[i][1] + dv[i][3];
        }
        if (ACC) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { dva[i][0] += dv[i][0]; dva[i][1] += dv[i][1]; dva[i][2] += dv[i][2]; dva[i][3] += dv[i][3]; }
        } else {
          const int k0r = 16 * km + g, k1r = k0r + 8;
          float* out = a.dkv + (int64_t)b * S * 2 * E + E + h * 32;
          store_rows_x(dv, out + (int64_t)min(k0r, S - 1) * 2 * E, out + (int64_t)min(k1r, S - 1) * 2 * E, k0r < S, k1r < S, 1.f, t);
        }
      }
    }
    // ---- dS replaces Pd in the tile ----
    __syncwarp();
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      *reinterpret_cast<float2*>(ps + g * PITCH + c) = make_float2(acc[nt][0], acc[nt][1]);
      *reinterpret_cast<float2*>(ps + (g + 8) * PITCH + c) = make_float2(acc[nt][2], acc[nt][3]);
    }
    __syncwarp();
    // ---- dK = dS^T Qs ----
    {
      float4 qn[2][2];
#pragma unroll
      for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = m0 + 8 * k2 + t + 4 * j;
          const float4 y = ldg4(q + (int64_t)min(r, L - 1) * E + h * 32 + 4 * g);
          const float s2 = r < L ? a.scale : 0.f;
          qn[k2][j] = make_float4(y.x * s2, y.y * s2, y.z * s2, y.w * s2);
        }
      for (int km = 0; km < n_km; ++km) {
        float dk[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f;
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          const int qa = 8 * k2 + t, ka = 16 * km + g;
          const uint32_t s0 = fu(ps[qa * PITCH + ka]), s1 = fu(ps[qa * PITCH + ka + 8]), s2 = fu(ps[(qa + 4) * PITCH + ka]),
                         s3 = fu(ps[(qa + 4) * PITCH + ka + 8]);
          const float4 q0 = qn[k2][0], q1 = qn[k2][1];
          mma_tf32(dk[0], s0, s1, s2, s3, fu(q0.x), fu(q1.x)); mma_tf32(dk[1], s0, s1, s2, s3, fu(q0.y), fu(q1.y));
          mma_tf32(dk[2], s0, s1, s2, s3, fu(q0.z), fu(q1.z)); mma_tf32(dk[3], s0, s1, s2, s3, fu(q0.w), fu(q1.w));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          bk[i] += dk[i][0] + dk[i][2];
          bk[4 + i] += dk[i][1] + dk[i][3];
        }
        if (ACC) {
#pragma unroll
          for (int i = 0; i < 4; ++i) { dka[i][0] += dk[i][0]; dka[i][1] += dk[i][1]; dka[i][2] += dk[i][2]; dka[i][3] += dk[i][3]; }
        } else {
          const int k0r = 16 * km + g, k1r = k0r + 8;
          float* out = a.dkv + (int64_t)b * S * 2 * E + h * 32;
          store_rows_x(dk, out + (int64_t)min(k0r, S - 1) * 2 * E, out + (int64_t)min(k1r, S - 1) * 2 * E, k0r < S, k1r < S, 1.f, t);
        }
      }
    }
    // ---- dQ = scale · dS K ----
    {
      float dq[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
      if (HOIST) {
#pragma unroll
        for (int kstep = 0; kstep < NH; ++kstep)
          if (kstep * 8 < S) mma_tile_x_frag<PITCH>(dq, ps, kstep, kfr[kstep][0], kfr[kstep][1], g, t);
      } else {
#pragma unroll 4
        for (int kstep = 0; kstep < NT; ++kstep) {
          if (kstep * 8 >= S) break;
          const float* k0 = kv + (int64_t)min(kstep * 8 + t, S - 1) * 2 * E + h * 32;      // (dS is zero past the last key)
          const float* k1 = kv + (int64_t)min(kstep * 8 + t + 4, S - 1) * 2 * E + h * 32;
          mma_tile_x<PITCH>(dq, ps, kstep, k0, k1, g, t);
        }
      }
      float* out = a.dq + (int64_t)b * L * E + h * 32;
      store_rows_x(dq, out + (int64_t)min(r0, L - 1) * E, out + (int64_t)min(r1, L - 1) * E, ok0, ok1, a.scale, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) {   // (rows past the last query are exact zeros: dS is zero there)
        bq[i] += (dq[i][0] + dq[i][2]) * a.scale;
        bq[4 + i] += (dq[i][1] + dq[i][3]) * a.scale;
      }
    }
  }
  flush();
  __syncthreads();
  if (ACC) {   // dK | dV of the clip, summed over its query tiles
    for (int i = tid; i < H * S * 16; i += NW * 32) {   // float4 pieces: [h][key][16 pieces of 4]
      const int piece = i & 15, key = (i >> 4) % S, h = (i >> 4) / S;
      float4 v = *reinterpret_cast<const float4*>(dkv_acc + ((int64_t)h * NT * 8 + key) * 64 + 4 * piece);
      for (int sl = 1; sl < n_slot; ++sl) {
        const float4 u = *reinterpret_cast<const float4*>(dkv_acc + (((int64_t)sl * H + h) * NT * 8 + key) * 64 + 4 * piece);
        v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
      }
      float* dst = a.dkv + ((int64_t)b * S + key) * 2 * E + (piece >= 8 ? E : 0) + h * 32 + 4 * (piece & 7);
      *reinterpret_cast<float4*>(dst) = v;
    }
  }
  if (prior) {
    for (int l = tid; l < L; l += NW * 32) a.dqs[(int64_t)b * L + l] = dqs_acc[l];
    for (int s = tid; s < S; s += NW * 32) a.dks[(int64_t)b * S + s] = dks_acc[s];
    if (tid == 0) atomicAdd(a.dbias_scale, dscale_acc[0]);
  }
  if (a.dbq != nullptr) {
    for (int i = tid; i < 3 * E; i += NW * 32) {
      const int which = i / E, c = i - which * E;
      atomicAdd(which == 0 ? a.dbq + c : a.dbkv + (which - 1) * E + c, col_acc[i]);
    }
  }
}

template <int NT, int NW>
size_t fwd_smem(int L, bool prior) {
  const int pitch = NT * 8 + 4, Lp = ((L + 15) / 16) * 16;
  return sizeof(float) * ((prior ? (size_t)Lp * pitch : 0) + (size_t)NW * 16 * pitch);
}
template <int NT, int NW, bool ACC>
size_t bwd_smem(int L, int E, int H, bool prior) {
  const int pitch = NT * 8 + 4, Lp = ((L + 15) / 16) * 16;
  return sizeof(float) * ((prior ? (size_t)Lp * pitch : 0) + (size_t)NW * 16 * pitch + Lp + NT * 8 + 3 * E + 4 +
                          (ACC ? (size_t)(NW % H == 0 ? NW / H : 1) * H * NT * 8 * 64 : 0));
}

template <typename K>
int set_smem(K kernel, size_t bytes, size_t*) {
  return ensure_dyn_smem(kernel, bytes);
}

}  // namespace

// Shapes the tensor-core tiles cover: head width 32, E = 32·H <= 128 wide accumulators are not assumed — only dh == 32 — and
// either (L <= 16, S <= 160) [video queries audio] or (L <= 160, S <= 16) [audio queries video].
bool attn_mma_ok(int L, int S, int E, int H) {
  if (H <= 0 || E != 32 * H || E % 4 != 0) return false;
  return (L <= 16 && S <= 160 && S > 16) || (L <= 160 && S <= 16);
}

int launch_attn_fwd_mma(cudaStream_t st, const AttnArgs& a, int B) {
  const bool prior = a.qs != nullptr;
  if (a.S > 16) {
    static size_t conf = 0;
    const size_t smem = fwd_smem<20, 4>(a.L, prior);
    XF_TRY(set_smem(attn_fwd_mma_kernel<20, 4>, smem, &conf));
    attn_fwd_mma_kernel<20, 4><<<B, 128, smem, st>>>(a);
  } else {
    static size_t conf = 0;
    const size_t smem = fwd_smem<2, 8>(a.L, prior);
    XF_TRY(set_smem(attn_fwd_mma_kernel<2, 8>, smem, &conf));
    attn_fwd_mma_kernel<2, 8><<<B, 256, smem, st>>>(a);
  }
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

int launch_attn_bwd_mma(cudaStream_t st, const AttnBwdArgs& a, int B) {
  const bool prior = a.qs != nullptr;
  if (a.S > 16) {
    static size_t conf = 0;
    const size_t smem = bwd_smem<20, 4, false>(a.L, a.E, a.H, prior);
    XF_TRY(set_smem(attn_bwd_mma_kernel<20, 4, false>, smem, &conf));
    attn_bwd_mma_kernel<20, 4, false><<<B, 128, smem, st>>>(a);
  } else {
    static size_t conf = 0;
    const size_t smem = bwd_smem<2, 8, true>(a.L, a.E, a.H, prior);
    XF_TRY(set_smem(attn_bwd_mma_kernel<2, 8, true>, smem, &conf));
    attn_bwd_mma_kernel<2, 8, true><<<B, 256, smem, st>>>(a);
  }
  count_launch();
  XF_LAUNCH_CHECK();
  return XATTN_OK;
}

}  // namespace xf
