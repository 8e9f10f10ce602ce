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
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

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
// dropout factors of indices 4c .. 4c+3 come from ONE Philox evaluation (drop_factor4), so lanes walk aligned groups of four.
// fn(r, s, factor) is called for every element of the valid rows.
template <typename F>
__device__ __forceinline__ void for_each_with_dropout(int rows_valid, int S, int64_t base0, int64_t row_stride, float p, uint64_t seed,
                                                      uint64_t offset, uint32_t stream, int lane, F&& fn) {
  for (int r = 0; r < rows_valid; ++r) {
    const int64_t base = base0 + (int64_t)r * row_stride;
    const int rem = (int)(base & 3);
    const int ngroups = (S + rem + 3) >> 2;
    for (int grp = lane; grp < ngroups; grp += 32) {
      float f[4] = {1.f, 1.f, 1.f, 1.f};
      if (p > 0.f) drop_factor4(p, seed, offset, stream, (uint64_t)(base - rem) + 4u * (uint64_t)grp, f);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int s = 4 * grp + j - rem;
        if (s >= 0 && s < S) fn(r, s, f[j]);
      }
    }
  }
}

// --------------------------------------------------------------------------------------------------------------------------
template <int NT, int NW>
__global__ void __launch_bounds__(NW * 32) attn_fwd_mma_kernel(AttnArgs a) {
  constexpr int PITCH = NT * 8 + 4;   // (≡ 4 mod 32: the canonical A-fragment reads of a tile are conflict-free)
  extern __shared__ __align__(16) float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int L = a.L, S = a.S, E = a.E, H = a.H;
  const int n_mt = (L + 15) >> 4, Lp = n_mt * 16;
  float* bt = sm;                                           // [Lp][PITCH]: tanh(qs_l + ks_s)·bias_scale, shared by the heads
  float* ps = sm + (a.qs ? Lp * PITCH : 0) + warp * 16 * PITCH;
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  if (a.qs) {
    const float bscale = a.bias_scale[0];
    const float* qs = a.qs + (int64_t)b * L;
    const float* ks = a.ks + (int64_t)b * S;
    for (int i = tid; i < Lp * NT * 8; i += NW * 32) {
      const int l = i / (NT * 8), s = i - l * (NT * 8);
      bt[l * PITCH + s] = (l < L && s < S) ? tanhf(qs[l] + ks[s]) * bscale : 0.f;
    }
    __syncthreads();
  }
  for (int item = warp; item < H * n_mt; item += NW) {
    const int h = item % H, m0 = (item / H) * 16;
    const int r0 = m0 + g, r1 = r0 + 8;
    const bool ok0 = r0 < L, ok1 = r1 < L;
    RowFrag fq;   // q scaled before the product, as torch does
    load_rowfrag(fq, q + (int64_t)min(r0, L - 1) * E + h * 32, q + (int64_t)min(r1, L - 1) * E + h * 32, ok0, ok1, a.scale, t);
    float acc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
      if (nt * 8 < S) mma_rows_nt(acc[nt], fq, kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + h * 32, t);
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
    // save P (before dropout) for the backward, apply the dropout mask in place
    const int rows_valid = min(16, L - m0);
    const int64_t base0 = (((int64_t)b * H + h) * L + m0) * S;
    if (a.P != nullptr || a.p_drop > 0.f) {
      for_each_with_dropout(rows_valid, S, base0, S, a.p_drop, a.seed, a.offset, a.rng_stream, lane, [&](int r, int s, float f) {
        const float p = ps[r * PITCH + s];
        if (a.P != nullptr) a.P[base0 + (int64_t)r * S + s] = p;
        ps[r * PITCH + s] = p * f;
      });
      __syncwarp();
    }
    // O = Pd V
    float o[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
#pragma unroll 4
    for (int kstep = 0; kstep < NT; ++kstep) {
      if (kstep * 8 >= S) break;
      const float* v0 = kv + (int64_t)min(kstep * 8 + t, S - 1) * 2 * E + E + h * 32;       // (P is zero past the last key)
      const float* v1 = kv + (int64_t)min(kstep * 8 + t + 4, S - 1) * 2 * E + E + h * 32;
      mma_tile_x<PITCH>(o, ps, kstep, v0, v1, g, t);
    }
    float* orow = a.O + (int64_t)b * L * E + h * 32;
    store_rows_x(o, orow + (int64_t)min(r0, L - 1) * E, orow + (int64_t)min(r1, L - 1) * E, ok0, ok1, 1.f, t);
  }
}

// --------------------------------------------------------------------------------------------------------------------------
// Backward.  Per item: dPd = dO V^T (accumulators), P and the dropout factors through the two warp tiles (row-major pass),
// dS = P ⊙ (dP − rowsum(dP ⊙ P)) in registers, the bias gradients from dS at the same positions, then Pd and dS go back to the
// tiles and feed dV = Pd^T dO, dK = dS^T Qs (A read transposed from the tile), dQ = scale · dS K (A read canonically).
// ACC: dK / dV are summed over the query tiles of a head in shared memory (a2v: 10 tiles of 16 queries, 16 keys); otherwise the
// single query tile writes them directly (v2a).
template <int NT, int NW, bool ACC>
__global__ void __launch_bounds__(NW * 32) attn_bwd_mma_kernel(AttnBwdArgs a) {
  constexpr int PITCH = NT * 8 + 4;
  extern __shared__ __align__(16) float sm[];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int L = a.L, S = a.S, E = a.E, H = a.H;
  const int n_mt = (L + 15) >> 4, Lp = n_mt * 16;
  const bool prior = a.qs != nullptr;
  float* bt = sm;                                           // [Lp][PITCH] tanh(qs_l + ks_s) (NOT scaled here)
  float* tiles = sm + (prior ? Lp * PITCH : 0);
  float* ps = tiles + warp * 2 * 16 * PITCH;                // P, then Pd
  float* fs = ps + 16 * PITCH;                              // dropout factors, then dS
  float* accs = tiles + NW * 2 * 16 * PITCH;
  float* dqs_acc = accs;                                    // [Lp]
  float* dks_acc = dqs_acc + Lp;                            // [NT*8]
  float* col_acc = dks_acc + NT * 8;                        // [3][E]: column sums of dq | dk | dv
  float* dscale_acc = col_acc + 3 * E;                      // [1] (+3 pad)
  float* dkv_acc = dscale_acc + 4;                          // ACC: [H][NT*8][64] (dK | dV per key)
  const float* q = a.q + (int64_t)b * L * E;
  const float* kv = a.kv + (int64_t)b * S * 2 * E;
  const float* dO = a.dO + (int64_t)b * L * E;
  const float bscale = prior ? a.bias_scale[0] : 0.f;
  {
    const int nacc = Lp + NT * 8 + 3 * E + 4 + (ACC ? H * NT * 8 * 64 : 0);
    for (int i = tid; i < nacc; i += NW * 32) accs[i] = 0.f;
    if (prior) {
      const float* qs = a.qs + (int64_t)b * L;
      const float* ks = a.ks + (int64_t)b * S;
      for (int i = tid; i < Lp * NT * 8; i += NW * 32) {
        const int l = i / (NT * 8), s = i - l * (NT * 8);
        bt[l * PITCH + s] = (l < L && s < S) ? tanhf(qs[l] + ks[s]) : 0.f;
      }
    }
    __syncthreads();
  }
  for (int item = warp; item < H * n_mt; item += NW) {
    const int h = item % H, m0 = (item / H) * 16;
    const int r0 = m0 + g, r1 = r0 + 8;
    const bool ok0 = r0 < L, ok1 = r1 < L;
    const int rows_valid = min(16, L - m0);
    const int64_t base0 = (((int64_t)b * H + h) * L + m0) * S;
    // ---- dPd = dO V^T ----
    float acc[NT][4];
    {
      RowFrag fo;
      load_rowfrag(fo, dO + (int64_t)min(r0, L - 1) * E + h * 32, dO + (int64_t)min(r1, L - 1) * E + h * 32, ok0, ok1, 1.f, t);
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
        if (nt * 8 < S) mma_rows_nt(acc[nt], fo, kv + (int64_t)min(nt * 8 + g, S - 1) * 2 * E + E + h * 32, t);
      }
    }
    // ---- P and the dropout factors into the tiles (row-major pass); rows / keys past the end are zero ----
    __syncwarp();   // the previous item's products have finished reading the tiles
    for (int i = lane; i < 16 * NT * 8; i += 32) {
      const int r = i / (NT * 8), s = i - r * (NT * 8);
      if (r >= rows_valid || s >= S) ps[r * PITCH + s] = 0.f, fs[r * PITCH + s] = 0.f;
    }
    for_each_with_dropout(rows_valid, S, base0, S, a.p_drop, a.seed, a.offset, a.rng_stream, lane, [&](int r, int s, float f) {
      ps[r * PITCH + s] = a.P[base0 + (int64_t)r * S + s];
      fs[r * PITCH + s] = f;
    });
    __syncwarp();
    // ---- dS at the accumulator positions; bias gradients ----
    float rd0 = 0.f, rd1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      const float2 p0 = *reinterpret_cast<const float2*>(ps + g * PITCH + c), p1 = *reinterpret_cast<const float2*>(ps + (g + 8) * PITCH + c);
      const float2 f0 = *reinterpret_cast<const float2*>(fs + g * PITCH + c), f1 = *reinterpret_cast<const float2*>(fs + (g + 8) * PITCH + c);
      acc[nt][0] *= f0.x; acc[nt][1] *= f0.y; acc[nt][2] *= f1.x; acc[nt][3] *= f1.y;       // dP = dPd ⊙ mask
      rd0 = fmaf(acc[nt][0], p0.x, fmaf(acc[nt][1], p0.y, rd0));
      rd1 = fmaf(acc[nt][2], p1.x, fmaf(acc[nt][3], p1.y, rd1));
    }
    rd0 += __shfl_xor_sync(0xffffffffu, rd0, 1); rd0 += __shfl_xor_sync(0xffffffffu, rd0, 2);
    rd1 += __shfl_xor_sync(0xffffffffu, rd1, 1); rd1 += __shfl_xor_sync(0xffffffffu, rd1, 2);
    float dq0 = 0.f, dq1 = 0.f, dsc = 0.f;
    __syncwarp();   // every lane has read its P / factor values: the tiles may be overwritten
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      const float2 p0 = *reinterpret_cast<const float2*>(ps + g * PITCH + c), p1 = *reinterpret_cast<const float2*>(ps + (g + 8) * PITCH + c);
      const float2 f0 = *reinterpret_cast<const float2*>(fs + g * PITCH + c), f1 = *reinterpret_cast<const float2*>(fs + (g + 8) * PITCH + c);
      const float d00 = p0.x * (acc[nt][0] - rd0), d01 = p0.y * (acc[nt][1] - rd0);   // dS = P ⊙ (dP − rowdot)
      const float d10 = p1.x * (acc[nt][2] - rd1), d11 = p1.y * (acc[nt][3] - rd1);
      if (prior) {
        const float2 t0 = *reinterpret_cast<const float2*>(bt + (m0 + g) * PITCH + c), t1 = *reinterpret_cast<const float2*>(bt + (m0 + g + 8) * PITCH + c);
        dsc = fmaf(d00, t0.x, fmaf(d01, t0.y, fmaf(d10, t1.x, fmaf(d11, t1.y, dsc))));
        const float u00 = d00 * (1.f - t0.x * t0.x), u01 = d01 * (1.f - t0.y * t0.y);
        const float u10 = d10 * (1.f - t1.x * t1.x), u11 = d11 * (1.f - t1.y * t1.y);
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
      // Pd and dS replace P and the factors (same positions, same lane)
      *reinterpret_cast<float2*>(ps + g * PITCH + c) = make_float2(p0.x * f0.x, p0.y * f0.y);
      *reinterpret_cast<float2*>(ps + (g + 8) * PITCH + c) = make_float2(p1.x * f1.x, p1.y * f1.y);
      *reinterpret_cast<float2*>(fs + g * PITCH + c) = make_float2(d00, d01);
      *reinterpret_cast<float2*>(fs + (g + 8) * PITCH + c) = make_float2(d10, d11);
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
    // ---- dV = Pd^T dO and dK = dS^T Qs: rows = keys, reduction over the tile's 16 queries (two k-steps), N-permuted columns ----
    float bk[8], bv[8], bq[8];   // column sums of this item's dk / dv / dq rows (physical columns 8t .. 8t+7), for the bias gradients
#pragma unroll
    for (int i = 0; i < 8; ++i) bk[i] = bv[i] = bq[i] = 0.f;
    {
      float4 on[2][2], qn[2][2];   // B operands: dO / scaled q rows of queries 8 k2 + t (+4), physical columns 4g .. 4g+3
#pragma unroll
      for (int k2 = 0; k2 < 2; ++k2)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int r = m0 + 8 * k2 + t + 4 * j;
          const bool ok = r < L;
          const int rr = min(r, L - 1);
          float4 x = ldg4(dO + (int64_t)rr * E + h * 32 + 4 * g), y = ldg4(q + (int64_t)rr * E + h * 32 + 4 * g);
          const float s1 = ok ? 1.f : 0.f, s2 = ok ? a.scale : 0.f;
          on[k2][j] = make_float4(x.x * s1, x.y * s1, x.z * s1, x.w * s1);
          qn[k2][j] = make_float4(y.x * s2, y.y * s2, y.z * s2, y.w * s2);
        }
      const int n_km = (S + 15) >> 4;
      for (int km = 0; km < n_km; ++km) {
        float dv[4][4], dk[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f;
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          // A[key][query] read transposed from the tiles: a0 (key g, query t)  a1 (key g+8, t)  a2 (g, t+4)  a3 (g+8, t+4)
          const int qa = 8 * k2 + t, ka = 16 * km + g;
          const uint32_t p0 = fu(ps[qa * PITCH + ka]), p1 = fu(ps[qa * PITCH + ka + 8]), p2 = fu(ps[(qa + 4) * PITCH + ka]),
                         p3 = fu(ps[(qa + 4) * PITCH + ka + 8]);
          const uint32_t s0 = fu(fs[qa * PITCH + ka]), s1 = fu(fs[qa * PITCH + ka + 8]), s2 = fu(fs[(qa + 4) * PITCH + ka]),
                         s3 = fu(fs[(qa + 4) * PITCH + ka + 8]);
          const float4 o0 = on[k2][0], o1 = on[k2][1], q0 = qn[k2][0], q1 = qn[k2][1];
          mma_tf32(dv[0], p0, p1, p2, p3, fu(o0.x), fu(o1.x)); mma_tf32(dv[1], p0, p1, p2, p3, fu(o0.y), fu(o1.y));
          mma_tf32(dv[2], p0, p1, p2, p3, fu(o0.z), fu(o1.z)); mma_tf32(dv[3], p0, p1, p2, p3, fu(o0.w), fu(o1.w));
          mma_tf32(dk[0], s0, s1, s2, s3, fu(q0.x), fu(q1.x)); mma_tf32(dk[1], s0, s1, s2, s3, fu(q0.y), fu(q1.y));
          mma_tf32(dk[2], s0, s1, s2, s3, fu(q0.z), fu(q1.z)); mma_tf32(dk[3], s0, s1, s2, s3, fu(q0.w), fu(q1.w));
        }
        const int k0r = 16 * km + g, k1r = k0r + 8;
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // (rows past the last key hold exact zeros: Pd and dS are zero there)
          bk[i] += dk[i][0] + dk[i][2]; bk[4 + i] += dk[i][1] + dk[i][3];
          bv[i] += dv[i][0] + dv[i][2]; bv[4 + i] += dv[i][1] + dv[i][3];
        }
        if (ACC) {
          float* d0 = dkv_acc + ((int64_t)h * NT * 8 + k0r) * 64 + 8 * t;
          float* d1 = dkv_acc + ((int64_t)h * NT * 8 + k1r) * 64 + 8 * t;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            atomicAdd(d0 + i, dk[i][0]); atomicAdd(d0 + 4 + i, dk[i][1]); atomicAdd(d0 + 32 + i, dv[i][0]); atomicAdd(d0 + 36 + i, dv[i][1]);
            atomicAdd(d1 + i, dk[i][2]); atomicAdd(d1 + 4 + i, dk[i][3]); atomicAdd(d1 + 32 + i, dv[i][2]); atomicAdd(d1 + 36 + i, dv[i][3]);
          }
        } else {
          float* out = a.dkv + (int64_t)b * S * 2 * E + h * 32;
          store_rows_x(dk, out + (int64_t)min(k0r, S - 1) * 2 * E, out + (int64_t)min(k1r, S - 1) * 2 * E, k0r < S, k1r < S, 1.f, t);
          store_rows_x(dv, out + (int64_t)min(k0r, S - 1) * 2 * E + E, out + (int64_t)min(k1r, S - 1) * 2 * E + E, k0r < S, k1r < S, 1.f, t);
        }
      }
    }
    // ---- dQ = scale · dS K ----
    {
      float dq[4][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
#pragma unroll 4
      for (int kstep = 0; kstep < NT; ++kstep) {
        if (kstep * 8 >= S) break;
        const float* k0 = kv + (int64_t)min(kstep * 8 + t, S - 1) * 2 * E + h * 32;      // (dS is zero past the last key)
        const float* k1 = kv + (int64_t)min(kstep * 8 + t + 4, S - 1) * 2 * E + h * 32;
        mma_tile_x<PITCH>(dq, fs, kstep, k0, k1, g, t);
      }
      float* out = a.dq + (int64_t)b * L * E + h * 32;
      store_rows_x(dq, out + (int64_t)min(r0, L - 1) * E, out + (int64_t)min(r1, L - 1) * E, ok0, ok1, a.scale, t);
#pragma unroll
      for (int i = 0; i < 4; ++i) {   // (rows past the last query are exact zeros: dS is zero there)
        bq[i] += (dq[i][0] + dq[i][2]) * a.scale;
        bq[4 + i] += (dq[i][1] + dq[i][3]) * a.scale;
      }
    }
    if (a.dbq != nullptr) {   // in_proj bias gradients: sum the rows over g, one shared-memory add per column and item
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
          atomicAdd(col_acc + h * 32 + 8 * t + i, bq[i]);
          atomicAdd(col_acc + E + h * 32 + 8 * t + i, bk[i]);
          atomicAdd(col_acc + 2 * E + h * 32 + 8 * t + i, bv[i]);
        }
      }
    }
  }
  __syncthreads();
  if (ACC) {   // dK | dV of the clip, summed over its query tiles
    for (int i = tid; i < H * S * 16; i += NW * 32) {   // float4 pieces: [h][key][16 pieces of 4]
      const int piece = i & 15, key = (i >> 4) % S, h = (i >> 4) / S;
      const float4 v = *reinterpret_cast<const float4*>(dkv_acc + ((int64_t)h * NT * 8 + key) * 64 + 4 * piece);
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
  return sizeof(float) * ((prior ? (size_t)Lp * pitch : 0) + (size_t)NW * 2 * 16 * pitch + Lp + NT * 8 + 3 * E + 4 +
                          (ACC ? (size_t)H * NT * 8 * 64 : 0));
}

template <typename K>
int set_smem(K kernel, size_t bytes, size_t* configured) {
  if (bytes > 48 * 1024 && bytes > *configured) {
    XF_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    *configured = bytes;
  }
  return XATTN_OK;
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
