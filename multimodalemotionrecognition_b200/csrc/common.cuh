// Shared helpers for the xattn fusion kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/xattn_fusion.h"

namespace xf {

// ---- error plumbing (thread-local text behind xattn_last_error) -------------------------------
void set_error(const char* fmt, ...);
const char* last_error();

#define XF_FAIL(code, ...)     \
  do {                         \
    xf::set_error(__VA_ARGS__); \
    return (code);             \
  } while (0)

#define XF_CUDA(expr)                                                                         \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      xf::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return XATTN_ERR_CUDA;                                                                  \
    }                                                                                         \
  } while (0)

#define XF_TRY(expr)            \
  do {                          \
    int _s = (expr);            \
    if (_s != XATTN_OK) return _s; \
  } while (0)

#define XF_LAUNCH_CHECK()                                                                  \
  do {                                                                                     \
    cudaError_t _e = cudaGetLastError();                                                   \
    if (_e != cudaSuccess) {                                                               \
      xf::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return XATTN_ERR_CUDA;                                                               \
    }                                                                                      \
  } while (0)

int check_device_sm100();           // XATTN_ERR_ARCH unless the current device is sm_100
int validate_config(const xattn_config* cfg);
int num_sms();
// A cached, timing-disabled event of the current device (slot < 4), created once under a lock; for cross-stream edges the
// library inserts itself (record on one stream, cudaStreamWaitEvent on another: the wait captures the record made before it).
int device_event(int slot, cudaEvent_t* out);

// Fork / join of a per-device side stream for the fp32 / tf32 kernel family (backward_f32 runs the latency-bound per-clip prior
// backward beside HBM-bound tensor-core weight-gradient GEMMs that nothing downstream reads).  The side stream has the highest priority the device
// offers, so its kernels' CTAs are placed first when both streams have work pending.  begin(): work enqueued on side() afterwards
// starts after what is already on `st`; join(): `st` waits for the side work.  Stream and events are created once per device under
// a lock and shared by every caller, so begin() .. join() holds a mutex.  Disabled (begin() leaves active() false, side() == st) by
// XATTN_NO_FORK or while per-kernel timing brackets are on.
struct SideForkRes;
class SideFork {
 public:
  explicit SideFork(cudaStream_t st) : st_(st), side_(st) {}   // until begin() succeeds side() IS the caller's stream
  ~SideFork() { join(); }
  SideFork(const SideFork&) = delete;
  SideFork& operator=(const SideFork&) = delete;
  int begin();
  int join();
  bool active() const { return res_ != nullptr; }
  cudaStream_t side() const { return side_; }

 private:
  SideForkRes* res_ = nullptr;
  cudaStream_t st_ = nullptr, side_ = nullptr;
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of a kernel ON ONE DEVICE: remembered per (function, device) — not per
// process — so a process that drives several GPUs configures each of them.  No-op up to 48 KB (the default limit).
int ensure_dyn_smem_impl(const void* func, size_t bytes);
template <typename K>
inline int ensure_dyn_smem(K kernel, size_t bytes) { return ensure_dyn_smem_impl(reinterpret_cast<const void*>(kernel), bytes); }

// Count of kernels this library launched since process start (bench.py's gpu_launches claim).
extern unsigned long long g_launch_count;
inline void count_launch(int n = 1) { g_launch_count += (unsigned long long)n; }

// Optional per-stage CUDA-event timing of the bf16 forward (xattn_stage_timing): bench.py's live per-kernel durations.
enum { ST_GEMM_V0 = 0, ST_GEMM_A0, ST_VIDEO_PRE, ST_PRIOR, ST_V2A, ST_VIDEO_UPD, ST_A2V, ST_HEAD, ST_COUNT };
extern bool g_stage_timing;
// precision = XATTN_PREC_TF32: the Linear helpers of gemm_simt.cuh route to the tensor-core GEMMs (gemm_tf32.cu) while this is
// set (by forward_f32 / backward_f32 for the duration of one call); a failed launch is remembered in g_tc_status.
void cvt_f32_to_bf16(cudaStream_t st, const float* x, void* y_bf16, int64_t n);   // n % 4 == 0 or any n; 16-byte aligned
extern thread_local bool g_tc_gemm;
extern thread_local int g_tc_status;
// Kernels of the bf16 forward chain are launched with programmatic stream serialisation (each one starts with
// griddepcontrol.wait, tc.cuh) unless XATTN_NO_PDL is set or per-kernel timing brackets are on.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
void stage_mark(int stage, int end, cudaStream_t st);   // records an event on st (no-op unless timing is on)
// NVTX range on the host thread that issues the launches (the reference has no tracing hooks; nsys / ncu --nvtx show
// these).  nvtx3 is header-only and a no-op unless a tool is attached.  Defined in common.cu.
void nvtx_push(const char* name);
void nvtx_pop();
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtx_push(name); }
  ~NvtxRange() { nvtx_pop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};
const char* stage_name(int i);
struct StageScope {
  int s;
  cudaStream_t st;
  StageScope(int stage, cudaStream_t stream) : s(stage), st(stream) {
    nvtx_push(stage_name(s));
    if (g_stage_timing) stage_mark(s, 0, st);
  }
  ~StageScope() {
    if (g_stage_timing) stage_mark(s, 1, st);
    nvtx_pop();
  }
};
int stage_timing_enable(int on);
int stage_timing_read(double* ms, int64_t* launches, int cap);

// ---- parameter layout ----------------------------------------------------------------------------
struct ParamLayout {
  int64_t off[XATTN_PARAM_ID_MAX];
  int64_t size[XATTN_PARAM_ID_MAX];
  int64_t total;
  int count;   // ids in use: XP_COUNT + 24 per transformer-pooler layer beyond the first
};
inline int pool_layers(const xattn_config& c) { return c.pool_layers > 0 ? c.pool_layers : 1; }
// Row stride (floats) of the saved attention probabilities P [B,H,L,stride]: S, rounded up to 4 under precision = TF32 so that the
// tensor-core attention kernels move whole 16-byte pieces of a row (S = 149 audio keys -> 152).  The fp32 SIMT kernels always
// use S (inside the same buffer when they stand in under TF32).
inline int p_row_stride(const xattn_config& c, int S) { return c.precision == XATTN_PREC_TF32 ? (S + 3) & ~3 : S; }
void compute_param_layout(const xattn_config& c, ParamLayout& L);

// ---- fp32 workspace layout (activations saved for backward; floats) -------------------------------
struct WsF32 {
  // token tensors
  int64_t v0, aseq, a0, q1, kv1, p1, o1, s1, mean1, rstd1, v1, q2, kv2, p2, o2, s2, mean2, rstd2, a1;
  // per-clip vectors
  int64_t pooled0, phid, prior, qv, ka, qa, kvs, pooled, hhid, gate, fusedg, dp1, dp2;
  // backward scratch (only laid out when training / save_for_backward)
  int64_t d_pooled, d_hh, d_z, d_fused, d_a0, d_br2, d_o2, d_q2, d_kv2, d_v1, d_v0, d_br1, d_o1, d_q1, d_kv1, d_aseq;
  int64_t d_qv, d_ka, d_qa, d_kvs, d_prior, d_phid;
  int64_t wbf;    // bf16 scratch of the TF32 precision (see compute_ws_f32)
  int64_t poolv, poola;   // pool_mode != MEAN: pooled tokens [B,E] each
  int64_t tx_x, tx_h, tx_q, tx_kv, tx_f;   // pool_mode = TRANSFORMER: encoder-layer activations of one side at a time
  // ... with training / save_for_backward: what the encoder layers keep for their backward, per side s (0 video, 1 audio) and
  // layer l, at tx_save[s] + l * tx_layer_floats[s] (carved up by tx_saved() below); the output of the last layer at tx_out[s];
  // and the backward's scratch (token gradient, two [M,E] temporaries, the pooler's d_h rows)
  int64_t tx_save[2], tx_layer_floats[2], tx_out[2];
  int64_t txb_dx, txb_t1, txb_t2, txb_dh;
  int64_t total;  // floats
};
void compute_ws_f32(const xattn_config& c, WsF32& w);
// Saved activations of encoder layer `layer` of side `side` (M = B·L rows, F = 4E, PH = pool_heads):
// x_in [M,E] layer input, q [M,E] | kv [M,2E] projected LN1(x_in), P [B,PH,L,L] attention probabilities (sign = dropped),
// o [M,E] attention output before out_proj, x_mid [M,E] after the first residual, f_pre / f_act [M,F] around the GELU.
struct TxSaved {
  float *x_in, *q, *kv, *o, *x_mid, *f_pre, *f_act, *P;
};
inline int64_t tx_layer_floats(const xattn_config& c, int L) {
  const int64_t ME = ((int64_t)c.B * L * c.E + 3) & ~(int64_t)3;
  return 14 * ME + (((int64_t)c.B * c.pool_heads * L * L + 3) & ~(int64_t)3);
}
inline TxSaved tx_saved(float* ws, const WsF32& w, const xattn_config& c, int side, int layer) {
  const int L = side == 0 ? c.T : c.Ta;
  const int64_t ME = ((int64_t)c.B * L * c.E + 3) & ~(int64_t)3;
  float* p = ws + w.tx_save[side] + layer * w.tx_layer_floats[side];
  TxSaved s;
  s.x_in = p; s.q = p + ME; s.kv = p + 2 * ME; s.o = p + 4 * ME; s.x_mid = p + 5 * ME; s.f_pre = p + 6 * ME; s.f_act = p + 10 * ME;
  s.P = p + 14 * ME;
  return s;
}

// ---- device helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }

// Philox4x32-10 counter RNG: uniform in [0,1) for (seed, stream, index).  Used for every dropout so
// the backward kernels regenerate the forward's masks instead of storing them.
__device__ __forceinline__ uint4 philox4x32(uint64_t seed, uint64_t ctr_hi, uint64_t ctr_lo) {
  uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
  uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ float rng_uniform(uint64_t seed, uint64_t offset, uint32_t stream, uint64_t idx) {
  uint4 r = philox4x32(seed, (offset << 8) | stream, idx >> 2);
  uint32_t x = (idx & 3) == 0 ? r.x : (idx & 3) == 1 ? r.y : (idx & 3) == 2 ? r.z : r.w;
  return (float)(x >> 8) * (1.0f / 16777216.0f);
}
// multiplicative dropout factor: 0 or 1/keep
__device__ __forceinline__ float drop_factor(float p, uint64_t seed, uint64_t offset, uint32_t stream, uint64_t idx) {
  if (p <= 0.f) return 1.f;
  if (p >= 1.f) return 0.f;
  return rng_uniform(seed, offset, stream, idx) < p ? 0.f : 1.f / (1.f - p);
}
// Factors of the four elements idx0 .. idx0+3 (idx0 % 4 == 0) from ONE Philox evaluation — the same values drop_factor gives
// element by element (rng_uniform maps idx to counter idx >> 2, word idx & 3).
__device__ __forceinline__ void drop_factor4(float p, uint64_t seed, uint64_t offset, uint32_t stream, uint64_t idx0, float (&f)[4]) {
  if (p <= 0.f) { f[0] = f[1] = f[2] = f[3] = 1.f; return; }
  if (p >= 1.f) { f[0] = f[1] = f[2] = f[3] = 0.f; return; }
  const uint4 r = philox4x32(seed, (offset << 8) | stream, idx0 >> 2);
  const float keep = 1.f / (1.f - p), sc = 1.0f / 16777216.0f;
  f[0] = (float)(r.x >> 8) * sc < p ? 0.f : keep;
  f[1] = (float)(r.y >> 8) * sc < p ? 0.f : keep;
  f[2] = (float)(r.z >> 8) * sc < p ? 0.f : keep;
  f[3] = (float)(r.w >> 8) * sc < p ? 0.f : keep;
}
enum : uint32_t { RS_ATTN1 = 1, RS_ATTN2 = 2, RS_PATH1 = 3, RS_PATH2 = 4, RS_HEAD = 5, RS_PRIOR = 6, RS_POOLV = 7, RS_POOLA = 8,
                  RS_FLAT = 9,   // hidden-layer Dropout of the flat concat / gated heads
                  RS_TX_BASE = 16 };   // transformer pooler: RS_TX_BASE + (side·XATTN_MAX_POOL_LAYERS + layer)·4 + {attn, drop1, ffn, drop2}
__host__ __device__ inline uint32_t rs_tx(int side, int layer, int which) {
  return RS_TX_BASE + (uint32_t)((side * XATTN_MAX_POOL_LAYERS + layer) * 4 + which);
}

}  // namespace xf
