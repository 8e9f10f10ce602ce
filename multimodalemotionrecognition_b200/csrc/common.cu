#include <stdlib.h>
// Host-side plumbing: error text, config validation, parameter + workspace layouts.
#include "common.cuh"

#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <vector>

#include <nvtx3/nvToolsExt.h>

namespace xf {

static thread_local char g_err[512] = "";
unsigned long long g_launch_count = 0;

bool g_stage_timing = false;
thread_local bool g_tc_gemm = false;
thread_local int g_tc_status = 0;
bool pdl_enabled() {
  static const bool off = getenv("XATTN_NO_PDL") != nullptr;
  return !off && !g_stage_timing;
}
int ensure_dyn_smem_impl(const void* func, size_t bytes) {
  if (bytes <= 48 * 1024) return XATTN_OK;
  struct Rec { const void* func; size_t bytes[64]; };
  static std::vector<Rec> recs;
  static std::mutex mu;
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) XF_FAIL(XATTN_ERR_CUDA, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  Rec* r = nullptr;
  for (Rec& x : recs)
    if (x.func == func) { r = &x; break; }
  if (r == nullptr) {
    recs.push_back(Rec{func, {}});
    r = &recs.back();
  }
  if (bytes > r->bytes[dev]) {
    XF_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    r->bytes[dev] = bytes;
  }
  return XATTN_OK;
}

struct SideForkRes {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  std::mutex enqueue;
};
int SideFork::begin() {
  static SideForkRes res[64];
  static std::mutex mu;
  static const bool off = getenv("XATTN_NO_FORK") != nullptr;
  cudaStream_t st = st_;
  if (off || g_stage_timing || res_ != nullptr) return XATTN_OK;
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) XF_FAIL(XATTN_ERR_CUDA, "device index %d out of range", dev);
  SideForkRes* r = &res[dev];
  {
    std::lock_guard<std::mutex> lock(mu);
    if (r->side == nullptr) {
      int lo = 0, hi = 0;
      XF_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      XF_CUDA(cudaStreamCreateWithPriority(&r->side, cudaStreamNonBlocking, hi));
    }
    if (r->fork == nullptr) XF_CUDA(cudaEventCreateWithFlags(&r->fork, cudaEventDisableTiming));
    if (r->join == nullptr) XF_CUDA(cudaEventCreateWithFlags(&r->join, cudaEventDisableTiming));
  }
  r->enqueue.lock();
  cudaError_t e = cudaEventRecord(r->fork, st);
  if (e == cudaSuccess) e = cudaStreamWaitEvent(r->side, r->fork, 0);
  if (e != cudaSuccess) {
    r->enqueue.unlock();
    set_error("side-stream fork failed: %s", cudaGetErrorString(e));
    return XATTN_ERR_CUDA;
  }
  res_ = r;
  side_ = r->side;
  return XATTN_OK;
}
int SideFork::join() {
  if (res_ == nullptr) return XATTN_OK;
  SideForkRes* r = res_;
  res_ = nullptr;
  side_ = st_;
  cudaError_t e = cudaEventRecord(r->join, r->side);
  if (e == cudaSuccess) e = cudaStreamWaitEvent(st_, r->join, 0);
  r->enqueue.unlock();
  if (e != cudaSuccess) {
    set_error("side-stream join failed: %s", cudaGetErrorString(e));
    return XATTN_ERR_CUDA;
  }
  return XATTN_OK;
}

static std::vector<cudaEvent_t> g_stage_ev[ST_COUNT][2];
static const char* const kStageNames[ST_COUNT] = {"gemm_v0", "gemm_a0", "video_pre", "prior", "v2a_attn", "video_update", "a2v_attn", "head"};

static void stage_drop() {
  for (int s = 0; s < ST_COUNT; ++s)
    for (int e = 0; e < 2; ++e) {
      for (cudaEvent_t ev : g_stage_ev[s][e]) cudaEventDestroy(ev);
      g_stage_ev[s][e].clear();
    }
}
void stage_mark(int stage, int end, cudaStream_t st) {
  cudaEvent_t ev;
  if (cudaEventCreate(&ev) != cudaSuccess) return;
  cudaEventRecord(ev, st);
  g_stage_ev[stage][end].push_back(ev);
}
int stage_timing_enable(int on) {
  stage_drop();
  g_stage_timing = on != 0;
  return XATTN_OK;
}
const char* stage_name(int i) { return i >= 0 && i < ST_COUNT ? kStageNames[i] : ""; }
// Sums the elapsed time of every bracket recorded since the last enable/read; waits for the events.
int stage_timing_read(double* ms, int64_t* launches, int cap) {
  for (int s = 0; s < ST_COUNT && s < cap; ++s) {
    double tot = 0.0;
    const size_t n = g_stage_ev[s][1].size() < g_stage_ev[s][0].size() ? g_stage_ev[s][1].size() : g_stage_ev[s][0].size();
    for (size_t i = 0; i < n; ++i) {
      float t = 0.f;
      XF_CUDA(cudaEventSynchronize(g_stage_ev[s][1][i]));
      XF_CUDA(cudaEventElapsedTime(&t, g_stage_ev[s][0][i], g_stage_ev[s][1][i]));
      tot += t;
    }
    if (ms) ms[s] = tot;
    if (launches) launches[s] = (int64_t)n;
  }
  stage_drop();
  return ST_COUNT;
}

void nvtx_push(const char* name) { nvtxRangePushA(name); }
void nvtx_pop() { nvtxRangePop(); }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }

namespace {
__global__ void __launch_bounds__(256) cvt_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x * 8;
  for (int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < n; i += stride) {
    if (i + 8 <= n) {
      const float4 a = *reinterpret_cast<const float4*>(x + i), b = *reinterpret_cast<const float4*>(x + i + 4);
      __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w), p2 = __floats2bfloat162_rn(b.x, b.y),
                     p3 = __floats2bfloat162_rn(b.z, b.w);
      *reinterpret_cast<uint4*>(y + i) = make_uint4(*reinterpret_cast<uint32_t*>(&p0), *reinterpret_cast<uint32_t*>(&p1),
                                                    *reinterpret_cast<uint32_t*>(&p2), *reinterpret_cast<uint32_t*>(&p3));
    } else {
      for (int64_t k = i; k < n; ++k) y[k] = __float2bfloat16_rn(x[k]);
    }
  }
}
}  // namespace

void cvt_f32_to_bf16(cudaStream_t st, const float* x, void* y, int64_t n) {
  if (n <= 0) return;
  int64_t blocks = (n / 8 + 255) / 256 + 1;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cvt_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(x, reinterpret_cast<__nv_bfloat16*>(y), n);
  count_launch();
}

int check_device_sm100() {
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  static int cached_dev = -1, cached_major = 0;
  if (cached_dev != dev) {
    int major = 0;
    XF_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    cached_major = major;
    cached_dev = dev;
  }
  if (cached_major != 10) XF_FAIL(XATTN_ERR_ARCH, "device %d is sm_%d0; this library is built for sm_100a only (no fallback)", dev, cached_major);
  return XATTN_OK;
}

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    // experiments: cap the grid of every persistent kernel (SM partitioning between concurrent streams)
    if (const char* e = getenv("XATTN_MAX_CTAS")) {
      const int cap = atoi(e);
      if (cap > 0 && cap < n) n = cap;
    }
  }
  return n;
}

int device_event(int slot, cudaEvent_t* out) {
  static cudaEvent_t ev[64][4] = {};
  static std::mutex mu;
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || slot < 0 || slot >= 4) XF_FAIL(XATTN_ERR_CUDA, "device_event: device %d / slot %d out of range", dev, slot);
  std::lock_guard<std::mutex> lock(mu);
  if (ev[dev][slot] == nullptr) XF_CUDA(cudaEventCreateWithFlags(&ev[dev][slot], cudaEventDisableTiming));
  *out = ev[dev][slot];
  return XATTN_OK;
}

int validate_config(const xattn_config* c) {
  if (c == nullptr) XF_FAIL(XATTN_ERR_NULL, "cfg is NULL");
  if (c->B < 0 || c->T <= 0 || c->Ta <= 0 || c->Dv <= 0 || c->Da <= 0 || c->E <= 0 || c->H <= 0 || c->C <= 0)
    XF_FAIL(XATTN_ERR_SHAPE, "non-positive dimension in cfg (B=%d T=%d Ta=%d Dv=%d Da=%d E=%d H=%d C=%d)", c->B, c->T, c->Ta,
            c->Dv, c->Da, c->E, c->H, c->C);
  if (c->E % c->H != 0) XF_FAIL(XATTN_ERR_SHAPE, "embed_dim must be divisible by num_heads (E=%d H=%d)", c->E, c->H);
  if (c->E > 512) XF_FAIL(XATTN_ERR_SHAPE, "d_model %d > 512 is not supported", c->E);
  if (c->head_kind != XATTN_HEAD_CONCAT && c->head_kind != XATTN_HEAD_GATED)
    XF_FAIL(XATTN_ERR_SHAPE, "unknown head_kind %d", c->head_kind);
  if (c->head_kind == XATTN_HEAD_CONCAT && c->common_dim <= 0) XF_FAIL(XATTN_ERR_SHAPE, "common_dim must be positive");
  if (c->use_prior && (c->prior_dim <= 0 || c->prior_hidden <= 0))
    XF_FAIL(XATTN_ERR_SHAPE, "emotion prior needs positive prior_dim / prior_hidden");
  if (c->io_dtype != XATTN_DTYPE_F32 && c->io_dtype != XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "unknown io_dtype %d", c->io_dtype);
  if (c->pool_mode != XATTN_POOL_MEAN && c->pool_mode != XATTN_POOL_ATTN && c->pool_mode != XATTN_POOL_TRANSFORMER)
    XF_FAIL(XATTN_ERR_SHAPE, "unknown pool_mode %d", c->pool_mode);
  if (c->pool_mode == XATTN_POOL_TRANSFORMER) {
    if (c->pool_layers < 0 || c->pool_layers > XATTN_MAX_POOL_LAYERS)
      XF_FAIL(XATTN_ERR_SHAPE, "pool_layers=%d outside 1..%d", c->pool_layers, XATTN_MAX_POOL_LAYERS);
    if (c->pool_heads <= 0 || c->E % c->pool_heads != 0) XF_FAIL(XATTN_ERR_SHAPE, "pool_heads=%d must divide E=%d", c->pool_heads, c->E);
  }
  if (c->pool_mode == XATTN_POOL_TRANSFORMER && c->precision == XATTN_DTYPE_BF16)
    XF_FAIL(XATTN_ERR_SHAPE, "temporal_pooling=transformer runs in the fp32 kernel family (precision f32 / tf32)");
  if (c->precision != XATTN_DTYPE_F32 && c->precision != XATTN_DTYPE_BF16 && c->precision != XATTN_PREC_TF32)
    XF_FAIL(XATTN_ERR_DTYPE, "unknown precision %d", c->precision);
  if ((int64_t)c->B * c->Ta * c->Da >= ((int64_t)1 << 40)) XF_FAIL(XATTN_ERR_SHAPE, "input too large");
  return XATTN_OK;
}

static inline int64_t al4(int64_t n) { return (n + 3) & ~(int64_t)3; }

void compute_param_layout(const xattn_config& c, ParamLayout& L) {
  const int64_t E = c.E, P = c.prior_dim, Hd = c.prior_hidden;
  const bool gated = c.head_kind == XATTN_HEAD_GATED;
  const int64_t HW = gated ? E : c.common_dim;
  int64_t sz[XATTN_PARAM_ID_MAX];
  memset(sz, 0, sizeof(sz));
  L.count = XP_COUNT;
  sz[XP_V_IN_W] = E * c.Dv; sz[XP_V_IN_B] = E;
  sz[XP_A_IN_W] = E * E; sz[XP_A_IN_B] = E;
  sz[XP_SEQ_W] = E * c.Da; sz[XP_SEQ_B] = E;
  sz[XP_V2A_IN_W] = 3 * E * E; sz[XP_V2A_IN_B] = 3 * E; sz[XP_V2A_OUT_W] = E * E; sz[XP_V2A_OUT_B] = E;
  sz[XP_A2V_IN_W] = 3 * E * E; sz[XP_A2V_IN_B] = 3 * E; sz[XP_A2V_OUT_W] = E * E; sz[XP_A2V_OUT_B] = E;
  sz[XP_VN_W] = E; sz[XP_VN_B] = E; sz[XP_AN_W] = E; sz[XP_AN_B] = E;
  sz[XP_HEAD0_W] = HW * 2 * E; sz[XP_HEAD0_B] = HW;
  sz[XP_HEAD1_W] = gated ? HW : (int64_t)c.C * HW; sz[XP_HEAD1_B] = gated ? 1 : c.C;
  sz[XP_CLS_W] = gated ? (int64_t)c.C * E : 0; sz[XP_CLS_B] = gated ? c.C : 0;
  if (c.use_prior) {
    sz[XP_PRIOR0_W] = Hd * 2 * E; sz[XP_PRIOR0_B] = Hd; sz[XP_PRIOR3_W] = P * Hd; sz[XP_PRIOR3_B] = P;
    sz[XP_VQ_W] = sz[XP_AK_W] = sz[XP_AQ_W] = sz[XP_VK_W] = E + P;
    sz[XP_VQ_B] = sz[XP_AK_B] = sz[XP_AQ_B] = sz[XP_VK_B] = 1;
    sz[XP_BIAS_SCALE] = 1;
  }
  if (c.pool_mode == XATTN_POOL_TRANSFORMER) {
    const int64_t F = 4 * E;   // max(2E, int(4.0·E)), temporal.py:58
    const int layers = pool_layers(c);
    L.count = XP_COUNT + 2 * XATTN_TX_PARAMS_PER_LAYER * (layers - 1);
    for (int layer = 0; layer < layers; ++layer)
      for (int side = 0; side < 2; ++side) {
        const int base = XATTN_TX_PARAM_ID(side, layer, 0);
        sz[base + 0] = 3 * E * E; sz[base + 1] = 3 * E; sz[base + 2] = E * E; sz[base + 3] = E;
        sz[base + 4] = F * E; sz[base + 5] = F; sz[base + 6] = E * F; sz[base + 7] = E;
        sz[base + 8] = sz[base + 9] = sz[base + 10] = sz[base + 11] = E;
      }
  }
  if (c.pool_mode == XATTN_POOL_ATTN || c.pool_mode == XATTN_POOL_TRANSFORMER) {
    const int64_t H2 = E / 2 > 0 ? E / 2 : 1;
    for (int base : {(int)XP_VPOOL_LN_W, (int)XP_APOOL_LN_W}) {
      sz[base + 0] = E; sz[base + 1] = E; sz[base + 2] = H2 * E; sz[base + 3] = H2; sz[base + 4] = H2; sz[base + 5] = 1;
    }
  }
  int64_t off = 0;
  for (int i = 0; i < XATTN_PARAM_ID_MAX; ++i) {
    L.off[i] = off;
    L.size[i] = sz[i];
    off += al4(sz[i]);
  }
  L.total = off;
}

void compute_ws_f32(const xattn_config& c, WsF32& w) {
  const int64_t B = c.B, T = c.T, Ta = c.Ta, E = c.E, H = c.H, P = c.prior_dim > 0 ? c.prior_dim : 1,
                Hd = c.prior_hidden > 0 ? c.prior_hidden : 1;
  const int64_t HW = c.head_kind == XATTN_HEAD_GATED ? E : c.common_dim;
  const bool bwd = c.training || c.save_for_backward;
  int64_t off = 0;
  auto take = [&](int64_t n) {
    int64_t o = off;
    off += al4(n);
    return o;
  };
  w.v0 = take(B * T * E); w.aseq = take(B * Ta * E); w.a0 = take(B * Ta * E);
  w.q1 = take(B * T * E); w.kv1 = take(B * Ta * 2 * E);
  w.p1 = bwd ? take(B * H * T * p_row_stride(c, c.Ta)) : 0;
  w.o1 = take(B * T * E); w.s1 = take(B * T * E); w.mean1 = take(B * T); w.rstd1 = take(B * T); w.v1 = take(B * T * E);
  w.q2 = take(B * Ta * E); w.kv2 = take(B * T * 2 * E);
  w.p2 = bwd ? take(B * H * Ta * p_row_stride(c, c.T)) : 0;
  w.o2 = take(B * Ta * E); w.s2 = take(B * Ta * E); w.mean2 = take(B * Ta); w.rstd2 = take(B * Ta); w.a1 = take(B * Ta * E);
  w.pooled0 = take(B * 2 * E); w.phid = take(B * Hd); w.prior = take(B * P);
  w.qv = take(B * T); w.ka = take(B * Ta); w.qa = take(B * Ta); w.kvs = take(B * T);
  w.pooled = take(B * 2 * E); w.hhid = take(B * HW); w.gate = take(B); w.fusedg = take(B * E);
  w.dp1 = take(B); w.dp2 = take(B);
  if (bwd) {
    w.d_pooled = take(B * 2 * E); w.d_hh = take(B * HW); w.d_z = take(B); w.d_fused = take(B * E);
    w.d_a0 = take(B * Ta * E); w.d_br2 = take(B * Ta * E); w.d_o2 = take(B * Ta * E); w.d_q2 = take(B * Ta * E);
    w.d_kv2 = take(B * T * 2 * E); w.d_v1 = take(B * T * E); w.d_v0 = take(B * T * E); w.d_br1 = take(B * T * E);
    w.d_o1 = take(B * T * E); w.d_q1 = take(B * T * E); w.d_kv1 = take(B * Ta * 2 * E); w.d_aseq = take(B * Ta * E);
    w.d_qv = take(B * T); w.d_ka = take(B * Ta); w.d_qa = take(B * Ta); w.d_kvs = take(B * T);
    w.d_prior = take(B * P); w.d_phid = take(B * Hd);
  }
  // precision = TF32 with bf16 embeddings: bf16 copies of the two first-layer weights (forward), and of d_aseq / d_v0 for the
  // first-layer weight gradients (backward; those live in d_o2 / d_o1, dead by then)
  w.wbf = take((E * (int64_t)c.Da + E * (int64_t)c.Dv + 1) / 2 + 8);
  w.poolv = take(c.pool_mode != XATTN_POOL_MEAN ? B * E : 0);
  w.poola = take(c.pool_mode != XATTN_POOL_MEAN ? B * E : 0);
  {
    const int64_t M = c.pool_mode == XATTN_POOL_TRANSFORMER ? B * (T > Ta ? T : Ta) : 0;   // rows of the longer side
    w.tx_x = take(M * E); w.tx_h = take(M * E); w.tx_q = take(M * E); w.tx_kv = take(M * 2 * E); w.tx_f = take(M * 4 * E);
    const bool txb = bwd && M > 0;
    for (int side = 0; side < 2; ++side) {
      const int L = side == 0 ? c.T : c.Ta;
      w.tx_layer_floats[side] = txb ? tx_layer_floats(c, L) : 0;
      w.tx_save[side] = take(w.tx_layer_floats[side] * pool_layers(c));
      w.tx_out[side] = take(txb ? B * L * E : 0);
    }
    w.txb_dx = take(txb ? M * E : 0); w.txb_t1 = take(txb ? M * E : 0); w.txb_t2 = take(txb ? M * E : 0);
    w.txb_dh = take(txb ? M * (E / 2 > 0 ? E / 2 : 1) : 0);
  }
  w.total = off;
}

}  // namespace xf
