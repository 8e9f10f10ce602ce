// xattn_forward_noalloc: the forward of the fusion head for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  Same kernels and results as xattn_forward; the only difference is where the
// scratch workspace lives: a file-scope __device__ array of XATTN_NOALLOC_MAX_WORKSPACE_BYTES instead of a caller buffer.
// No counterpart in the reference.
//
//  - every GPU that loads the library holds the array until the process exits;
//  - nothing in it has to start at zero: the forward writes every workspace region before it reads it (xattn_forward is
//    given torch.empty memory), so the previous call's contents are simply overwritten;
//  - every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) race.  Calls on
//    one stream are ordered by the stream and are safe back to back.
//  - no cudaMalloc / cudaFree / stream or event creation: the video branch stays on the caller's stream (fork = false).
#include <mutex>

#include "kernels_bf16.cuh"
#include "kernels_f32.cuh"

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[XATTN_NOALLOC_MAX_WORKSPACE_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
void* g_scratch_ptr[kMaxDevices] = {};   // device address of g_noalloc_scratch on each device, resolved once per device

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel can use: the device address comes
// from cudaGetSymbolAddress, asked for with the device current.
int scratch_address(void** out) {
  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) XF_FAIL(XATTN_ERR_CUDA, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    XF_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (p == nullptr || ((uintptr_t)p & 255) != 0) XF_FAIL(XATTN_ERR_CUDA, "static scratch resolved to a bad address %p", p);
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return XATTN_OK;
}

}  // namespace

using namespace xf;

extern "C" XATTN_API int64_t xattn_noalloc_max_workspace_bytes(void) { return (int64_t)XATTN_NOALLOC_MAX_WORKSPACE_BYTES; }

extern "C" XATTN_API int xattn_forward_noalloc(const xattn_config* cfg, const float* params, const void* packed, const void* video_emb,
                                               const void* audio_emb, float* logits, void* stream) {
  XF_TRY(validate_config(cfg));
  if (!params || !video_emb || !audio_emb || !logits) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (((uintptr_t)params & 15) || ((uintptr_t)video_emb & 15) || ((uintptr_t)audio_emb & 15) || ((uintptr_t)logits & 15))
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  if (cfg->training || cfg->save_for_backward)
    XF_FAIL(XATTN_ERR_SHAPE, "xattn_forward_noalloc is inference-only (training / save_for_backward need a caller workspace)");
  if (cfg->precision == XATTN_DTYPE_BF16 && !packed)
    XF_FAIL(XATTN_ERR_NULL, "precision=bf16 needs the packed weights (xattn_pack_weights)");
  XF_TRY(check_device_sm100());
  int64_t need = 0;
  XF_TRY(xattn_workspace_bytes(cfg, &need));
  if (need > (int64_t)XATTN_NOALLOC_MAX_WORKSPACE_BYTES)
    XF_FAIL(XATTN_ERR_WORKSPACE,
            "xattn_forward_noalloc: this config needs %lld bytes of scratch, the static array holds %lld (bf16 path: B <= 66 at "
            "T=16, Ta=149); use xattn_forward with a caller workspace",
            (long long)need, (long long)XATTN_NOALLOC_MAX_WORKSPACE_BYTES);
  if (cfg->B == 0) return XATTN_OK;
  void* ws = nullptr;
  XF_TRY(scratch_address(&ws));
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->precision == XATTN_DTYPE_BF16) return forward_bf16(*cfg, params, packed, video_emb, audio_emb, logits, ws, st, /*fork=*/false);
  return forward_f32(*cfg, params, video_emb, audio_emb, logits, ws, st);
}
