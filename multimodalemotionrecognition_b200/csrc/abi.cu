// extern "C" entry points declared in include/xattn_fusion.h.
#include "kernels_bf16.cuh"
#include "kernels_f32.cuh"
#include "gemm_tc.cuh"

namespace xf {
int emit_embeddings_bf16(cudaStream_t st, const float* x, void* out, const int64_t* dst, int n_clips, int rows_per_clip, int D, int R);
int ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int B, int C, float label_smoothing, float grad_scale,
                    float* loss_out, float* dlogits, cudaStream_t st);
int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, float wd,
              int step, int decoupled, cudaStream_t st);
}  // namespace xf

using namespace xf;

static inline bool misaligned(const void* p) { return ((uintptr_t)p & 15) != 0; }

extern "C" XATTN_API int xattn_abi_version(void) { return XATTN_ABI_VERSION; }
extern "C" XATTN_API const char* xattn_last_error(void) { return xf::last_error(); }
// Not part of the reference-facing surface: kernel launches issued by this library so far (bench.py's
// gpu_launches evidence).
extern "C" XATTN_API unsigned long long xattn_launch_count(void) { return xf::g_launch_count; }

extern "C" XATTN_API int xattn_stage_timing(int32_t enable) { return xf::stage_timing_enable(enable); }
extern "C" XATTN_API int xattn_stage_timing_read(double* ms_total, int64_t* launches, int32_t cap) {
  return xf::stage_timing_read(ms_total, launches, cap);
}
extern "C" XATTN_API const char* xattn_stage_name(int32_t i) { return xf::stage_name(i); }

extern "C" XATTN_API int xattn_param_count(const xattn_config* cfg) {
  if (cfg == nullptr || cfg->pool_mode != XATTN_POOL_TRANSFORMER || cfg->pool_layers <= 1) return XP_COUNT;
  const int layers = cfg->pool_layers > XATTN_MAX_POOL_LAYERS ? XATTN_MAX_POOL_LAYERS : cfg->pool_layers;
  return XP_COUNT + 2 * XATTN_TX_PARAMS_PER_LAYER * (layers - 1);
}

extern "C" XATTN_API int xattn_param_layout(const xattn_config* cfg, int64_t* offsets, int64_t* sizes, int64_t* total_floats) {
  XF_TRY(validate_config(cfg));
  ParamLayout L;
  compute_param_layout(*cfg, L);
  for (int i = 0; i < L.count; ++i) {
    if (offsets) offsets[i] = L.off[i];
    if (sizes) sizes[i] = L.size[i];
  }
  if (total_floats) *total_floats = L.total;
  return XATTN_OK;
}

extern "C" XATTN_API int xattn_workspace_bytes(const xattn_config* cfg, int64_t* bytes) {
  XF_TRY(validate_config(cfg));
  if (!bytes) XF_FAIL(XATTN_ERR_NULL, "bytes is NULL");
  if (cfg->precision == XATTN_DTYPE_BF16) return workspace_bytes_bf16(*cfg, bytes);
  WsF32 w;
  compute_ws_f32(*cfg, w);
  *bytes = w.total * (int64_t)sizeof(float) + 16;
  return XATTN_OK;
}

extern "C" XATTN_API int xattn_packed_bytes(const xattn_config* cfg, int64_t* bytes) {
  XF_TRY(validate_config(cfg));
  if (!bytes) XF_FAIL(XATTN_ERR_NULL, "bytes is NULL");
  return packed_bytes_bf16(*cfg, bytes);
}

extern "C" XATTN_API int xattn_pack_weights(const xattn_config* cfg, const float* params, void* packed, void* stream) {
  XF_TRY(validate_config(cfg));
  if (!params || !packed) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(params) || misaligned(packed)) XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  return pack_weights_bf16(*cfg, params, packed, (cudaStream_t)stream);
}

extern "C" XATTN_API int xattn_forward(const xattn_config* cfg, const float* params, const void* packed, const void* video_emb,
                             const void* audio_emb, float* logits, void* workspace, int64_t workspace_bytes, void* stream) {
  XF_TRY(validate_config(cfg));
  if (!params || !video_emb || !audio_emb || !logits || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(params) || misaligned(video_emb) || misaligned(audio_emb) || misaligned(workspace) || misaligned(logits))
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  int64_t need = 0;
  XF_TRY(xattn_workspace_bytes(cfg, &need));
  if (workspace_bytes < need)
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)need);
  if (cfg->B == 0) return XATTN_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->precision == XATTN_DTYPE_BF16) {
    if (!packed) XF_FAIL(XATTN_ERR_NULL, "precision=bf16 needs the packed weights (xattn_pack_weights)");
    if (cfg->training || cfg->save_for_backward)
      XF_FAIL(XATTN_ERR_DTYPE, "the bf16 tcgen05 path is inference-only; train with precision=f32");
    return forward_bf16(*cfg, params, packed, video_emb, audio_emb, logits, workspace, st);
  }
  return forward_f32(*cfg, params, video_emb, audio_emb, logits, workspace, st);
}

extern "C" XATTN_API int xattn_backward(const xattn_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                              const float* dlogits, float* grads, void* workspace, int64_t workspace_bytes, void* stream) {
  XF_TRY(validate_config(cfg));
  if (!params || !video_emb || !audio_emb || !dlogits || !grads || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(params) || misaligned(video_emb) || misaligned(audio_emb) || misaligned(workspace) || misaligned(grads))
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  if (cfg->precision == XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "xattn_backward runs with precision=f32 or tf32");
  if (!cfg->training && !cfg->save_for_backward)
    XF_FAIL(XATTN_ERR_WORKSPACE, "forward was not run with training or save_for_backward set: no saved activations");
  XF_TRY(check_device_sm100());
  int64_t need = 0;
  XF_TRY(xattn_workspace_bytes(cfg, &need));
  if (workspace_bytes < need)
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)need);
  if (cfg->B == 0) {
    ParamLayout L;
    compute_param_layout(*cfg, L);
    XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * L.total, (cudaStream_t)stream));
    return XATTN_OK;
  }
  return backward_f32(*cfg, params, video_emb, audio_emb, dlogits, grads, workspace, (cudaStream_t)stream);
}

extern "C" XATTN_API int xattn_backward_ex(const xattn_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                                           const float* dlogits, float* grads, void* workspace, int64_t workspace_bytes, void* stream,
                                           void* tail_stream, int64_t* tail_offset_floats, float* d_video_emb, float* d_audio_emb) {
  XF_TRY(validate_config(cfg));
  if (tail_offset_floats) {
    ParamLayout L;
    compute_param_layout(*cfg, L);
    *tail_offset_floats = L.off[XP_V2A_IN_W];
  }
  if (!params || !video_emb || !audio_emb || !dlogits || !grads || !workspace) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(params) || misaligned(video_emb) || misaligned(audio_emb) || misaligned(workspace) || misaligned(grads) ||
      misaligned(d_video_emb) || misaligned(d_audio_emb))
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  if (cfg->precision == XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "xattn_backward runs with precision=f32 or tf32");
  if (!cfg->training && !cfg->save_for_backward)
    XF_FAIL(XATTN_ERR_WORKSPACE, "forward was not run with training or save_for_backward set: no saved activations");
  XF_TRY(check_device_sm100());
  int64_t need = 0;
  XF_TRY(xattn_workspace_bytes(cfg, &need));
  if (workspace_bytes < need)
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes, (long long)need);
  cudaStream_t st = (cudaStream_t)stream;
  if (cfg->B == 0) {
    ParamLayout L;
    compute_param_layout(*cfg, L);
    XF_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * L.total, st));
    if (tail_stream) {
      cudaEvent_t ev = nullptr;
      XF_TRY(device_event(0, &ev));
      XF_CUDA(cudaEventRecord(ev, st));
      XF_CUDA(cudaStreamWaitEvent((cudaStream_t)tail_stream, ev, 0));
    }
    return XATTN_OK;
  }
  BackwardOptions opt;
  opt.tail_stream = (cudaStream_t)tail_stream;
  opt.d_video = d_video_emb;
  opt.d_audio = d_audio_emb;
  return backward_f32(*cfg, params, video_emb, audio_emb, dlogits, grads, workspace, st, opt);
}

extern "C" XATTN_API int xattn_ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t B, int32_t C, float label_smoothing,
                                     float grad_scale, float* loss_out, float* dlogits, void* stream) {
  if (!logits || !labels || !loss_out) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (B < 0 || C <= 0) XF_FAIL(XATTN_ERR_SHAPE, "bad CE shape B=%d C=%d", B, C);
  XF_TRY(check_device_sm100());
  return ce_loss_fwd_bwd(logits, labels, B, C, label_smoothing, grad_scale, loss_out, dlogits, (cudaStream_t)stream);
}

extern "C" XATTN_API int xattn_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr, float beta1,
                               float beta2, float eps, float weight_decay, int32_t step, int32_t decoupled, void* stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(params) || misaligned(grads) || misaligned(exp_avg) || misaligned(exp_avg_sq))
    XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  return adam_step(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay, step, decoupled, (cudaStream_t)stream);
}

extern "C" XATTN_API int xattn_emit_embeddings_bf16(const float* x, void* out_bf16, const int64_t* dst_clip, int32_t n_clips,
                                                    int32_t rows_per_clip, int32_t D, int32_t R, void* stream) {
  if (!x || !out_bf16) XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  if (misaligned(x) || misaligned(out_bf16)) XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  XF_TRY(check_device_sm100());
  return emit_embeddings_bf16((cudaStream_t)stream, x, out_bf16, dst_clip, n_clips, rows_per_clip, D, R);
}
