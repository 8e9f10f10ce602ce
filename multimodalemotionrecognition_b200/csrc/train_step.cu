// xattn_stage1_step: forward + cross entropy + backward + Adam of one rank's Stage-1 step (train.py:204-228) behind one call,
// optionally replayed as a CUDA graph.  The step's kernels take the Philox offset and the Adam step count by value, so a graph
// cannot simply be replayed: every call captures the launches again (host work only) and patches the cached executable graph
// with cudaGraphExecUpdate — the topology of a step is fixed by the config, only kernel arguments change.
#include <mutex>

#include "kernels_f32.cuh"

namespace xf {
int ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int B, int C, float label_smoothing, float grad_scale,
                    float* loss_out, float* dlogits, cudaStream_t st);
int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, float wd,
              int step, int decoupled, cudaStream_t st);
}  // namespace xf

using namespace xf;

namespace {

struct StepGraph {
  cudaStream_t capture = nullptr;   // private stream the step is captured on
  cudaGraphExec_t exec = nullptr;
  std::mutex mu;
};
StepGraph g_step_graph[64][3];   // per device and phase

int enqueue_step(const xattn_step_args& a, int64_t n_params, cudaStream_t st) {
  const xattn_config& c = *a.cfg;
  BackwardOptions opt;
  opt.phase = a.phase;
  if (a.phase != 2) {
    XF_TRY(forward_f32(c, a.params, a.video_emb, a.audio_emb, a.logits, a.workspace, st));
    XF_TRY(ce_loss_fwd_bwd(a.logits, a.labels, c.B, c.C, a.label_smoothing, a.grad_scale > 0.f ? a.grad_scale : 1.f, a.loss, a.dlogits, st));
  }
  XF_TRY(backward_f32(c, a.params, a.video_emb, a.audio_emb, a.dlogits, a.grads, a.workspace, st, opt));
  if (a.phase != 0) return XATTN_OK;
  return adam_step(a.params, a.grads, a.exp_avg, a.exp_avg_sq, n_params, a.lr, a.beta1, a.beta2, a.eps, a.weight_decay, a.step,
                   a.decoupled, st);
}

}  // namespace

extern "C" XATTN_API int xattn_stage1_step(const xattn_step_args* a, void* stream) {
  if (a == nullptr) XF_FAIL(XATTN_ERR_NULL, "args is NULL");
  XF_TRY(validate_config(a->cfg));
  const xattn_config& c = *a->cfg;
  if (a->phase < 0 || a->phase > 2) XF_FAIL(XATTN_ERR_SHAPE, "phase %d outside 0..2", a->phase);
  if (!a->params || !a->video_emb || !a->audio_emb || !a->labels || !a->logits || !a->loss || !a->dlogits || !a->grads || !a->workspace ||
      (a->phase == 0 && (!a->exp_avg || !a->exp_avg_sq)))
    XF_FAIL(XATTN_ERR_NULL, "NULL pointer argument");
  for (const void* p : {(const void*)a->params, a->video_emb, a->audio_emb, (const void*)a->logits, (const void*)a->dlogits,
                        (const void*)a->grads, (const void*)a->exp_avg, (const void*)a->exp_avg_sq, (const void*)a->workspace})
    if (((uintptr_t)p & 15) != 0) XF_FAIL(XATTN_ERR_ALIGN, "pointers must be 16-byte aligned");
  if (c.precision == XATTN_DTYPE_BF16) XF_FAIL(XATTN_ERR_DTYPE, "the training step runs with precision=f32 or tf32");
  if (!c.training && !c.save_for_backward) XF_FAIL(XATTN_ERR_SHAPE, "cfg needs training or save_for_backward for a training step");
  if (a->phase == 0 && a->step < 1) XF_FAIL(XATTN_ERR_SHAPE, "Adam step count is 1-based (got %d)", a->step);
  XF_TRY(check_device_sm100());
  int64_t need = 0;
  XF_TRY(xattn_workspace_bytes(a->cfg, &need));
  if (a->workspace_bytes < need)
    XF_FAIL(XATTN_ERR_WORKSPACE, "workspace too small: %lld < %lld bytes", (long long)a->workspace_bytes, (long long)need);
  if (c.B == 0) XF_FAIL(XATTN_ERR_SHAPE, "a training step needs at least one clip");
  ParamLayout pl;
  compute_param_layout(c, pl);
  cudaStream_t st = (cudaStream_t)stream;
  if (!a->use_graph) return enqueue_step(*a, pl.total, st);

  int dev = 0;
  XF_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) XF_FAIL(XATTN_ERR_CUDA, "device index %d out of range", dev);
  StepGraph& sg = g_step_graph[dev][a->phase];
  std::lock_guard<std::mutex> lock(sg.mu);
  if (sg.capture == nullptr) XF_CUDA(cudaStreamCreateWithFlags(&sg.capture, cudaStreamNonBlocking));
  // thread-local capture mode: other threads of the process may keep calling CUDA while this one captures
  XF_CUDA(cudaStreamBeginCapture(sg.capture, cudaStreamCaptureModeThreadLocal));
  const int rc = enqueue_step(*a, pl.total, sg.capture);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(sg.capture, &graph);   // always end the capture, also after a failed launch
  if (rc != XATTN_OK) {
    if (graph != nullptr) cudaGraphDestroy(graph);
    cudaGetLastError();
    return rc;
  }
  if (ce != cudaSuccess || graph == nullptr) XF_FAIL(XATTN_ERR_CUDA, "capturing the training step failed: %s", cudaGetErrorString(ce));
  bool ready = false;
  if (sg.exec != nullptr) {
    cudaGraphExecUpdateResultInfo info;
    if (cudaGraphExecUpdate(sg.exec, graph, &info) == cudaSuccess) {
      ready = true;
    } else {   // another topology (batch size, config): build a new executable graph
      cudaGetLastError();
      cudaGraphExecDestroy(sg.exec);
      sg.exec = nullptr;
    }
  }
  if (!ready) {
    const cudaError_t ie = cudaGraphInstantiate(&sg.exec, graph, 0);
    if (ie != cudaSuccess) {
      cudaGraphDestroy(graph);
      sg.exec = nullptr;
      XF_FAIL(XATTN_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ie));
    }
  }
  cudaGraphDestroy(graph);
  XF_CUDA(cudaGraphLaunch(sg.exec, st));
  return XATTN_OK;
}
