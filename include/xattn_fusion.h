/*
 * xattn_fusion.h — C ABI of the B200-native (sm_100a) cross-attention fusion head.
 *
 * The reference (Wionerlol/MultimodalEmotionRecognition) has NO native/FFI interface: the path is a
 * Python nn.Module (src/models/fusion.py:187-437).  This header is therefore the boundary a
 * maintainer binds with ctypes; every entry point names the reference code
 * it replaces.  Plain pointers and sizes only; no torch types.
 *
 * Conventions
 *  - every entry point returns int: 0 = XATTN_OK, otherwise an xattn_status code; the text of the
 *    last failure on the calling thread is returned by xattn_last_error().  Nothing aborts, no C++
 *    exception crosses the ABI.  There is NO CPU fallback: a non-sm_100 device is XATTN_ERR_ARCH.
 *  - all data pointers are DEVICE pointers owned by the caller (PyTorch's allocator); the library
 *    allocates nothing persistent and frees nothing.  Calls are asynchronous on `stream`
 *    (a cudaStream_t passed as void*).
 *  - parameters live in ONE flat fp32 buffer whose layout (offset/size per xattn_param_id, in
 *    floats) is defined by xattn_param_layout(); gradients use the same layout.  The Python module
 *    keeps the reference's state_dict names as views into that buffer.
 */
#ifndef XATTN_FUSION_H_
#define XATTN_FUSION_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XATTN_ABI_VERSION 21

#if defined(__GNUC__)
#define XATTN_API __attribute__((visibility("default")))
#else
#define XATTN_API
#endif

typedef enum xattn_status {
  XATTN_OK = 0,
  XATTN_ERR_SHAPE = 1,     /* unsupported or inconsistent dimensions (-> ValueError) */
  XATTN_ERR_DTYPE = 2,     /* unknown dtype / precision code          (-> ValueError) */
  XATTN_ERR_ALIGN = 3,     /* pointer not 16-byte aligned             (-> ValueError) */
  XATTN_ERR_ARCH = 4,      /* device is not sm_100                    (-> RuntimeError) */
  XATTN_ERR_CUDA = 5,      /* a CUDA call failed                      (-> RuntimeError) */
  XATTN_ERR_WORKSPACE = 6, /* workspace too small                     (-> RuntimeError) */
  XATTN_ERR_NULL = 7       /* required pointer is NULL                (-> ValueError) */
} xattn_status;

enum { XATTN_HEAD_CONCAT = 0, XATTN_HEAD_GATED = 1 };
enum { XATTN_POOL_MEAN = 0, XATTN_POOL_ATTN = 1, XATTN_POOL_TRANSFORMER = 2 };
enum { XATTN_DTYPE_F32 = 0, XATTN_DTYPE_BF16 = 1 };
/* xattn_config.precision only: the fp32 kernel family (forward, backward, training) with every Linear whose dimensions allow it
 * on the tensor cores — tcgen05 kind::tf32 straight from the fp32 tensors (kind::f16 for bf16 embeddings), fp32 accumulation. */
enum { XATTN_PREC_TF32 = 2 };

/* Problem description.  Field meaning follows FusionModel.__init__ (fusion.py:188-213). */
typedef struct xattn_config {
  int32_t B;            /* clips in this call                                        */
  int32_t T;            /* video tokens per clip (frames)                            */
  int32_t Ta;           /* audio tokens per clip (WavLM frames, 149 for 3 s)         */
  int32_t Dv;           /* video embedding dim   (video_model.embedding_dim, 512)    */
  int32_t Da;           /* audio sequence dim    (audio_model.sequence_dim, 768)     */
  int32_t E;            /* d_model                                                   */
  int32_t H;            /* num_heads                                                 */
  int32_t prior_dim;    /* xattn_emotion_prior_dim                                   */
  int32_t prior_hidden; /* xattn_emotion_prior_hidden_dim                            */
  int32_t common_dim;   /* hidden width of the concat head (fusion.py:313)           */
  int32_t C;            /* num_classes                                               */
  int32_t head_kind;    /* XATTN_HEAD_CONCAT | XATTN_HEAD_GATED (fusion.py:311/318)  */
  int32_t use_prior;    /* xattn_use_emotion_prior                                   */
  int32_t io_dtype;     /* dtype of video_emb / audio_emb: XATTN_DTYPE_*             */
  int32_t precision;    /* XATTN_DTYPE_F32: fp32 SIMT kernels (parity + training path)
                           XATTN_DTYPE_BF16: tcgen05 bf16 kernels, fp32 accumulate
                           XATTN_PREC_TF32: fp32 family, Linears on tcgen05 tf32     */
  int32_t training;     /* 1: dropouts active (Philox) and activations saved for
                           xattn_backward; 0: eval                                   */
  int32_t save_for_backward; /* 1: keep activations in the workspace even in eval mode
                                (gradient parity in eval mode)            */
  int32_t pool_mode;    /* XATTN_POOL_MEAN (temporal.py:108-109) | XATTN_POOL_ATTN (TemporalAttentionPooling,
                           temporal.py:9-26: LN -> Linear -> GELU -> Dropout -> Linear -> softmax over time -> weighted
                           sum; forward and backward in the fp32 / tf32 family, forward also on the bf16 tensor-core path)
                           XATTN_POOL_TRANSFORMER (TemporalTransformerPooling, temporal.py:46-75: sinusoidal positions,
                           pool_layers pre-norm TransformerEncoderLayers (GELU, FFN = 4·E, dropout p_pool), then the attention
                           pooling; fp32 / tf32 family, forward and backward) */
  int32_t pool_heads;   /* temporal_num_heads: heads of the pooler's self-attention (XATTN_POOL_TRANSFORMER) */
  int32_t pool_layers;  /* temporal_num_layers: encoder layers of XATTN_POOL_TRANSFORMER, 1..XATTN_MAX_POOL_LAYERS (0 reads as 1) */
  float p_attn;         /* attention dropout            (fusion.py:277,280)          */
  float p_path;         /* stochastic depth             (fusion.py:282-283)          */
  float p_head;         /* head Dropout(0.2)            (fusion.py:315,322)          */
  float p_prior;        /* prior_net Dropout            (fusion.py:161)              */
  float p_pool;         /* temporal_dropout: Dropout inside the attention pooler's score MLP (temporal.py:19) and the dropouts of
                           the transformer pooler's encoder layers (temporal.py:60-68: attention weights, both residual branches,
                           the FFN activation) */
  uint64_t seed;        /* Philox seed for the dropouts                               */
  uint64_t offset;      /* Philox offset (advance per step)                           */
} xattn_config;

/* Parameter ids: order of the flat parameter / gradient buffer.  Names are the reference's
 * state_dict keys.  audio_time_conv.* has no id of its own: on the WavLM path it is unused and receives no
 * gradient (fusion.py:273, :379-384); with the mel front end (xattn_mel_unfold) it is bound to the XP_SEQ_W / XP_SEQ_B slots. */
typedef enum xattn_param_id {
  XP_V_IN_W = 0,   /* v_in_proj.weight            [E,Dv]   */
  XP_V_IN_B,       /* v_in_proj.bias              [E]      */
  XP_A_IN_W,       /* a_in_proj.weight            [E,E]    */
  XP_A_IN_B,       /* a_in_proj.bias              [E]      */
  XP_SEQ_W,        /* audio_seq_proj.weight       [E,Da]   */
  XP_SEQ_B,        /* audio_seq_proj.bias         [E]      */
  XP_V2A_IN_W,     /* v2a_attn.in_proj_weight     [3E,E]   */
  XP_V2A_IN_B,     /* v2a_attn.in_proj_bias       [3E]     */
  XP_V2A_OUT_W,    /* v2a_attn.out_proj.weight    [E,E]    */
  XP_V2A_OUT_B,    /* v2a_attn.out_proj.bias      [E]      */
  XP_A2V_IN_W,     /* a2v_attn.in_proj_weight     [3E,E]   */
  XP_A2V_IN_B,     /* a2v_attn.in_proj_bias       [3E]     */
  XP_A2V_OUT_W,    /* a2v_attn.out_proj.weight    [E,E]    */
  XP_A2V_OUT_B,    /* a2v_attn.out_proj.bias      [E]      */
  XP_VN_W,         /* v_norm.weight               [E]      */
  XP_VN_B,         /* v_norm.bias                 [E]      */
  XP_AN_W,         /* a_norm.weight               [E]      */
  XP_AN_B,         /* a_norm.bias                 [E]      */
  XP_HEAD0_W,      /* xattn_mlp.0.weight [common,2E]   | xattn_gate.0.weight [E,2E] */
  XP_HEAD0_B,      /* xattn_mlp.0.bias   [common]      | xattn_gate.0.bias   [E]    */
  XP_HEAD1_W,      /* xattn_mlp.3.weight [C,common]    | xattn_gate.3.weight [1,E]  */
  XP_HEAD1_B,      /* xattn_mlp.3.bias   [C]           | xattn_gate.3.bias   [1]    */
  XP_CLS_W,        /* (gated) xattn_classifier.weight [C,E]; size 0 for concat      */
  XP_CLS_B,        /* (gated) xattn_classifier.bias   [C];   size 0 for concat      */
  XP_PRIOR0_W,     /* emotion_prior_bias.prior_net.0.weight [Hd,2E] (size 0 if !use_prior, also below) */
  XP_PRIOR0_B,     /* emotion_prior_bias.prior_net.0.bias   [Hd]     */
  XP_PRIOR3_W,     /* emotion_prior_bias.prior_net.3.weight [P,Hd]   */
  XP_PRIOR3_B,     /* emotion_prior_bias.prior_net.3.bias   [P]      */
  XP_VQ_W,         /* emotion_prior_bias.v_query_bias.weight [1,E+P] */
  XP_VQ_B,         /* emotion_prior_bias.v_query_bias.bias   [1]     */
  XP_AK_W,         /* emotion_prior_bias.a_key_bias.weight   [1,E+P] */
  XP_AK_B,         /* emotion_prior_bias.a_key_bias.bias     [1]     */
  XP_AQ_W,         /* emotion_prior_bias.a_query_bias.weight [1,E+P] */
  XP_AQ_B,         /* emotion_prior_bias.a_query_bias.bias   [1]     */
  XP_VK_W,         /* emotion_prior_bias.v_key_bias.weight   [1,E+P] */
  XP_VK_B,         /* emotion_prior_bias.v_key_bias.bias     [1]     */
  XP_BIAS_SCALE,   /* emotion_prior_bias.bias_scale          []      */
  /* temporal_pooling="attn" (size 0 for "mean"): {v,a}_temporal_pool.pool.score.{0,1,4}.{weight,bias} (temporal.py:15-21) */
  XP_VPOOL_LN_W,   /* v_temporal_pool.pool.score.0.weight [E]       */
  XP_VPOOL_LN_B,   /* v_temporal_pool.pool.score.0.bias   [E]       */
  XP_VPOOL_W1,     /* v_temporal_pool.pool.score.1.weight [E/2,E]   */
  XP_VPOOL_B1,     /* v_temporal_pool.pool.score.1.bias   [E/2]     */
  XP_VPOOL_W2,     /* v_temporal_pool.pool.score.4.weight [1,E/2]   */
  XP_VPOOL_B2,     /* v_temporal_pool.pool.score.4.bias   [1]       */
  XP_APOOL_LN_W,   /* a_temporal_pool.pool.score.* as above          */
  XP_APOOL_LN_B,
  XP_APOOL_W1,
  XP_APOOL_B1,
  XP_APOOL_W2,
  XP_APOOL_B2,
  /* XATTN_POOL_TRANSFORMER: {v,a}_temporal_pool.pool.encoder.layers.0.* (first layer); the XP_?POOL_* ids above then hold
   * ...pool.pool.score.*.  Layers l >= 1 follow XP_COUNT: see xattn_param_count / XATTN_TX_PARAM_ID below. */
  XP_VTX_IN_W,     /* self_attn.in_proj_weight [3E,E] */
  XP_VTX_IN_B,     /* self_attn.in_proj_bias   [3E]   */
  XP_VTX_OUT_W,    /* self_attn.out_proj.weight [E,E] */
  XP_VTX_OUT_B,    /* self_attn.out_proj.bias   [E]   */
  XP_VTX_FF1_W,    /* linear1.weight [F,E], F = max(2E, 4E) */
  XP_VTX_FF1_B,    /* linear1.bias   [F]   */
  XP_VTX_FF2_W,    /* linear2.weight [E,F] */
  XP_VTX_FF2_B,    /* linear2.bias   [E]   */
  XP_VTX_N1_W,     /* norm1.weight [E] */
  XP_VTX_N1_B,
  XP_VTX_N2_W,     /* norm2.weight [E] */
  XP_VTX_N2_B,
  XP_ATX_IN_W, XP_ATX_IN_B, XP_ATX_OUT_W, XP_ATX_OUT_B, XP_ATX_FF1_W, XP_ATX_FF1_B, XP_ATX_FF2_W, XP_ATX_FF2_B,
  XP_ATX_N1_W, XP_ATX_N1_B, XP_ATX_N2_W, XP_ATX_N2_B,
  XP_COUNT
} xattn_param_id;

/* temporal_num_layers > 1 (temporal.py:51, :69): the 12 tensors of encoder layer `layer` >= 1 of side 0 (video) / 1 (audio),
 * k = 0..11 in the order of XP_VTX_IN_W .. XP_VTX_N2_B, take the ids after XP_COUNT; layer 0 keeps XP_VTX_* / XP_ATX_*. */
#define XATTN_MAX_POOL_LAYERS 8
#define XATTN_TX_PARAMS_PER_LAYER 12
#define XATTN_TX_PARAM_ID(side, layer, k)                                                                      \
  ((layer) == 0 ? ((side) == 0 ? XP_VTX_IN_W : XP_ATX_IN_W) + (k)                                               \
                : XP_COUNT + (((layer)-1) * 2 + (side)) * XATTN_TX_PARAMS_PER_LAYER + (k))
#define XATTN_PARAM_ID_MAX (XP_COUNT + 2 * XATTN_TX_PARAMS_PER_LAYER * (XATTN_MAX_POOL_LAYERS - 1))

/* ---- library info ------------------------------------------------------------------------ */
XATTN_API int xattn_abi_version(void);
/* Text of the last error raised on this thread ("" if none). */
XATTN_API const char* xattn_last_error(void);

/* ---- layout queries (host only, no CUDA) --------------------------------------------------- */
/* Number of parameter ids of cfg: XP_COUNT, plus 24 per transformer-pooler layer beyond the first. */
XATTN_API int xattn_param_count(const xattn_config* cfg);
/* Fills offsets[n] / sizes[n], n = xattn_param_count(cfg) (in floats; each offset is a multiple of 4 floats so
 * every tensor is 16-byte aligned) and *total_floats.  Replaces: the nn.Parameter tree built at
 * fusion.py:269-326 (state_dict layout). */
XATTN_API int xattn_param_layout(const xattn_config* cfg, int64_t* offsets, int64_t* sizes, int64_t* total_floats);
/* Bytes of scratch/activation workspace xattn_forward needs for cfg (depends on B, precision,
 * training/save_for_backward). */
XATTN_API int xattn_workspace_bytes(const xattn_config* cfg, int64_t* bytes);
/* Bytes of the derived ("packed") bf16 weight buffer used by precision == BF16. */
XATTN_API int xattn_packed_bytes(const xattn_config* cfg, int64_t* bytes);

/* ---- forward ------------------------------------------------------------------------------- */
/* Derives the bf16 operand set (rounded weights, the audio_seq_proj∘a_in_proj fold of SURVEY 3.1
 * fact 2, pre-multiplied attention weights) from the flat fp32 parameters.  Call once per weight
 * update; inference callers cache the result. */
XATTN_API int xattn_pack_weights(const xattn_config* cfg, const float* params, void* packed, void* stream);

/* logits[B,C] (fp32) = FusionModel.forward for mode="xattn", from the point where the frozen
 * backbones have produced video_emb [B,T,Dv] and audio_emb [B,Ta,Da] (dtype cfg->io_dtype,
 * contiguous).  Replaces fusion.py:372-411 (+ :170-184, :351-354, temporal.py:108-109).
 * `packed` may be NULL when precision == F32. */
XATTN_API int xattn_forward(const xattn_config* cfg, const float* params, const void* packed, const void* video_emb,
                  const void* audio_emb, float* logits, void* workspace, int64_t workspace_bytes, void* stream);

/* xattn_forward for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same kernels and results as
 * xattn_forward (eval mode only), but the scratch workspace is a static __device__ array inside the library of
 * XATTN_NOALLOC_MAX_WORKSPACE_BYTES (held by every GPU that loads the library until the process exits), so the call itself
 * performs no cudaMalloc / cudaFree, creates no stream or event, and runs every kernel on `stream`.  A config whose
 * xattn_workspace_bytes exceeds the maximum is refused with XATTN_ERR_WORKSPACE (bf16 path at T=16, Ta=149: B <= 66;
 * fp32 path: B <= 8).  All calls on a device share the one array: it keeps the last call's contents (nothing needs to start
 * at zero — every region is written before it is read) and calls that OVERLAP on the GPU (different streams or threads) RACE;
 * calls ordered by one stream are safe. */
#define XATTN_NOALLOC_MAX_WORKSPACE_BYTES (8 * 1024 * 1024)
XATTN_API int64_t xattn_noalloc_max_workspace_bytes(void);
XATTN_API int xattn_forward_noalloc(const xattn_config* cfg, const float* params, const void* packed, const void* video_emb,
                                    const void* audio_emb, float* logits, void* stream);

/* Backward of the same head for the Stage-1 step (frozen backbones => no input gradients).
 * `workspace` must be the buffer a preceding xattn_forward (same cfg, training or
 * save_for_backward set) filled.  grads (flat, xattn_param_layout) is OVERWRITTEN.
 * Replaces: loss.backward() over the fusion parameters at train.py:227. */
XATTN_API int xattn_backward(const xattn_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                   const float* dlogits, float* grads, void* workspace, int64_t workspace_bytes, void* stream);

/* xattn_backward with two options.
 *  - tail_stream (a cudaStream_t, may be NULL): the library makes it wait for the point of `stream` where every gradient except
 *    the input projections' — the flat buffer from *tail_offset_floats (= the offset of XP_V2A_IN_W) to its end — is final, so a
 *    data-parallel caller can all-reduce that part on tail_stream while the input-projection backward is still running
 *    ("launch it on a side stream as soon as the last backward kernel writes the bucket").
 *  - d_video_emb [B,T,Dv] / d_audio_emb [B,Ta,Da] (fp32, may be NULL): gradients w.r.t. the embeddings, for Stage 2 of the
 *    reference where backbone layers are trainable (train.py:815-830; loss.backward() at :227 then reaches v_feat / a_seq). */
XATTN_API int xattn_backward_ex(const xattn_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                      const float* dlogits, float* grads, void* workspace, int64_t workspace_bytes, void* stream,
                      void* tail_stream, int64_t* tail_offset_floats, float* d_video_emb, float* d_audio_emb);

/* ---- loss / optimizer ------------------------------------------------------------------------ */
/* Mean-reduced cross entropy with label smoothing + its gradient w.r.t. logits.
 * Replaces nn.CrossEntropyLoss(label_smoothing) at train.py:1033 / :218.  loss_out: 1 float.
 * labels: int64[B], each in [0, C) — as with torch (which raises a device-side assert) anything else is a caller error and is
 * not checked here.  grad_scale multiplies dlogits (1/world_size under data parallelism). */
XATTN_API int xattn_ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t B, int32_t C, float label_smoothing,
                          float grad_scale, float* loss_out, float* dlogits, void* stream);

/* One fused Adam step over n floats.  decoupled == 0: torch.optim.Adam(weight_decay) — coupled L2,
 * what the reference runs (train.py:872); decoupled == 1: AdamW.  step is 1-based. */
XATTN_API int xattn_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                    float beta1, float beta2, float eps, float weight_decay, int32_t step, int32_t decoupled,
                    void* stream);

/* One whole Stage-1 step in ONE call: xattn_forward (cfg as given: training / save_for_backward, precision f32 or tf32) ->
 * xattn_ce_loss_fwd_bwd -> xattn_backward -> xattn_adam_step, the loop body of train_one_epoch (train.py:204-228) for one rank.
 * use_graph != 0: the launches of the step are stream-captured and run as one CUDA graph (the ~58 kernels of a step are 10-60 us
 * each: replayed as a graph the gaps between them go, 1.73 -> 1.62 ms at B=1024).  The Philox offset and the Adam step count are
 * kernel arguments, so every call re-captures (host work only, ~0.3 ms, overlapping the previous step's GPU time) and UPDATES the
 * cached executable graph in place (cudaGraphExecUpdate: same topology, new arguments); it is re-instantiated only when the
 * topology changes (another B, another config).  One cached graph per device; calls on a device are serialised by a mutex.
 * All buffers are the caller's (nothing is allocated): logits [B,C], loss [1], dlogits [B,C], grads / exp_avg / exp_avg_sq
 * [xattn_param_layout total], workspace as xattn_workspace_bytes(cfg).  Data-parallel training splits the step at the
 * gradient all-reduce (phase 1 / phase 2 below) and runs Adam through xattn_adam_step. */
typedef struct xattn_step_args {
  const xattn_config* cfg;
  float* params;
  const void* video_emb;
  const void* audio_emb;
  const int64_t* labels;
  float* logits;
  float* loss;
  float* dlogits;
  float* grads;
  float* exp_avg;
  float* exp_avg_sq;
  void* workspace;
  int64_t workspace_bytes;
  float lr, beta1, beta2, eps, weight_decay, label_smoothing;
  int32_t step;       /* 1-based Adam step */
  int32_t decoupled;  /* 0: torch.optim.Adam(weight_decay), 1: AdamW */
  int32_t use_graph;
  int32_t phase;      /* 0: the whole step.  Data parallel: 1 = forward + CE + the backward up to the point where every gradient
                         but the input projections' (the flat buffer from XP_V2A_IN_W on) is final; 2 = the input-projection
                         backward.  The caller all-reduces the tail bucket on another stream after phase 1, beside phase 2, then
                         the head bucket, then calls xattn_adam_step.  One cached graph per device and phase. */
  float grad_scale;   /* multiplies dlogits: 1 / world_size under data parallelism (0 reads as 1) */
} xattn_step_args;
XATTN_API int xattn_stage1_step(const xattn_step_args* args, void* stream);

/* ---- flat concat / gated heads (BASELINE config 5) -------------------------------------------- */
typedef struct flat_head_config {
  int32_t B, T, Ta, Dv, Da; /* sequence inputs [B,T,Dv], [B,Ta,Da]; T == Ta == 1 for pre-pooled inputs */
  int32_t align_dim;        /* ClipStyleAlignment width (fusion.py:132-133); 0 = no alignment           */
  int32_t common_dim;       /* audio_proj/video_proj width (fusion.py:240-241)                          */
  int32_t C;
  int32_t gated;            /* 0: mode="concat" (fusion.py:243-248), 1: mode="gated" (:252-259)         */
  int32_t io_dtype;
  int32_t precision;        /* XATTN_DTYPE_F32: fp32 SIMT arithmetic; XATTN_DTYPE_BF16: temporal mean -> bf16, every Linear
                               of width >= 128 on the tcgen05 projection GEMM (bf16 embeddings, dims %% 64 / %% 128);
                               XATTN_PREC_TF32: fp32 arithmetic with the Linears on tcgen05 tf32 where dims allow  */
  /* training (precision F32 / TF32 only; fusion.py:246, :255, :430): */
  int32_t training;         /* 1: Dropout(p_drop) behind the hidden layer's ReLU (fusion.2 / gate.2) is active (Philox)       */
  int32_t drop_audio;       /* ModalityDropout (fusion.py:29-55, gated mode, training): the caller draws the two host-side   */
  int32_t drop_video;       /*   Bernoulli decisions as the reference does (torch.rand(1).item() < p) and passes them here    */
  float p_drop;             /* 0.2 in the reference                                                                           */
  uint64_t seed;            /* Philox seed / offset of the hidden-layer dropout                                               */
  uint64_t offset;
} flat_head_config;

typedef enum flat_param_id {
  FP_ALIGN_A_W = 0, /* semantic_alignment.audio_proj.weight [align,Da] */
  FP_ALIGN_A_B,
  FP_ALIGN_V_W,     /* semantic_alignment.video_proj.weight [align,Dv] */
  FP_ALIGN_V_B,
  FP_AUDIO_W,       /* audio_proj.weight [common, align|Da] */
  FP_AUDIO_B,
  FP_VIDEO_W,       /* video_proj.weight [common, align|Dv] */
  FP_VIDEO_B,
  FP_H0_W,          /* fusion.0.weight | gate.0.weight [common, 2*common] */
  FP_H0_B,
  FP_H1_W,          /* fusion.3.weight [C,common] | gate.3.weight [1,common] */
  FP_H1_B,
  FP_CLS_W,         /* (gated) classifier.weight [C,common] */
  FP_CLS_B,
  FP_COUNT
} flat_param_id;

XATTN_API int flat_head_param_layout(const flat_head_config* cfg, int64_t* offsets, int64_t* sizes, int64_t* total_floats);
XATTN_API int flat_head_workspace_bytes(const flat_head_config* cfg, int64_t* bytes);
/* logits[B,C] for mode in {"concat","gated"} with the temporal mean (the parameter-free last step
 * of audio_model.encode / video_model.encode, fusion.py:414-415) fused in.  Replaces
 * fusion.py:413-435 (+ :137-139).  cfg->training = 0: eval mode (ModalityDropout / Dropout identity); 1 (precision F32 / TF32):
 * the hidden Dropout and the caller's ModalityDropout decisions apply, and the workspace keeps what flat_head_backward reads. */
XATTN_API int flat_head_forward(const flat_head_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                      float* logits, void* workspace, int64_t workspace_bytes, void* stream);
/* Backward of flat_head_forward (precision F32 / TF32; `workspace` is the buffer that forward filled, same cfg): the part of
 * loss.backward() (train.py:227) that runs through fusion.py:413-435.  grads (flat, flat_head_param_layout) is OVERWRITTEN.
 * d_video_emb [B,T,Dv] / d_audio_emb [B,Ta,Da] (fp32, may be NULL): gradients w.r.t. the embeddings — the temporal mean spreads
 * d_pooled / L over the tokens — for trainable backbones. */
XATTN_API int flat_head_backward(const flat_head_config* cfg, const float* params, const void* video_emb, const void* audio_emb,
                       const float* dlogits, float* grads, float* d_video_emb, float* d_audio_emb, void* workspace,
                       int64_t workspace_bytes, void* stream);

/* Serving epilogue (optimized_runtime.py:99-108 softmax(dim=1); inference_worker.py:136-143 row.argmax() and the probability at
 * it) in one launch: probs[B,C] = softmax(logits) (may alias logits; NULL to skip), top1[B] = first maximal class, top1_prob[B]
 * = its probability (either may be NULL). */
XATTN_API int xattn_softmax_top1(const float* logits, int32_t B, int32_t C, float* probs, int32_t* top1, float* top1_prob, void* stream);

/* Mel front end (fusion.py:379-384: audio_time_conv = Conv1d(n_mels -> a_dim, kernel 3, padding 1) over time, then permute to
 * [B, Ta, a_dim]) for audio encoders without encode_sequence().  mel [B, n_mels, Ta] -> unfolded [B, Ta, 3*n_mels] (same
 * dtype), unfolded[b,t,c*3+k] = mel[b,c,t+k-1] (0 outside [0,Ta)).  The convolution then IS the audio_seq_proj Linear of the
 * head: call xattn_forward / xattn_backward with cfg.Da = 3*n_mels and the XP_SEQ_W / XP_SEQ_B slots holding
 * audio_time_conv.weight (viewed [a_dim, n_mels*3], its own memory order) / audio_time_conv.bias. */
XATTN_API int xattn_mel_unfold(const void* mel, void* unfolded, int32_t io_dtype, int32_t B, int32_t n_mels, int32_t Ta, void* stream);

/* The ClipStyleAlignment loss of the flat heads (fusion.py:141-149; what FusionModel.pop_alignment_loss() returns, train.py:221-222):
 * 0.5·(CE(s·â v̂ᵀ, arange(B)) + CE(s·v̂ âᵀ, arange(B))) with â, v̂ the L2-normalised alignment projections of the temporally
 * pooled embeddings and s = min(exp(*logit_scale), 100).  A B x B cross-sample term: under batch sharding gather the batch first.
 * `logit_scale`: device pointer to semantic_alignment.logit_scale; `loss`: device scalar.  precision != F32 runs the three
 * contractions on tcgen05 (tf32 products). */
XATTN_API int flat_head_alignment_workspace_bytes(const flat_head_config* cfg, int64_t* bytes);
XATTN_API int flat_head_alignment_loss(const flat_head_config* cfg, const float* params, const float* logit_scale, const void* video_emb,
                                       const void* audio_emb, float* loss, void* workspace, int64_t workspace_bytes, void* stream);
/* Backward of flat_head_alignment_loss: `dloss` is a DEVICE scalar (the upstream gradient, fusion_align_weight · dL, train.py:225).
 * grads (flat, flat_head_param_layout) is OVERWRITTEN: the four semantic_alignment.*_proj entries hold their gradients, everything
 * else zero; *d_logit_scale (device scalar) the gradient of semantic_alignment.logit_scale (zero where exp(logit_scale) is clamped
 * at 100).  d_video_emb / d_audio_emb as in flat_head_backward (may be NULL).  Recomputes the forward quantities: needs only the
 * workspace size of flat_head_alignment_workspace_bytes, not the forward call's buffer. */
XATTN_API int flat_head_alignment_backward(const flat_head_config* cfg, const float* params, const float* logit_scale,
                                 const void* video_emb, const void* audio_emb, const float* dloss, float* grads,
                                 float* d_logit_scale, float* d_video_emb, float* d_audio_emb, void* workspace,
                                 int64_t workspace_bytes, void* stream);

/* Backbone tail -> head hand-off (fusion.py:368-370, :377; video.py:23 AdaptiveAvgPool2d; wavlm_audio.py:183 last_hidden_state).
 * x fp32 [n_clips · rows_per_clip, D, R] -> the mean over the R innermost values (R = h·w of ResNet18's last feature map, R = 1
 * for WavLM hidden states), rounded to bf16, written to rows (dst_clip[c] · rows_per_clip + j) of out_bf16 [*, D] (dst_clip NULL:
 * c itself; device int64).  The embeddings reach the tensor-core kernels without an fp32 round trip and land directly in a
 * resident embedding table (FrozenEmbeddingCache). */
XATTN_API int xattn_emit_embeddings_bf16(const float* x, void* out_bf16, const int64_t* dst_clip, int32_t n_clips, int32_t rows_per_clip,
                               int32_t D, int32_t R, void* stream);

/* ---- diagnostics ---------------------------------------------------------------------------------- */
/* Kernel launches issued by this library since process start (bench.py's gpu_launches evidence). */
XATTN_API unsigned long long xattn_launch_count(void);
/* Per-kernel timing of xattn_forward(precision = BF16).  While enabled, every kernel of the forward is bracketed by a
 * pair of CUDA events recorded on the caller's stream; xattn_stage_timing_read waits for them, writes the summed
 * milliseconds and bracket counts of up to `cap` stages (order: xattn_stage_name(i)), clears the record and returns
 * the number of stages.  bench.py's roofline durations come from here.  Not thread-safe; leave off in production. */
XATTN_API int xattn_stage_timing(int32_t enable);
XATTN_API int xattn_stage_timing_read(double* ms_total, int64_t* launches, int32_t cap);
XATTN_API const char* xattn_stage_name(int32_t i);
#ifdef __cplusplus
}
#endif
#endif /* XATTN_FUSION_H_ */
