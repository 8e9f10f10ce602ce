"""ctypes binding of ``libxattn_fusion.so`` (C ABI: ``include/xattn_fusion.h``).

There is no CPU or PyTorch-eager fallback: if the shared library has not been built
(``python __graft_entry__.py build`` or ``make -C multimodalemotionrecognition_b200/csrc``) loading raises,
and every entry point refuses non-sm_100 devices with ``XATTN_ERR_ARCH``.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

LIB_NAME = "libxattn_fusion.so"
LIB_PATH = Path(__file__).resolve().parent / LIB_NAME
DEBUG_LIB_PATH = Path(__file__).resolve().parent / "libxattn_fusion_debug.so"
ABI_VERSION = 21

# xattn_status
OK, ERR_SHAPE, ERR_DTYPE, ERR_ALIGN, ERR_ARCH, ERR_CUDA, ERR_WORKSPACE, ERR_NULL = range(8)
HEAD_CONCAT, HEAD_GATED = 0, 1
DTYPE_F32, DTYPE_BF16 = 0, 1
POOL_MEAN, POOL_ATTN, POOL_TRANSFORMER = 0, 1, 2
PREC_TF32 = 2   # xattn_config.precision only: fp32 kernel family with the Linears on tcgen05 tf32 (gemm_tf32.cu)

# xattn_param_id -> reference state_dict key (None: resolved per head kind)
XP_NAMES = [
    "v_in_proj.weight", "v_in_proj.bias", "a_in_proj.weight", "a_in_proj.bias",
    "audio_seq_proj.weight", "audio_seq_proj.bias",
    "v2a_attn.in_proj_weight", "v2a_attn.in_proj_bias", "v2a_attn.out_proj.weight", "v2a_attn.out_proj.bias",
    "a2v_attn.in_proj_weight", "a2v_attn.in_proj_bias", "a2v_attn.out_proj.weight", "a2v_attn.out_proj.bias",
    "v_norm.weight", "v_norm.bias", "a_norm.weight", "a_norm.bias",
    "HEAD0.weight", "HEAD0.bias", "HEAD1.weight", "HEAD1.bias", "xattn_classifier.weight", "xattn_classifier.bias",
    "emotion_prior_bias.prior_net.0.weight", "emotion_prior_bias.prior_net.0.bias",
    "emotion_prior_bias.prior_net.3.weight", "emotion_prior_bias.prior_net.3.bias",
    "emotion_prior_bias.v_query_bias.weight", "emotion_prior_bias.v_query_bias.bias",
    "emotion_prior_bias.a_key_bias.weight", "emotion_prior_bias.a_key_bias.bias",
    "emotion_prior_bias.a_query_bias.weight", "emotion_prior_bias.a_query_bias.bias",
    "emotion_prior_bias.v_key_bias.weight", "emotion_prior_bias.v_key_bias.bias",
    "emotion_prior_bias.bias_scale",
    # temporal_pooling="attn" (temporal.py:15-21); absent (size 0) for "mean"
    "v_temporal_pool.pool.score.0.weight", "v_temporal_pool.pool.score.0.bias", "v_temporal_pool.pool.score.1.weight",
    "v_temporal_pool.pool.score.1.bias", "v_temporal_pool.pool.score.4.weight", "v_temporal_pool.pool.score.4.bias",
    "a_temporal_pool.pool.score.0.weight", "a_temporal_pool.pool.score.0.bias", "a_temporal_pool.pool.score.1.weight",
    "a_temporal_pool.pool.score.1.bias", "a_temporal_pool.pool.score.4.weight", "a_temporal_pool.pool.score.4.bias",
]
# temporal_pooling="transformer" (temporal.py:46-75): the 12 tensors of one encoder layer, first layer inside XP_COUNT, layers >= 1
# after it (XATTN_TX_PARAM_ID in the header); the score ids above then name ...pool.pool.score.*
TX_LAYER_NAMES = ("self_attn.in_proj_weight", "self_attn.in_proj_bias", "self_attn.out_proj.weight", "self_attn.out_proj.bias",
                  "linear1.weight", "linear1.bias", "linear2.weight", "linear2.bias", "norm1.weight", "norm1.bias", "norm2.weight",
                  "norm2.bias")
MAX_POOL_LAYERS = 8
XP_NAMES += [f"{side}_temporal_pool.pool.encoder.layers.0.{n}" for side in ("v", "a") for n in TX_LAYER_NAMES]
XP_COUNT = len(XP_NAMES)

FP_NAMES = [
    "semantic_alignment.audio_proj.weight", "semantic_alignment.audio_proj.bias",
    "semantic_alignment.video_proj.weight", "semantic_alignment.video_proj.bias",
    "audio_proj.weight", "audio_proj.bias", "video_proj.weight", "video_proj.bias",
    "H0.weight", "H0.bias", "H1.weight", "H1.bias", "classifier.weight", "classifier.bias",
]
FP_COUNT = len(FP_NAMES)


def xattn_param_names(head: str, pool_layers: int = 1) -> list[str]:
    """state_dict key of every parameter id for the given head kind (reference fusion.py:311-326); ``pool_layers`` > 1 appends the
    transformer pooler's further encoder layers (layer-major, video before audio: XATTN_TX_PARAM_ID)."""
    h0, h1 = ("xattn_mlp.0", "xattn_mlp.3") if head == "concat" else ("xattn_gate.0", "xattn_gate.3")
    names = [n.replace("HEAD0", h0).replace("HEAD1", h1) for n in XP_NAMES]
    for layer in range(1, max(1, pool_layers)):
        names += [f"{side}_temporal_pool.pool.encoder.layers.{layer}.{n}" for side in ("v", "a") for n in TX_LAYER_NAMES]
    return names


def flat_param_names(gated: bool) -> list[str]:
    h0, h1 = ("gate.0", "gate.3") if gated else ("fusion.0", "fusion.3")
    return [n.replace("H0", h0).replace("H1", h1) for n in FP_NAMES]


class XAttnConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "B", "T", "Ta", "Dv", "Da", "E", "H", "prior_dim", "prior_hidden", "common_dim", "C", "head_kind",
        "use_prior", "io_dtype", "precision", "training", "save_for_backward", "pool_mode", "pool_heads", "pool_layers")] + [
        ("p_attn", C.c_float), ("p_path", C.c_float), ("p_head", C.c_float), ("p_prior", C.c_float), ("p_pool", C.c_float),
        ("seed", C.c_uint64), ("offset", C.c_uint64)]


class XAttnStepArgs(C.Structure):
    """include/xattn_fusion.h: struct xattn_step_args (xattn_stage1_step)."""
    _fields_ = [("cfg", C.POINTER(XAttnConfig)), ("params", C.c_void_p), ("video_emb", C.c_void_p), ("audio_emb", C.c_void_p),
                ("labels", C.c_void_p), ("logits", C.c_void_p), ("loss", C.c_void_p), ("dlogits", C.c_void_p), ("grads", C.c_void_p),
                ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_int64),
                ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("weight_decay", C.c_float),
                ("label_smoothing", C.c_float), ("step", C.c_int32), ("decoupled", C.c_int32), ("use_graph", C.c_int32),
                ("phase", C.c_int32), ("grad_scale", C.c_float)]


class FlatHeadConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("B", "T", "Ta", "Dv", "Da", "align_dim", "common_dim", "C", "gated", "io_dtype",
                                                 "precision", "training", "drop_audio", "drop_video")] + [
        ("p_drop", C.c_float), ("seed", C.c_uint64), ("offset", C.c_uint64)]


_EXPORTS = {
    # name: (restype, argtypes)
    "xattn_abi_version": (C.c_int, []),
    "xattn_last_error": (C.c_char_p, []),
    "xattn_launch_count": (C.c_ulonglong, []),
    "xattn_stage_timing": (C.c_int, [C.c_int32]),
    "xattn_stage_timing_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_int32]),
    "xattn_stage_name": (C.c_char_p, [C.c_int32]),
    "xattn_param_count": (C.c_int, [C.POINTER(XAttnConfig)]),
    "xattn_param_layout": (C.c_int, [C.POINTER(XAttnConfig), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "xattn_workspace_bytes": (C.c_int, [C.POINTER(XAttnConfig), C.POINTER(C.c_int64)]),
    "xattn_packed_bytes": (C.c_int, [C.POINTER(XAttnConfig), C.POINTER(C.c_int64)]),
    "xattn_pack_weights": (C.c_int, [C.POINTER(XAttnConfig), C.c_void_p, C.c_void_p, C.c_void_p]),
    "xattn_forward": (C.c_int, [C.POINTER(XAttnConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.c_int64, C.c_void_p]),
    "xattn_noalloc_max_workspace_bytes": (C.c_int64, []),
    "xattn_forward_noalloc": (C.c_int, [C.POINTER(XAttnConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
    "xattn_backward": (C.c_int, [C.POINTER(XAttnConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_int64, C.c_void_p]),
    "xattn_backward_ex": (C.c_int, [C.POINTER(XAttnConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p]),
    "xattn_stage1_step": (C.c_int, [C.POINTER(XAttnStepArgs), C.c_void_p]),
    "xattn_ce_loss_fwd_bwd": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_void_p,
                                        C.c_void_p, C.c_void_p]),
    "xattn_adam_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float,
                                  C.c_float, C.c_float, C.c_float, C.c_int32, C.c_int32, C.c_void_p]),
    "flat_head_param_layout": (C.c_int, [C.POINTER(FlatHeadConfig), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                         C.POINTER(C.c_int64)]),
    "flat_head_workspace_bytes": (C.c_int, [C.POINTER(FlatHeadConfig), C.POINTER(C.c_int64)]),
    "flat_head_alignment_workspace_bytes": (C.c_int, [C.POINTER(FlatHeadConfig), C.POINTER(C.c_int64)]),
    "flat_head_alignment_loss": (C.c_int, [C.POINTER(FlatHeadConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_void_p, C.c_int64, C.c_void_p]),
    "xattn_emit_embeddings_bf16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "xattn_softmax_top1": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "xattn_mel_unfold": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "flat_head_forward": (C.c_int, [C.POINTER(FlatHeadConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int64, C.c_void_p]),
    "flat_head_backward": (C.c_int, [C.POINTER(FlatHeadConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "flat_head_alignment_backward": (C.c_int, [C.POINTER(FlatHeadConfig), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
}
# test-only entry points of libxattn_fusion_debug.so (include/xattn_fusion_debug.h); the product library does not export them
_DEBUG_EXPORTS = {
    "xattn_debug_dropout_factors": (C.c_int, [C.c_float, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint64, C.c_int64, C.c_void_p,
                                              C.c_void_p]),
    "xattn_debug_gemm_tc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_void_p]),
    "xattn_debug_gemm_tf32": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                        C.c_int32, C.c_int32, C.c_void_p]),
    "xattn_debug_umma_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                         C.c_void_p]),
}
EXPORTED_SYMBOLS = tuple(_EXPORTS)
DEBUG_SYMBOLS = tuple(_DEBUG_EXPORTS)

_lib = None


def load() -> C.CDLL:
    """Load the extension once.  Raises RuntimeError when it is missing — by design there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = Path(os.environ.get("XATTN_FUSION_LIB", str(LIB_PATH)))
    if not path.exists():
        raise RuntimeError(
            f"{path} not found: the CUDA extension is not built. Run `python __graft_entry__.py build` "
            "(or `make -C multimodalemotionrecognition_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(str(path))
    for name, (res, args) in _EXPORTS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.xattn_abi_version() != ABI_VERSION:
        raise RuntimeError(f"{path}: ABI version {lib.xattn_abi_version()} != expected {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


_dbg = None


def load_debug() -> C.CDLL:
    """The test library: the product objects plus the ``xattn_debug_*`` entry points (tests only; a second, self-contained
    library with its own error text — read it with ``debug_check``)."""
    global _dbg
    if _dbg is not None:
        return _dbg
    if not DEBUG_LIB_PATH.exists():
        raise RuntimeError(f"{DEBUG_LIB_PATH} not found: run `python __graft_entry__.py build`")
    lib = C.CDLL(str(DEBUG_LIB_PATH))
    for name, (res, args) in {**_EXPORTS, **_DEBUG_EXPORTS}.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _dbg = lib
    return lib


def debug_check(status: int) -> None:
    if status != OK:
        msg = load_debug().xattn_last_error().decode("utf-8", "replace")
        raise (ValueError if status in (ERR_SHAPE, ERR_DTYPE, ERR_ALIGN, ERR_NULL) else RuntimeError)(
            f"xattn_fusion_debug: {msg} (status {status})")


def last_error() -> str:
    return load().xattn_last_error().decode("utf-8", "replace")


def check(status: int) -> None:
    """Map an xattn_status to the reference's error convention (ValueError for bad arguments)."""
    if status == OK:
        return
    msg = f"xattn_fusion: {last_error()} (status {status})"
    if status in (ERR_SHAPE, ERR_DTYPE, ERR_ALIGN, ERR_NULL):
        raise ValueError(msg)
    raise RuntimeError(msg)


def stage_timing(enable: bool) -> None:
    """Turn the per-kernel CUDA-event brackets of the bf16 forward on or off (diagnostics; see xattn_stage_timing)."""
    check(load().xattn_stage_timing(1 if enable else 0))


def stage_timing_read() -> dict:
    """{stage name: (total ms, brackets)} since the last enable/read; waits for the recorded events."""
    lib = load()
    ms = (C.c_double * 16)()
    n = (C.c_int64 * 16)()
    k = lib.xattn_stage_timing_read(ms, n, 16)
    return {lib.xattn_stage_name(i).decode(): (float(ms[i]), int(n[i])) for i in range(k)}


def launch_count() -> int:
    return int(load().xattn_launch_count())


def xattn_param_layout(cfg: XAttnConfig):
    n = load().xattn_param_count(C.byref(cfg))
    off = (C.c_int64 * n)()
    size = (C.c_int64 * n)()
    total = C.c_int64(0)
    check(load().xattn_param_layout(C.byref(cfg), off, size, C.byref(total)))
    return list(off), list(size), int(total.value)


def flat_param_layout(cfg: FlatHeadConfig):
    off = (C.c_int64 * FP_COUNT)()
    size = (C.c_int64 * FP_COUNT)()
    total = C.c_int64(0)
    check(load().flat_head_param_layout(C.byref(cfg), off, size, C.byref(total)))
    return list(off), list(size), int(total.value)


def workspace_bytes(cfg: XAttnConfig) -> int:
    n = C.c_int64(0)
    check(load().xattn_workspace_bytes(C.byref(cfg), C.byref(n)))
    return int(n.value)


def packed_bytes(cfg: XAttnConfig) -> int:
    n = C.c_int64(0)
    check(load().xattn_packed_bytes(C.byref(cfg), C.byref(n)))
    return int(n.value)


def flat_workspace_bytes(cfg: FlatHeadConfig) -> int:
    n = C.c_int64(0)
    check(load().flat_head_workspace_bytes(C.byref(cfg), C.byref(n)))
    return int(n.value)
