"""Host-side mirror of the reference's fusion module, backed by the sm_100a kernels.

``XAttnFusionHead`` is the hot path proper: ``forward(video_emb [B,T,512], audio_emb [B,Ta,768]) -> logits [B,C]``
(reference ``src/models/fusion.py:372-411``).  ``FusionModel`` keeps the reference constructor and
``forward(video, audio)`` signature (``fusion.py:188-213``, ``:356``) so it drops into ``src/train.py:211``,
``src/eval.py:50``, ``src/optimized_runtime.py:106`` and ``backend/app/infer.py:98``: the two backbones run in
PyTorch exactly as the reference calls them (``fusion.py:368-370``, ``:377``) and everything after runs in the
CUDA extension.

Parameters are created by the same ``nn`` containers, in the same order and under the same names as
``fusion.py:269-326``, so ``state_dict()`` is key- and shape-identical and same-seed initialisation is
bit-identical.  At first use they are re-pointed into one flat fp32 buffer (layout: ``xattn_param_layout`` in
``include/xattn_fusion.h``), which is what the kernels, the gradient all-reduce and the fused Adam consume.

No fallback: inputs must be CUDA tensors on an sm_100 device, otherwise the call raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch
from torch import nn

from . import _lib

_XATTN_MODES = {"xattn", "xattn_concat", "xattn_gated"}


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _io_dtype(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return _lib.DTYPE_F32
    if t.dtype == torch.bfloat16:
        return _lib.DTYPE_BF16
    raise ValueError(f"embeddings must be float32 or bfloat16, got {t.dtype}")


class StochasticDepth(nn.Module):
    """Placeholder kept so module names match the reference (``fusion.py:11-26``); the drop-path itself is
    applied inside the residual+LayerNorm kernel from ``drop_prob``."""

    def __init__(self, drop_prob: float = 0.0) -> None:
        super().__init__()
        self.drop_prob = float(max(0.0, min(1.0, drop_prob)))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # pragma: no cover - not on the kernel path
        raise RuntimeError("StochasticDepth is fused into the CUDA residual+LayerNorm kernel")


class TemporalAttentionPooling(nn.Module):
    """Parameter container of the learned pooling (``temporal.py:9-26``): LayerNorm -> Linear(dim, dim/2) -> GELU -> Dropout ->
    Linear(dim/2, 1); the arithmetic runs in ``attn_pool_kernel``."""

    def __init__(self, dim: int, dropout: float = 0.1) -> None:
        super().__init__()
        hidden_dim = max(1, dim // 2)
        self.score = nn.Sequential(nn.LayerNorm(dim), nn.Linear(dim, hidden_dim), nn.GELU(), nn.Dropout(dropout),
                                   nn.Linear(hidden_dim, 1))

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # pragma: no cover - not on the kernel path
        raise RuntimeError("TemporalAttentionPooling runs inside the CUDA forward (attn_pool_kernel)")


class TemporalTransformerPooling(nn.Module):
    """Parameter container of ``temporal.py:46-75``: sinusoidal positions (no parameters; the reference's ``pe`` buffer is
    non-persistent), ``nn.TransformerEncoder`` of pre-norm GELU layers, then ``TemporalAttentionPooling``.  Built from the same
    torch modules as the reference so keys, shapes and same-seed initialisation are identical.  Up to 8 layers."""

    def __init__(self, dim: int, num_heads: int = 4, num_layers: int = 1, dropout: float = 0.1, mlp_ratio: float = 4.0) -> None:
        super().__init__()
        if not 1 <= num_layers <= _lib.MAX_POOL_LAYERS:
            raise ValueError(f"the B200 transformer pooler runs 1..{_lib.MAX_POOL_LAYERS} encoder layers, got {num_layers}")
        if max(dim * 2, int(dim * mlp_ratio)) != 4 * dim:
            raise ValueError("the B200 transformer pooler is built for mlp_ratio=4")
        layer = nn.TransformerEncoderLayer(d_model=dim, nhead=num_heads, dim_feedforward=4 * dim, dropout=dropout, activation="gelu",
                                           batch_first=True, norm_first=True)
        self.encoder = nn.TransformerEncoder(layer, num_layers=num_layers)
        self.pool = TemporalAttentionPooling(dim=dim, dropout=dropout)
        self.num_heads = num_heads
        self.num_layers = num_layers

    def forward(self, x: torch.Tensor) -> torch.Tensor:  # pragma: no cover - not on the kernel path
        raise RuntimeError("TemporalTransformerPooling runs inside the CUDA forward")


class TemporalPooler(nn.Module):
    """``mean`` (``temporal.py:91-92, 108-109``; no parameters), ``attn`` (``temporal.py:93-94``: the learned pooling, forward
    and backward in the fp32 / tf32 kernel family) and ``transformer`` (``temporal.py:95-102``: ``num_layers`` pre-norm encoder
    layers in front of the learned pooling; forward and backward in the same family)."""

    def __init__(self, dim: int, mode: str = "mean", dropout: float = 0.1, num_heads: int = 4, num_layers: int = 1, **_unused) -> None:
        super().__init__()
        if mode == "mean":
            self.pool = None
        elif mode == "attn":
            self.pool = TemporalAttentionPooling(dim=dim, dropout=dropout)
        elif mode == "transformer":
            self.pool = TemporalTransformerPooling(dim=dim, num_heads=num_heads, num_layers=num_layers, dropout=dropout)
        else:
            raise ValueError(f"Unsupported temporal pooling mode: {mode}")
        self.mode = mode


class EmotionPriorBiasAdapter(nn.Module):
    """Parameter container with the reference's names and init order (``fusion.py:156-168``)."""

    def __init__(self, token_dim: int, prior_dim: int, hidden_dim: int, dropout: float = 0.1) -> None:
        super().__init__()
        self.prior_net = nn.Sequential(
            nn.Linear(token_dim * 2, hidden_dim), nn.ReLU(inplace=True), nn.Dropout(dropout), nn.Linear(hidden_dim, prior_dim))
        self.v_query_bias = nn.Linear(token_dim + prior_dim, 1)
        self.a_key_bias = nn.Linear(token_dim + prior_dim, 1)
        self.a_query_bias = nn.Linear(token_dim + prior_dim, 1)
        self.v_key_bias = nn.Linear(token_dim + prior_dim, 1)
        self.bias_scale = nn.Parameter(torch.tensor(1.0, dtype=torch.float32))
        self.dropout_p = float(dropout)


class _XAttnFunction(torch.autograd.Function):
    """Autograd bridge: forward saves activations in the kernel workspace, backward calls ``xattn_backward``."""

    @staticmethod
    def forward(ctx, owner, video_emb, audio_emb, *params):
        logits, state = owner._launch_forward(video_emb, audio_emb, save=True)
        ctx.owner = owner
        ctx.state = state
        ctx.params = params
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        owner = ctx.owner
        # ctx.needs_input_grad: (owner, video_emb, audio_emb, *params).  Stage 1 freezes the backbones (train.py:808-812), so the
        # embeddings need no gradient; Stage 2 (train.py:815-830) unfreezes backbone layers and loss.backward() must reach them.
        need_v, need_a = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        gflat, d_video, d_audio = owner._launch_backward(ctx.state, dlogits.contiguous().float(), input_grads=(need_v, need_a))
        grads = []
        for p, (off, n) in zip(ctx.params, owner._xattn_slices):
            grads.append(gflat[off:off + n].view_as(p) if n > 0 else None)
        ctx.state = None
        return (None, d_video, d_audio, *grads)


def softmax_top1(logits: torch.Tensor, probs: Optional[torch.Tensor] = None, top1: Optional[torch.Tensor] = None,
                 top1_prob: Optional[torch.Tensor] = None):
    """``softmax(dim=1)`` (optimized_runtime.py:107), the row argmax and the probability at it (inference_worker.py:137-142)
    in one launch (``xattn_softmax_top1``).  ``logits`` [B,C] fp32 CUDA; outputs are allocated unless given (``probs`` may be
    ``logits`` itself).  Returns ``(probs, top1 int32 [B], top1_prob [B])``."""
    if not logits.is_cuda or logits.dtype != torch.float32 or logits.ndim != 2 or not logits.is_contiguous():
        raise ValueError("softmax_top1 expects a contiguous fp32 CUDA tensor [B,C]")
    B, C_ = logits.shape
    if probs is None:
        probs = torch.empty_like(logits)
    if top1 is None:
        top1 = torch.empty((B,), dtype=torch.int32, device=logits.device)
    if top1_prob is None:
        top1_prob = torch.empty((B,), dtype=torch.float32, device=logits.device)
    with torch.cuda.device(logits.device):
        _lib.check(_lib.load().xattn_softmax_top1(_ptr(logits), B, C_, _ptr(probs), _ptr(top1), _ptr(top1_prob),
                                                  _stream_ptr(logits.device)))
    return probs, top1, top1_prob


class HostResult:
    """Pending result of ``predict_host_async``: ``result()`` waits for the call's device->host copy and returns fresh host
    tensors (logits / probabilities ``[B,C]``, or ``(probs, top1, top1_prob)``)."""

    def __init__(self, event: "torch.cuda.Event", host: torch.Tensor, top1) -> None:
        self._event, self._host, self._top1 = event, host, top1

    def done(self) -> bool:
        return self._event.query()

    def result(self):
        self._event.synchronize()
        out = self._host.clone()
        return out if self._top1 is None else (out, self._top1[0].clone(), self._top1[1].clone())


class GraphedForward:
    """One CUDA graph of the bf16 forward at a fixed batch size ("CUDA-graph per batch size").

    Below a few hundred clips a forward is bound by the host issuing its 9 launches, not by the device.  The graph is
    captured once over static device buffers (``video``, ``audio`` -> ``logits`` / ``probs``); a call is then one
    ``cudaGraphLaunch``.  Fill ``video`` / ``audio`` (e.g. ``copy_(host, non_blocking=True)``), call the object, read
    ``logits`` — all on the current stream.  The graph holds the packed-weight pointers: it is only valid while ``key``
    matches the head's ``_xattn_packed_key`` (``XAttnKernelsMixin.graphed_forward`` re-captures when parameters change).
    """

    def __init__(self, head: "XAttnKernelsMixin", B: int, T: int, Ta: int, dtype: torch.dtype, device: torch.device) -> None:
        self.B, self.key = B, None
        with torch.cuda.device(device), torch.no_grad():
            self.video = torch.zeros((B, T, head.v_dim), dtype=dtype, device=device)
            self.audio = torch.zeros((B, Ta, head._kernel_audio_dim), dtype=dtype, device=device)
            self.logits = torch.empty((B, head.num_classes), dtype=torch.float32, device=device)
            side = torch.cuda.Stream(device)
            side.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(side):     # warm-up outside the capture: packs weights, sets kernel attributes
                for _ in range(2):
                    head._launch_forward(self.video, self.audio, save=False, training=False, out=self.logits)
            torch.cuda.current_stream(device).wait_stream(side)
            self.key = head._xattn_packed_key
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                head._launch_forward(self.video, self.audio, save=False, training=False, out=self.logits)
                self.probs = torch.empty_like(self.logits)
                self.top1 = torch.empty((B,), dtype=torch.int32, device=device)
                self.top1_prob = torch.empty((B,), dtype=torch.float32, device=device)
                softmax_top1(self.logits, self.probs, self.top1, self.top1_prob)

    def __call__(self) -> torch.Tensor:
        self.graph.replay()
        return self.logits


class XAttnKernelsMixin:
    """Builds the xattn parameters under the reference's names and runs them through the C ABI."""

    # ---- construction (mirrors fusion.py:263-327) -------------------------------------------------------------
    def _build_xattn(self, *, v_dim: int, audio_sequence_dim: int, num_classes: int, common_dim: int, xattn_head: str,
                     d_model: int, num_heads: int, audio_n_mels: int, xattn_attn_dropout: float,
                     xattn_stochastic_depth: float, temporal_pooling: str, use_emotion_prior: bool, prior_dim: int,
                     prior_hidden: int, prior_dropout: float, precision: str = "auto", temporal_dropout: float = 0.1,
                     audio_frontend: str = "seq", temporal_num_heads: int = 4, temporal_num_layers: int = 1) -> None:
        if xattn_head not in ("concat", "gated"):
            raise ValueError(f"Unknown xattn head: {xattn_head}")
        if precision not in ("auto", "fp32", "bf16", "tf32"):
            raise ValueError(f"precision must be auto|fp32|bf16|tf32, got {precision}")
        if audio_frontend not in ("seq", "mel"):
            raise ValueError(f"audio_frontend must be seq|mel, got {audio_frontend}")
        self.v_dim = v_dim
        self.audio_sequence_dim = audio_sequence_dim
        # "seq": audio_seq_proj over encoder states (fusion.py:374-378).  "mel": audio_time_conv over a mel spectrogram
        # (:379-384) — run as the same Linear over time-unfolded rows (csrc/mel_frontend.cu), so the kernels see
        # Da = 3*n_mels and the audio_seq_proj slots of the flat buffer hold audio_time_conv.{weight,bias}.
        self.audio_frontend = audio_frontend
        self.audio_n_mels = audio_n_mels
        self._kernel_audio_dim = 3 * audio_n_mels if audio_frontend == "mel" else audio_sequence_dim
        self.a_dim = d_model
        self.d_model = d_model
        self.num_heads = num_heads
        self.num_classes = num_classes
        self.common_dim = common_dim
        self.precision = precision
        self.v_in_proj = nn.Linear(self.v_dim, d_model)
        self.a_in_proj = nn.Linear(self.a_dim, d_model)
        # both front ends exist in every checkpoint (fusion.py:273-274); the one not selected is never touched
        self.audio_time_conv = nn.Conv1d(in_channels=audio_n_mels, out_channels=self.a_dim, kernel_size=3, padding=1)
        self.audio_seq_proj = nn.Linear(self.audio_sequence_dim, self.a_dim)
        self.v2a_attn = nn.MultiheadAttention(embed_dim=d_model, num_heads=num_heads, dropout=xattn_attn_dropout, batch_first=True)
        self.a2v_attn = nn.MultiheadAttention(embed_dim=d_model, num_heads=num_heads, dropout=xattn_attn_dropout, batch_first=True)
        self.v_drop_path = StochasticDepth(drop_prob=xattn_stochastic_depth)
        self.a_drop_path = StochasticDepth(drop_prob=xattn_stochastic_depth)
        self.v_norm = nn.LayerNorm(d_model)
        self.a_norm = nn.LayerNorm(d_model)
        if use_emotion_prior:
            self.emotion_prior_bias = EmotionPriorBiasAdapter(d_model, prior_dim, prior_hidden, prior_dropout)
        else:
            self.emotion_prior_bias = None
        self.v_temporal_pool = TemporalPooler(dim=d_model, mode=temporal_pooling, dropout=temporal_dropout,
                                              num_heads=temporal_num_heads, num_layers=temporal_num_layers)
        self.a_temporal_pool = TemporalPooler(dim=d_model, mode=temporal_pooling, dropout=temporal_dropout,
                                              num_heads=temporal_num_heads, num_layers=temporal_num_layers)
        self._temporal_pooling = temporal_pooling
        self._temporal_num_heads = int(temporal_num_heads)
        self._temporal_num_layers = int(temporal_num_layers)
        self.xattn_head = xattn_head
        if xattn_head == "concat":
            self.xattn_mlp = nn.Sequential(
                nn.Linear(d_model * 2, common_dim), nn.ReLU(inplace=True), nn.Dropout(0.2), nn.Linear(common_dim, num_classes))
        else:
            self.xattn_gate = nn.Sequential(
                nn.Linear(d_model * 2, d_model), nn.ReLU(inplace=True), nn.Dropout(0.2), nn.Linear(d_model, 1), nn.Sigmoid())
            self.xattn_classifier = nn.Linear(d_model, num_classes)
            with torch.no_grad():  # fusion.py:338-344 — every Linear of the gate except the final Sigmoid module
                for layer in self.xattn_gate:
                    if isinstance(layer, nn.Linear):
                        layer.bias.fill_(-1.0)
        self.xattn_attn_dropout = float(xattn_attn_dropout)
        self.xattn_stochastic_depth = float(max(0.0, min(1.0, xattn_stochastic_depth)))
        self.xattn_head_dropout = 0.2
        self.xattn_prior_dropout = float(prior_dropout)
        self.xattn_pool_dropout = float(temporal_dropout)   # Dropout inside TemporalAttentionPooling.score (temporal.py:19)
        self._prior_dim = int(prior_dim) if use_emotion_prior else 0
        self._prior_hidden = int(prior_hidden) if use_emotion_prior else 0
        # kernel-side state
        self._xattn_flat: Optional[torch.Tensor] = None
        self._xattn_grad_flat: Optional[torch.Tensor] = None
        self._xattn_slices: list[tuple[int, int]] = []
        self._xattn_packed: Optional[torch.Tensor] = None
        self._xattn_packed_key = None
        self._xattn_epoch = 0
        self._rng_offset = 0

    # ---- config / parameter plumbing ------------------------------------------------------------------------------
    def _xattn_param_names(self) -> list[str]:
        tx = getattr(self, "_temporal_pooling", "mean") == "transformer"
        names = _lib.xattn_param_names(self.xattn_head, getattr(self, "_temporal_num_layers", 1) if tx else 1)
        if getattr(self, "audio_frontend", "seq") == "mel":
            names = [n.replace("audio_seq_proj.", "audio_time_conv.") for n in names]
        if getattr(self, "_temporal_pooling", "mean") == "transformer":   # the score MLP sits one module deeper (temporal.py:70)
            names = [n.replace("_temporal_pool.pool.score.", "_temporal_pool.pool.pool.score.") for n in names]
        return names

    def _xattn_named_params(self) -> list[Optional[nn.Parameter]]:
        """The Parameters in flat-buffer order.  Resolved once and re-validated per call by identity against the owning
        module's ``_parameters`` dict (cheap), so replacing a Parameter object is still noticed."""
        cache = getattr(self, "_xattn_param_cache", None)
        if cache is not None:
            for owner, key, p in cache:
                if owner is not None and owner._parameters.get(key) is not p:
                    cache = None
                    break
        if cache is None:
            cache = []
            for name in self._xattn_param_names():
                owner, _, key = name.rpartition(".")
                try:
                    mod = self.get_submodule(owner) if owner else self
                except AttributeError:
                    mod = None
                p = mod._parameters.get(key) if mod is not None else None
                cache.append((mod if p is not None else None, key, p))
            self._xattn_param_cache = cache
            self._xattn_layout = None
        return [p for _, _, p in cache]

    def _make_cfg(self, B: int, T: int, Ta: int, io_dtype: int, precision: int, training: bool, save: bool) -> _lib.XAttnConfig:
        cfg = _lib.XAttnConfig()
        cfg.B, cfg.T, cfg.Ta, cfg.Dv, cfg.Da = B, T, Ta, self.v_dim, self._kernel_audio_dim
        cfg.E, cfg.H = self.d_model, self.num_heads
        cfg.prior_dim, cfg.prior_hidden = self._prior_dim, self._prior_hidden
        cfg.common_dim, cfg.C = self.common_dim, self.num_classes
        cfg.head_kind = _lib.HEAD_CONCAT if self.xattn_head == "concat" else _lib.HEAD_GATED
        cfg.use_prior = 1 if self.emotion_prior_bias is not None else 0
        cfg.io_dtype, cfg.precision = io_dtype, precision
        cfg.training, cfg.save_for_backward = int(training), int(save)
        cfg.pool_mode = {"mean": _lib.POOL_MEAN, "attn": _lib.POOL_ATTN, "transformer": _lib.POOL_TRANSFORMER}[
            getattr(self, "_temporal_pooling", "mean")]
        cfg.pool_heads = getattr(self, "_temporal_num_heads", 4)
        cfg.pool_layers = getattr(self, "_temporal_num_layers", 1)
        cfg.p_attn, cfg.p_path = self.xattn_attn_dropout, self.xattn_stochastic_depth
        cfg.p_head, cfg.p_prior = self.xattn_head_dropout, self.xattn_prior_dropout
        cfg.p_pool = self.xattn_pool_dropout
        seed = torch.initial_seed() & 0xFFFFFFFFFFFFFFFF
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            # data parallel: every rank holds different clips, so every rank must draw its own dropout masks (the same seed on
            # all ranks would give shard r's clip i the mask of shard 0's clip i)
            seed ^= ((torch.distributed.get_rank() + 1) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        cfg.seed = seed
        cfg.offset = self._rng_offset
        return cfg

    def flat_parameters(self) -> torch.Tensor:
        """The flat fp32 parameter buffer (creating / refreshing it if parameters were moved or replaced)."""
        params = self._xattn_named_params()
        ref = next(p for p in params if p is not None)
        dev = ref.device
        layout = getattr(self, "_xattn_layout", None)
        if layout is None:   # depends on the module's dimensions only
            layout = _lib.xattn_param_layout(self._make_cfg(0, 1, 1, _lib.DTYPE_F32, _lib.DTYPE_F32, False, False))
            self._xattn_layout = layout
        offs, sizes, total = layout
        flat = self._xattn_flat
        ok = flat is not None and flat.device == dev and flat.numel() == total
        if ok:
            base = flat.data_ptr()
            for p, off, n in zip(params, offs, sizes):
                if n == 0:
                    continue
                if p is None or p.dtype != torch.float32 or p.numel() != n or p.data_ptr() != base + 4 * off or not p.is_contiguous():
                    ok = False
                    break
        if not ok:
            flat = torch.zeros(total, dtype=torch.float32, device=dev)
            with torch.no_grad():
                for p, off, n, name in zip(params, offs, sizes, self._xattn_param_names()):
                    if n == 0:
                        continue
                    if p is None or p.numel() != n:
                        raise ValueError(f"parameter {name} missing or wrong size (expected {n} elements)")
                    view = flat[off:off + n].view(p.shape)
                    view.copy_(p.detach().to(torch.float32))
                    p.data = view
            self._xattn_flat = flat
            self._xattn_grad_flat = None
            self._xattn_epoch += 1
        self._xattn_slices = list(zip(offs, sizes))
        return self._xattn_flat

    def flat_gradients(self, attach: bool = True) -> torch.Tensor:
        """Flat gradient buffer (same layout); with ``attach`` every parameter's ``.grad`` is a view into it."""
        flat = self.flat_parameters()
        g = self._xattn_grad_flat
        if g is None or g.device != flat.device or g.numel() != flat.numel():
            g = torch.zeros_like(flat)
            self._xattn_grad_flat = g
        if attach:
            for p, (off, n) in zip(self._xattn_named_params(), self._xattn_slices):
                if p is not None and n > 0 and p.requires_grad:
                    p.grad = g[off:off + n].view(p.shape)
        return g

    def mark_parameters_updated(self) -> None:
        """Call after modifying the flat buffer directly (fused Adam) so derived bf16 weights are re-packed."""
        self._xattn_epoch += 1

    def _packed_weights(self, cfg: _lib.XAttnConfig, flat: torch.Tensor) -> torch.Tensor:
        params = self._xattn_named_params()
        key = (flat.data_ptr(), self._xattn_epoch, tuple(0 if p is None else p._version for p in params))
        if self._xattn_packed is None or self._xattn_packed_key != key or self._xattn_packed.device != flat.device:
            nbytes = _lib.packed_bytes(cfg)
            packed = torch.empty(nbytes, dtype=torch.uint8, device=flat.device)
            _lib.check(_lib.load().xattn_pack_weights(C.byref(cfg), _ptr(flat), _ptr(packed), _stream_ptr(flat.device)))
            self._xattn_packed, self._xattn_packed_key = packed, key
        return self._xattn_packed

    # ---- launches ------------------------------------------------------------------------------------------------------
    def _check_inputs(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> None:
        if video_emb.ndim != 3 or audio_emb.ndim != 3:
            raise ValueError(
                f"xattn head expects video_emb [B,T,{self.v_dim}] and audio_emb [B,Ta,{self._kernel_audio_dim}], got "
                f"shapes {tuple(video_emb.shape)} and {tuple(audio_emb.shape)}")
        if video_emb.shape[0] != audio_emb.shape[0]:
            raise ValueError("video_emb and audio_emb disagree on the batch size")
        if video_emb.shape[2] != self.v_dim or audio_emb.shape[2] != self._kernel_audio_dim:
            raise ValueError(
                f"embedding dims ({video_emb.shape[2]}, {audio_emb.shape[2]}) != ({self.v_dim}, {self._kernel_audio_dim})")
        if not (video_emb.is_cuda and audio_emb.is_cuda):
            raise RuntimeError("the B200 xattn fusion head has no CPU path: inputs must be CUDA tensors on an sm_100 device")
        if video_emb.dtype != audio_emb.dtype:
            raise ValueError("video_emb and audio_emb must share a dtype")

    def _bf16_supported(self, T: int, Ta: int) -> bool:
        """Shapes the bf16 tensor-core kernels are specialised for (csrc/bf16_pack.cu: bf16_supported)."""
        hw = self.d_model if self.xattn_head == "gated" else self.common_dim
        if getattr(self, "_temporal_pooling", "mean") == "transformer":   # the encoder layers live in the fp32 kernel family
            return False
        return (self.d_model == 128 and self.num_heads == 4 and T <= 16 and Ta <= 160 and self.v_dim % 64 == 0
                and self._kernel_audio_dim % 64 == 0 and hw <= 256 and hw % 64 == 0 and self.num_classes <= 16
                and self._prior_hidden <= 128 and self._prior_dim <= 32)

    def _pick_precision(self, io_dtype: int, need_grad: bool, T: int = 16, Ta: int = 149) -> int:
        """fp32 SIMT kernels for training / gradient parity and for shapes outside the bf16 specialisation; the bf16
        tcgen05 path for bf16 embeddings at inference (``precision="bf16"`` forces it and raises on unsupported shapes);
        ``precision="tf32"``: the fp32 kernel family with every Linear on the tensor cores (tcgen05 kind::tf32 / kind::f16,
        fp32 accumulate) — the fast Stage-1 training step, gradients within ~1e-2 relative of the fp32 ones.
        All are CUDA paths of this library — there is no eager fallback."""
        if self.precision == "tf32":
            return _lib.PREC_TF32
        if need_grad or self.precision == "fp32":
            return _lib.DTYPE_F32
        if self.precision == "bf16":
            return _lib.DTYPE_BF16
        return _lib.DTYPE_BF16 if (io_dtype == _lib.DTYPE_BF16 and self._bf16_supported(T, Ta)) else _lib.DTYPE_F32

    def _launch_forward(self, video_emb: torch.Tensor, audio_emb: torch.Tensor, save: bool, training: Optional[bool] = None,
                        out: Optional[torch.Tensor] = None):
        self._check_inputs(video_emb, audio_emb)
        video_emb = video_emb.contiguous()
        audio_emb = audio_emb.contiguous()
        dev = video_emb.device
        training = self.training if training is None else training
        B, T, _ = video_emb.shape
        Ta = audio_emb.shape[1]
        io = _io_dtype(video_emb)
        precision = self._pick_precision(io, save or training, T, Ta)
        if precision == _lib.DTYPE_BF16 and io != _lib.DTYPE_BF16:   # precision="bf16" forced on fp32 embeddings
            video_emb, audio_emb, io = video_emb.bfloat16(), audio_emb.bfloat16(), _lib.DTYPE_BF16
        with torch.cuda.device(dev):
            flat = self.flat_parameters()
            if flat.device != dev:
                raise RuntimeError(f"parameters live on {flat.device}, inputs on {dev}")
            if training:
                self._rng_offset += 1
            cfg = self._make_cfg(B, T, Ta, io, precision, training, save)
            packed = self._packed_weights(cfg, flat) if precision == _lib.DTYPE_BF16 else None
            ws = torch.empty(_lib.workspace_bytes(cfg), dtype=torch.uint8, device=dev)
            logits = out if out is not None else torch.empty((B, self.num_classes), dtype=torch.float32, device=dev)
            if B == 0:   # empty tensors have a NULL data_ptr; the ABI wants non-NULL pointers even when B == 0
                dummy = torch.empty(16, dtype=torch.uint8, device=dev)
                pv = pa = pl = _ptr(dummy)
            else:
                pv, pa, pl = _ptr(video_emb), _ptr(audio_emb), _ptr(logits)
            _lib.check(_lib.load().xattn_forward(C.byref(cfg), _ptr(flat), _ptr(packed), pv, pa, pl, _ptr(ws), ws.numel(),
                                                 _stream_ptr(dev)))
        return logits, (cfg, video_emb, audio_emb, ws)

    def _launch_backward(self, state, dlogits: torch.Tensor, grads: Optional[torch.Tensor] = None, input_grads=None,
                         tail_stream: Optional["torch.cuda.Stream"] = None):
        """``xattn_backward`` into the flat gradient buffer.  With ``input_grads=(need_video, need_audio)`` returns
        ``(grads, d_video_emb, d_audio_emb)`` (``xattn_backward_ex``; gradients in the embeddings' dtype, None where not asked
        for).  ``tail_stream``: made to wait for the point where every gradient but the input projections' is final; returns
        ``(grads, tail_offset)`` then (the flat buffer from ``tail_offset`` on may be all-reduced on that stream)."""
        cfg, video_emb, audio_emb, ws = state
        dev = video_emb.device
        with torch.cuda.device(dev):
            flat = self.flat_parameters()
            if grads is None:
                grads = torch.empty_like(flat)
            if input_grads is None and tail_stream is None:
                _lib.check(_lib.load().xattn_backward(C.byref(cfg), _ptr(flat), _ptr(video_emb), _ptr(audio_emb), _ptr(dlogits),
                                                      _ptr(grads), _ptr(ws), ws.numel(), _stream_ptr(dev)))
                return grads
            need_v, need_a = input_grads if input_grads is not None else (False, False)
            dv = torch.empty(video_emb.shape, dtype=torch.float32, device=dev) if need_v else None
            da = torch.empty(audio_emb.shape, dtype=torch.float32, device=dev) if need_a else None
            tail_off = C.c_int64(0)
            _lib.check(_lib.load().xattn_backward_ex(
                C.byref(cfg), _ptr(flat), _ptr(video_emb), _ptr(audio_emb), _ptr(dlogits), _ptr(grads), _ptr(ws), ws.numel(),
                _stream_ptr(dev), None if tail_stream is None else C.c_void_p(tail_stream.cuda_stream), C.byref(tail_off),
                _ptr(dv), _ptr(da)))
        if input_grads is None:
            return grads, int(tail_off.value)
        if dv is not None:
            dv = dv.to(video_emb.dtype)
        if da is not None:
            da = da.to(audio_emb.dtype)
        return grads, dv, da

    def graphed_forward(self, B: int, T: int = 16, Ta: int = 149, dtype: torch.dtype = torch.bfloat16,
                        device: Optional[torch.device] = None) -> GraphedForward:
        """The cached ``GraphedForward`` for this batch size (captured on first use, re-captured when the parameters or the
        packed weights changed).  Inference only; shapes must be ones the bf16 tensor-core path takes."""
        if device is None:
            device = next(p for p in self._xattn_named_params() if p is not None).device
        if self._pick_precision(_lib.DTYPE_BF16 if dtype == torch.bfloat16 else _lib.DTYPE_F32, False, T, Ta) != _lib.DTYPE_BF16:
            raise ValueError("graphed_forward covers the bf16 tensor-core inference path only")
        cache = self.__dict__.setdefault("_xattn_graphs", {})
        k = (B, T, Ta, dtype, device)
        g = cache.get(k)
        if g is not None:   # still the same weights?  (cheap: data_ptr + versions, as _packed_weights keys them)
            flat = self.flat_parameters()
            params = self._xattn_named_params()
            key = (flat.data_ptr(), self._xattn_epoch, tuple(0 if p is None else p._version for p in params))
            if key != g.key:
                g = None
        if g is None:
            if len(cache) >= 64:
                cache.pop(next(iter(cache)))
            g = GraphedForward(self, B, T, Ta, dtype, device)
            cache[k] = g
        return g

    @torch.no_grad()
    def predict_host(self, video_host: torch.Tensor, audio_host: torch.Tensor, chunk: int = 1024,
                     probs: bool = False, graph_max: int = 256, top1: bool = False):
        """Batch path with HOST buffers, mirroring ``TorchModelRunner.predict_probs`` (``optimized_runtime.py:99-108``:
        ``.to(device)`` -> model -> ``softmax`` -> ``.cpu()``) and the worker's ``torch.stack(...).to(device)``
        (``inference_worker.py:131-134``).  The host batch (pinned memory for full PCIe rate) is cut into ``chunk``-clip
        pieces that are copied on a side stream into two device staging slots while the previous piece is computed;
        returns host logits (or probabilities) ``[B,C]`` after synchronising on the final device->host copy — a fresh tensor,
        as the reference's ``.cpu()`` gives (the pinned staging buffers are reused by later calls).
        ``top1=True`` returns ``(probs [B,C], top1 int32 [B], top1_prob [B])`` on the host — the three things the worker
        puts into its result (``inference_worker.py:136-143``), computed on the device in the same launch as the softmax.
        ``predict_host_async`` is the same without the final wait."""
        return self.predict_host_async(video_host, audio_host, chunk=chunk, probs=probs, graph_max=graph_max, top1=top1).result()

    @torch.no_grad()
    def predict_host_async(self, video_host: torch.Tensor, audio_host: torch.Tensor, chunk: int = 1024,
                           probs: bool = False, graph_max: int = 256, top1: bool = False) -> "HostResult":
        """``predict_host`` without the final synchronisation: enqueues H2D, forward and D2H and returns a ``HostResult``
        whose ``result()`` waits for this call's D2H only.  Two calls may be in flight (the pinned output buffers and the
        device staging slots alternate between calls), so a serving loop can issue batch i+1 — its H2D runs on the copy
        stream beside the compute of batch i — before it collects batch i.  A third call waits for the first one."""
        probs = probs or top1
        if video_host.is_cuda or audio_host.is_cuda:
            raise ValueError("predict_host takes host tensors; call forward() for device tensors")
        if video_host.ndim != 3 or audio_host.ndim != 3 or video_host.shape[0] != audio_host.shape[0]:
            raise ValueError("expected video [B,T,Dv] and audio [B,Ta,Da] host tensors with the same B")
        if video_host.dtype != audio_host.dtype:
            raise ValueError("video_emb and audio_emb must share a dtype")
        dev = next(p for p in self._xattn_named_params() if p is not None).device
        if dev.type != "cuda":
            raise RuntimeError("the B200 xattn fusion head has no CPU path: move the module to an sm_100 device")
        B, T_, Dv = video_host.shape
        Ta_, Da = audio_host.shape[1], audio_host.shape[2]
        # per-call output slot: pinned logits / top-1 buffers and the event that marks their D2H
        slots = self.__dict__.setdefault("_host_slots", [None, None])
        calls = self.__dict__.get("_host_calls", 0)
        self._host_calls = calls + 1
        si = calls & 1
        slot = slots[si]
        if slot is not None:
            slot["event"].synchronize()          # the call that used this slot two calls ago has delivered
        if slot is None or slot["out"].shape[0] < B or slot["out"].shape[1] != self.num_classes:
            cap = max(B, 1)
            slot = {"out": torch.empty((cap, self.num_classes), dtype=torch.float32).pin_memory(),
                    "top1": torch.empty((cap,), dtype=torch.int32).pin_memory(),
                    "top1p": torch.empty((cap,), dtype=torch.float32).pin_memory(), "event": torch.cuda.Event()}
            slots[si] = slot
        host = slot["out"][:B]
        with torch.cuda.device(dev):
            cur = torch.cuda.current_stream(dev)
            if 0 < B <= min(graph_max, chunk) and video_host.dtype == torch.bfloat16 and self.precision in ("auto", "bf16") \
                    and self._bf16_supported(T_, Ta_):
                # worker-sized batches (inference_worker.py:33-34) are launch-bound: H2D, one graph launch, D2H.  Graphs are kept
                # per batch size rounded up to a multiple of 8 (clips are independent: the spare rows compute on stale data and
                # are not read), so a worker whose batches vary in size re-uses a handful of graphs.
                Bp = (B + 7) & ~7
                g = self.graphed_forward(Bp, T_, Ta_, video_host.dtype, dev)
                g.video[:B].copy_(video_host, non_blocking=True)
                g.audio[:B].copy_(audio_host, non_blocking=True)
                g()
                host.copy_((g.probs if probs else g.logits)[:B], non_blocking=True)
                if top1:
                    slot["top1"][:B].copy_(g.top1[:B], non_blocking=True)
                    slot["top1p"][:B].copy_(g.top1_prob[:B], non_blocking=True)
            else:
                key = (dev, chunk, T_, Dv, Ta_, Da, video_host.dtype)
                pipe = getattr(self, "_host_pipe", None)
                if pipe is None or pipe["key"] != key:
                    pipe = {"key": key, "stream": torch.cuda.Stream(dev), "n": 0,
                            "v": [torch.empty((chunk, T_, Dv), dtype=video_host.dtype, device=dev) for _ in range(2)],
                            "a": [torch.empty((chunk, Ta_, Da), dtype=audio_host.dtype, device=dev) for _ in range(2)],
                            "ready": [torch.cuda.Event() for _ in range(2)], "done": [torch.cuda.Event() for _ in range(2)]}
                    self._host_pipe = pipe
                side = pipe["stream"]
                logits = torch.empty((B, self.num_classes), dtype=torch.float32, device=dev)
                for s in range(0, B, chunk):
                    e = min(B, s + chunk)
                    n = pipe["n"]                 # pieces issued so far, over ALL calls: the staging slots alternate across calls
                    pipe["n"] = n + 1
                    st = n & 1
                    with torch.cuda.stream(side):
                        if n >= 2:
                            side.wait_event(pipe["done"][st])      # the forward that read this slot two pieces ago is done
                        pipe["v"][st][: e - s].copy_(video_host[s:e], non_blocking=True)
                        pipe["a"][st][: e - s].copy_(audio_host[s:e], non_blocking=True)
                        pipe["ready"][st].record(side)
                    cur.wait_event(pipe["ready"][st])
                    self._launch_forward(pipe["v"][st][: e - s], pipe["a"][st][: e - s], save=False, training=False,
                                         out=logits[s:e])
                    pipe["done"][st].record(cur)
                if probs:
                    _, t1, t1p = softmax_top1(logits, logits)
                host.copy_(logits, non_blocking=True)
                if top1:
                    slot["top1"][:B].copy_(t1, non_blocking=True)
                    slot["top1p"][:B].copy_(t1p, non_blocking=True)
            slot["event"].record(cur)
        return HostResult(slot["event"], host, (slot["top1"][:B], slot["top1p"][:B]) if top1 else None)

    def mel_unfold(self, mel: torch.Tensor) -> torch.Tensor:
        """mel [B, 1, n_mels, Ta] or [B, n_mels, Ta] -> time-unfolded rows [B, Ta, 3*n_mels] (csrc/mel_frontend.cu), the
        operand that turns ``audio_time_conv`` (fusion.py:381-384) into the head's first audio Linear."""
        if mel.ndim == 4 and mel.shape[1] == 1:
            mel = mel[:, 0]                                                                         # fusion.py:382 (squeeze(1))
        if mel.ndim != 3 or mel.shape[1] != self.audio_n_mels:
            raise ValueError(f"expected mel [B,1,{self.audio_n_mels},Ta] or [B,{self.audio_n_mels},Ta], got {tuple(mel.shape)}")
        if not mel.is_cuda:
            raise RuntimeError("the B200 xattn fusion head has no CPU path: inputs must be CUDA tensors on an sm_100 device")
        if mel.dtype not in (torch.float32, torch.bfloat16):
            mel = mel.float()
        mel = mel.contiguous()
        B, n_mels, Ta = mel.shape
        out = torch.empty((B, Ta, 3 * n_mels), dtype=mel.dtype, device=mel.device)
        with torch.cuda.device(mel.device):
            _lib.check(_lib.load().xattn_mel_unfold(_ptr(mel), _ptr(out), _lib.DTYPE_BF16 if mel.dtype == torch.bfloat16 else _lib.DTYPE_F32,
                                                    B, n_mels, Ta, _stream_ptr(mel.device)))
        return out

    def xattn_forward_mel(self, video_emb: torch.Tensor, mel: torch.Tensor) -> torch.Tensor:
        """logits from video embeddings and a mel spectrogram (``audio_frontend="mel"``, fusion.py:379-411)."""
        if self.audio_frontend != "mel":
            raise ValueError('xattn_forward_mel needs a head built with audio_frontend="mel"')
        if mel.requires_grad and torch.is_grad_enabled():   # fusion.py:379-384 feeds the loader's spectrogram: data, not a model output
            raise RuntimeError("the mel front end has no input gradient (its input is the spectrogram itself, fusion.py:379-384)")
        return self.xattn_forward_embeddings(video_emb, self.mel_unfold(mel))

    def xattn_forward_embeddings(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        """logits from backbone embeddings; differentiable w.r.t. the fusion parameters and — when they require grad (Stage 2
        of the reference unfreezes backbone layers, train.py:815-830) — w.r.t. the embeddings (``xattn_backward_ex``)."""
        params = self._xattn_named_params()
        need_grad = torch.is_grad_enabled() and (any(p is not None and p.requires_grad for p in params)
                                                 or video_emb.requires_grad or audio_emb.requires_grad)
        if need_grad:
            self.flat_parameters()
            live = [p if p is not None else torch.empty(0) for p in params]
            return _XAttnFunction.apply(self, video_emb, audio_emb, *live)
        logits, _ = self._launch_forward(video_emb, audio_emb, save=False)
        return logits


class XAttnFusionHead(XAttnKernelsMixin, nn.Module):
    """The hot path as a stand-alone module: ``forward(video_emb, audio_emb) -> logits``.

    Same parameter names/shapes as the non-backbone part of the reference's ``FusionModel`` in ``mode="xattn"``
    (SURVEY 8b), so a reference checkpoint's fusion keys load with ``load_state_dict(strict=True)`` after
    filtering out ``audio_model.*`` / ``video_model.*``.
    """

    def __init__(self, num_classes: int = 8, video_dim: int = 512, audio_dim: int = 768, common_dim: int = 256,
                 xattn_head: str = "concat", d_model: int = 128, num_heads: int = 4, audio_n_mels: int = 64,
                 xattn_attn_dropout: float = 0.1, xattn_stochastic_depth: float = 0.1, temporal_pooling: str = "mean", temporal_dropout: float = 0.1,
                 xattn_use_emotion_prior: bool = False, xattn_emotion_prior_dim: int = 8,
                 xattn_emotion_prior_hidden_dim: int = 64, xattn_emotion_prior_dropout: float = 0.1,
                 precision: str = "auto", audio_frontend: str = "seq", temporal_num_heads: int = 4,
                 temporal_num_layers: int = 1) -> None:
        super().__init__()
        self.mode = "xattn"
        self._build_xattn(v_dim=video_dim, audio_sequence_dim=audio_dim, num_classes=num_classes, common_dim=common_dim,
                          xattn_head=xattn_head, d_model=d_model, num_heads=num_heads, audio_n_mels=audio_n_mels,
                          xattn_attn_dropout=xattn_attn_dropout, xattn_stochastic_depth=xattn_stochastic_depth,
                          temporal_pooling=temporal_pooling, use_emotion_prior=xattn_use_emotion_prior,
                          prior_dim=xattn_emotion_prior_dim, prior_hidden=xattn_emotion_prior_hidden_dim,
                          prior_dropout=xattn_emotion_prior_dropout, precision=precision, temporal_dropout=temporal_dropout,
                          audio_frontend=audio_frontend, temporal_num_heads=temporal_num_heads,
                          temporal_num_layers=temporal_num_layers)

    def forward(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        """``audio_emb``: encoder states [B,Ta,audio_dim] (``audio_frontend="seq"``) or a mel spectrogram
        [B,1,n_mels,Ta] / [B,n_mels,Ta] (``"mel"``)."""
        if self.audio_frontend == "mel":
            return self.xattn_forward_mel(video_emb, audio_emb)
        return self.xattn_forward_embeddings(video_emb, audio_emb)

    @torch.no_grad()
    def predict_probs(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        """Batch-path helper mirroring ``TorchModelRunner.predict_probs`` (``optimized_runtime.py:99-108``)."""
        return softmax_top1(self.forward(video_emb, audio_emb).contiguous())[0]


class ClipStyleAlignment(nn.Module):
    """Parameter container for the CLIP-style alignment (``fusion.py:127-135``); the projections run in the flat-head
    kernels, the contrastive loss (B x B) in ``FlatHeadKernelsMixin.flat_alignment_loss``."""

    def __init__(self, audio_dim: int, video_dim: int, align_dim: int, init_temperature: float = 0.07) -> None:
        super().__init__()
        self.audio_proj = nn.Linear(audio_dim, align_dim)
        self.video_proj = nn.Linear(video_dim, align_dim)
        safe_temp = max(float(init_temperature), 1e-3)
        self.logit_scale = nn.Parameter(torch.tensor(math.log(1.0 / safe_temp), dtype=torch.float32))


class _FlatFunction(torch.autograd.Function):
    """Autograd bridge of the flat concat / gated heads: forward keeps its activations in the kernel workspace, backward calls
    ``flat_head_backward`` (flat gradient buffer -> one view per parameter; input gradients when the embeddings require them)."""

    @staticmethod
    def forward(ctx, owner, video_emb, audio_emb, *params):
        logits, state = owner._flat_launch_forward(video_emb, audio_emb, save=True)
        ctx.owner, ctx.state = owner, state
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        owner = ctx.owner
        cfg, video_emb, audio_emb, ws, flat = ctx.state
        dev = video_emb.device
        need_v, need_a = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        with torch.cuda.device(dev):
            grads = torch.empty_like(flat)
            dv = torch.empty(video_emb.shape, dtype=torch.float32, device=dev) if need_v else None
            da = torch.empty(audio_emb.shape, dtype=torch.float32, device=dev) if need_a else None
            dl = dlogits.contiguous().float()
            _lib.check(_lib.load().flat_head_backward(C.byref(cfg), _ptr(flat), _ptr(video_emb), _ptr(audio_emb), _ptr(dl), _ptr(grads),
                                                      _ptr(dv), _ptr(da), _ptr(ws), ws.numel(), _stream_ptr(dev)))
        out = [grads[off:off + n].view(shape) if n > 0 else None for off, n, shape in owner._flat_slices]
        ctx.state = None
        return (None, None if dv is None else dv.to(video_emb.dtype), None if da is None else da.to(audio_emb.dtype), *out)


class _FlatAlignFunction(torch.autograd.Function):
    """Autograd bridge of the ClipStyleAlignment loss (``fusion.py:141-149``): ``flat_head_alignment_loss`` forward,
    ``flat_head_alignment_backward`` backward (gradients of the two alignment projections, ``logit_scale`` and the embeddings)."""

    @staticmethod
    def forward(ctx, owner, video_emb, audio_emb, logit_scale, *align_params):
        loss, state = owner._flat_launch_alignment(video_emb, audio_emb)
        ctx.owner, ctx.state = owner, state
        return loss

    @staticmethod
    def backward(ctx, dloss):
        owner = ctx.owner
        cfg, video_emb, audio_emb, flat, scale = ctx.state
        dev = video_emb.device
        need_v, need_a = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        with torch.cuda.device(dev):
            n = C.c_int64()
            _lib.check(_lib.load().flat_head_alignment_workspace_bytes(C.byref(cfg), C.byref(n)))
            ws = torch.empty(n.value, dtype=torch.uint8, device=dev)
            grads = torch.empty_like(flat)
            dscale = torch.empty((), dtype=torch.float32, device=dev)
            dv = torch.empty(video_emb.shape, dtype=torch.float32, device=dev) if need_v else None
            da = torch.empty(audio_emb.shape, dtype=torch.float32, device=dev) if need_a else None
            g = dloss.detach().to(device=dev, dtype=torch.float32).contiguous()
            _lib.check(_lib.load().flat_head_alignment_backward(C.byref(cfg), _ptr(flat), _ptr(scale), _ptr(video_emb), _ptr(audio_emb),
                                                                _ptr(g), _ptr(grads), _ptr(dscale), _ptr(dv), _ptr(da), _ptr(ws),
                                                                ws.numel(), _stream_ptr(dev)))
        out = [grads[off:off + n_].view(shape) for off, n_, shape in owner._flat_slices[:4]]   # FP_ALIGN_A_W .. FP_ALIGN_V_B
        ctx.state = None
        return (None, None if dv is None else dv.to(video_emb.dtype), None if da is None else da.to(audio_emb.dtype), dscale, *out)


class FlatHeadKernelsMixin:
    """``mode in {"concat","gated"}`` heads (``fusion.py:226-260``, ``:413-435``) on the CUDA extension: eval and training forward
    (hidden-layer ``Dropout(0.2)``, ``ModalityDropout`` in gated mode), backward and the alignment loss with its backward."""

    def _build_flat(self, *, audio_dim: int, video_dim: int, num_classes: int, common_dim: int, mode: str,
                    fusion_align_mode: str, fusion_align_dim: int, fusion_align_temperature: float,
                    precision: str = "auto") -> None:
        if precision not in ("auto", "fp32", "bf16"):
            raise ValueError(f"precision must be auto|fp32|bf16, got {precision}")
        self._flat_precision = precision
        self._flat_audio_dim, self._flat_video_dim = audio_dim, video_dim
        self._flat_common, self._flat_classes = common_dim, num_classes
        fusion_audio_dim, fusion_video_dim = audio_dim, video_dim
        if fusion_align_mode == "clip":
            self.semantic_alignment = ClipStyleAlignment(audio_dim, video_dim, fusion_align_dim, fusion_align_temperature)
            fusion_audio_dim = fusion_video_dim = fusion_align_dim
            self._flat_align = fusion_align_dim
        else:
            self.semantic_alignment = None
            self._flat_align = 0
        self.audio_proj = nn.Linear(fusion_audio_dim, common_dim)
        self.video_proj = nn.Linear(fusion_video_dim, common_dim)
        if mode == "concat":
            self.fusion = nn.Sequential(
                nn.Linear(common_dim * 2, common_dim), nn.ReLU(inplace=True), nn.Dropout(0.2), nn.Linear(common_dim, num_classes))
        else:
            self.gate = nn.Sequential(
                nn.Linear(common_dim * 2, common_dim), nn.ReLU(inplace=True), nn.Dropout(0.2), nn.Linear(common_dim, 1), nn.Sigmoid())
            self.classifier = nn.Linear(common_dim, num_classes)
            with torch.no_grad():  # fusion.py:329-336
                for layer in self.gate:
                    if isinstance(layer, nn.Linear):
                        layer.bias.fill_(-1.0)
        self._flat_gated = mode == "gated"
        self._flat_buf: Optional[torch.Tensor] = None

    def _flat_cfg(self, B: int, T: int, Ta: int, io_dtype: int) -> _lib.FlatHeadConfig:
        cfg = _lib.FlatHeadConfig()
        cfg.B, cfg.T, cfg.Ta, cfg.Dv, cfg.Da = B, T, Ta, self._flat_video_dim, self._flat_audio_dim
        cfg.align_dim, cfg.common_dim, cfg.C = self._flat_align, self._flat_common, self._flat_classes
        cfg.gated, cfg.io_dtype = int(self._flat_gated), io_dtype
        cfg.precision = self._flat_pick_precision(io_dtype)
        return cfg

    flat_dropout_p = 0.2            # nn.Dropout(0.2) of fusion.2 / gate.2 (fusion.py:246, :255)
    modality_dropout_p = (0.2, 0.2)  # ModalityDropout(audio_dropout_p, video_dropout_p) (fusion.py:251)

    def _flat_train_cfg(self, cfg: _lib.FlatHeadConfig) -> _lib.FlatHeadConfig:
        """Training-mode fields of a config: fp32 kernels, the Philox state of the hidden-layer dropout, and — gated mode — the two
        ModalityDropout decisions, drawn exactly as the reference draws them (``torch.rand(1).item() < p``, audio first;
        ``fusion.py:47-53``), so the same torch seed drops the same modalities."""
        if cfg.precision == _lib.DTYPE_BF16:
            cfg.precision = _lib.DTYPE_F32
        if self.training:
            cfg.training = 1
            cfg.p_drop = float(self.flat_dropout_p)
            self._flat_rng_offset = getattr(self, "_flat_rng_offset", 0) + 1
            cfg.seed, cfg.offset = torch.initial_seed() & 0xFFFFFFFFFFFFFFFF, self._flat_rng_offset
            if self._flat_gated:
                cfg.drop_audio = int(torch.rand(1).item() < self.modality_dropout_p[0])
                cfg.drop_video = int(torch.rand(1).item() < self.modality_dropout_p[1])
        return cfg

    def _flat_tc_supported(self) -> bool:
        """Dimensions the tcgen05 projection GEMM takes (K % 64 == 0, N % 128 == 0)."""
        a, cm = self._flat_align, self._flat_common
        in_a = a if a > 0 else self._flat_audio_dim
        in_v = a if a > 0 else self._flat_video_dim
        return (self._flat_audio_dim % 64 == 0 and self._flat_video_dim % 64 == 0 and a % 128 == 0 and cm % 128 == 0
                and in_a % 64 == 0 and in_v % 64 == 0)

    def _flat_pick_precision(self, io_dtype: int) -> int:
        """``auto``: bf16 embeddings at supported dimensions run the tensor-core path (bf16 operands, fp32 accumulate);
        everything else the fp32 kernels.  ``bf16`` forces the tensor-core path (the library raises if it cannot)."""
        mode = getattr(self, "_flat_precision", "auto")
        if mode == "fp32":
            return _lib.DTYPE_F32
        if mode == "bf16":
            return _lib.DTYPE_BF16
        return _lib.DTYPE_BF16 if io_dtype == _lib.DTYPE_BF16 and self._flat_tc_supported() else _lib.DTYPE_F32

    def _flat_params(self, dev) -> torch.Tensor:
        own = dict(self.named_parameters(recurse=True))
        names = _lib.flat_param_names(self._flat_gated)
        offs, sizes, total = _lib.flat_param_layout(self._flat_cfg(0, 1, 1, _lib.DTYPE_F32))
        flat = self._flat_buf
        ok = flat is not None and flat.device == dev and flat.numel() == total
        if ok:
            for name, off, n in zip(names, offs, sizes):
                if n and (own[name].data_ptr() != flat.data_ptr() + 4 * off or own[name].dtype != torch.float32):
                    ok = False
                    break
        if not ok:
            flat = torch.zeros(total, dtype=torch.float32, device=dev)
            with torch.no_grad():
                for name, off, n in zip(names, offs, sizes):
                    if n == 0:
                        continue
                    p = own[name]
                    view = flat[off:off + n].view(p.shape)
                    view.copy_(p.detach().to(device=dev, dtype=torch.float32))
                    p.data = view
            self._flat_buf = flat
        self._flat_slices = [(off, n, tuple(own[name].shape) if n else ()) for name, off, n in zip(names, offs, sizes)]
        return self._flat_buf

    def _flat_named_params(self) -> list:
        own = dict(self.named_parameters(recurse=True))
        return [own[name] if n else None for name, (off, n, _) in zip(_lib.flat_param_names(self._flat_gated), self._flat_slices)]

    def flat_forward_sequences(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        """logits from sequence embeddings [B,T,Dv] / [B,Ta,Da]; the temporal mean is fused into the kernels."""
        if video_emb.ndim != 3 or audio_emb.ndim != 3:
            raise ValueError("expected video_emb [B,T,Dv] and audio_emb [B,Ta,Da]")
        if not (video_emb.is_cuda and audio_emb.is_cuda):
            raise RuntimeError("the B200 flat fusion heads have no CPU path: inputs must be CUDA tensors")
        if video_emb.shape[2] != self._flat_video_dim or audio_emb.shape[2] != self._flat_audio_dim:
            raise ValueError("embedding dims do not match the head")
        if video_emb.dtype != audio_emb.dtype or video_emb.shape[0] != audio_emb.shape[0]:
            raise ValueError("video_emb / audio_emb must share dtype and batch size")
        flat = self._flat_params(video_emb.device)
        params = self._flat_named_params()
        need_grad = torch.is_grad_enabled() and (video_emb.requires_grad or audio_emb.requires_grad or
                                                 any(p is not None and p.requires_grad for p in params))
        if need_grad:
            return _FlatFunction.apply(self, video_emb, audio_emb, *[p for p in params])
        if self.training:
            return self._flat_launch_forward(video_emb, audio_emb, save=True)[0]
        return self._flat_launch_forward(video_emb, audio_emb, save=False)[0]

    def _flat_launch_forward(self, video_emb: torch.Tensor, audio_emb: torch.Tensor, save: bool):
        """``flat_head_forward``.  ``save``: the fp32 kernels with the training-mode fields (their workspace is what
        ``flat_head_backward`` reads); else the inference path (bf16 tensor cores where the dimensions allow)."""
        video_emb, audio_emb = video_emb.detach().contiguous(), audio_emb.detach().contiguous()
        dev = video_emb.device
        with torch.cuda.device(dev):
            cfg = self._flat_cfg(video_emb.shape[0], video_emb.shape[1], audio_emb.shape[1], _io_dtype(video_emb))
            if save:
                cfg = self._flat_train_cfg(cfg)
            flat = self._flat_params(dev)
            ws = torch.empty(_lib.flat_workspace_bytes(cfg), dtype=torch.uint8, device=dev)
            logits = torch.empty((cfg.B, cfg.C), dtype=torch.float32, device=dev)
            _lib.check(_lib.load().flat_head_forward(C.byref(cfg), _ptr(flat), _ptr(video_emb), _ptr(audio_emb), _ptr(logits),
                                                     _ptr(ws), ws.numel(), _stream_ptr(dev)))
        return logits, (cfg, video_emb, audio_emb, ws, flat)

    def flat_alignment_loss(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        """The CLIP-style alignment loss of ``ClipStyleAlignment.forward`` (``fusion.py:141-149``) for a batch of sequence
        ``[B,T,Dv]`` / ``[B,Ta,Da]`` (or pooled ``[B,1,D]``) embeddings: a 0-dim CUDA tensor, no host sync.  It is a B x B
        cross-sample term — under batch sharding gather the embeddings first.  Differentiable (alignment projections,
        ``logit_scale``, embeddings) when autograd is recording."""
        sa = self.semantic_alignment
        if sa is not None and torch.is_grad_enabled() and (
                video_emb.requires_grad or audio_emb.requires_grad or any(p.requires_grad for p in sa.parameters())):
            self._flat_params(video_emb.device)
            return _FlatAlignFunction.apply(self, video_emb, audio_emb, sa.logit_scale, sa.audio_proj.weight, sa.audio_proj.bias,
                                            sa.video_proj.weight, sa.video_proj.bias)
        return self._flat_launch_alignment(video_emb, audio_emb)[0]

    @torch.no_grad()
    def _flat_launch_alignment(self, video_emb: torch.Tensor, audio_emb: torch.Tensor):
        if self.semantic_alignment is None:
            raise ValueError("the alignment loss needs fusion_align_mode=\"clip\"")
        if video_emb.ndim != 3 or audio_emb.ndim != 3 or video_emb.shape[0] != audio_emb.shape[0] or video_emb.shape[0] == 0:
            raise ValueError("expected non-empty video_emb [B,T,Dv] and audio_emb [B,Ta,Da] with the same B")
        if not (video_emb.is_cuda and audio_emb.is_cuda):
            raise RuntimeError("the B200 flat fusion heads have no CPU path: inputs must be CUDA tensors")
        if video_emb.dtype != audio_emb.dtype:
            raise ValueError("video_emb / audio_emb must share dtype")
        video_emb, audio_emb = video_emb.detach().contiguous(), audio_emb.detach().contiguous()
        dev = video_emb.device
        with torch.cuda.device(dev):
            cfg = self._flat_cfg(video_emb.shape[0], video_emb.shape[1], audio_emb.shape[1], _io_dtype(video_emb))
            if self._flat_precision == "auto":   # tensor cores (tf32 products) whenever the three contractions fit them
                cfg.precision = _lib.PREC_TF32 if self._flat_tc_supported() else _lib.DTYPE_F32
            elif self._flat_precision != "fp32":
                cfg.precision = _lib.PREC_TF32
            flat = self._flat_params(dev)
            n = C.c_int64()
            _lib.check(_lib.load().flat_head_alignment_workspace_bytes(C.byref(cfg), C.byref(n)))
            ws = torch.empty(n.value, dtype=torch.uint8, device=dev)
            loss = torch.empty((), dtype=torch.float32, device=dev)
            scale = self.semantic_alignment.logit_scale.detach().to(device=dev, dtype=torch.float32).contiguous()
            _lib.check(_lib.load().flat_head_alignment_loss(C.byref(cfg), _ptr(flat), _ptr(scale), _ptr(video_emb), _ptr(audio_emb),
                                                            _ptr(loss), _ptr(ws), ws.numel(), _stream_ptr(dev)))
        return loss, (cfg, video_emb, audio_emb, flat, scale)

    def flat_forward_pooled(self, audio_pooled: torch.Tensor, video_pooled: torch.Tensor) -> torch.Tensor:
        """logits from already pooled [B,Da] / [B,Dv] embeddings (what ``audio_model.encode`` returns)."""
        return self.flat_forward_sequences(video_pooled.unsqueeze(1), audio_pooled.unsqueeze(1))


class FlatFusionHead(FlatHeadKernelsMixin, nn.Module):
    """Stand-alone concat / gated head with the reference's parameter names (BASELINE config 5)."""

    def __init__(self, num_classes: int = 8, video_dim: int = 512, audio_dim: int = 768, common_dim: int = 256,
                 mode: str = "concat", fusion_align_mode: str = "clip", fusion_align_dim: int = 256,
                 fusion_align_temperature: float = 0.07, precision: str = "auto") -> None:
        super().__init__()
        if mode not in ("concat", "gated"):
            raise ValueError(f"Unknown fusion mode: {mode}")
        self.mode = mode
        self._build_flat(audio_dim=audio_dim, video_dim=video_dim, num_classes=num_classes, common_dim=common_dim, mode=mode,
                         fusion_align_mode=fusion_align_mode, fusion_align_dim=fusion_align_dim,
                         fusion_align_temperature=fusion_align_temperature, precision=precision)

    def forward(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        return self.flat_forward_sequences(video_emb, audio_emb)

    def alignment_loss(self, video_emb: torch.Tensor, audio_emb: torch.Tensor) -> torch.Tensor:
        return self.flat_alignment_loss(video_emb, audio_emb)


class FusionModel(XAttnKernelsMixin, FlatHeadKernelsMixin, nn.Module):
    """Drop-in for the reference ``FusionModel`` (``src/models/fusion.py:187-437``): same constructor arguments, same
    ``forward(video, audio)``, same ``state_dict`` keys, ``pop_alignment_loss()`` (``:346-349``).

    The backbones (``video_model.backbone``, ``audio_model.encode_sequence`` / ``.encode``) stay PyTorch modules the
    caller supplies; everything downstream runs in the sm_100a kernels.
    """

    def __init__(self, audio_model: nn.Module, video_model: nn.Module, num_classes: int, mode: str = "late",
                 common_dim: int = 256, xattn_head: str = "concat", d_model: int = 128, num_heads: int = 4,
                 audio_n_mels: int = 64, xattn_attn_dropout: float = 0.1, xattn_stochastic_depth: float = 0.1,
                 temporal_pooling: str = "mean", temporal_num_heads: int = 4, temporal_num_layers: int = 1,
                 temporal_dropout: float = 0.1, fusion_align_mode: str = "none", fusion_align_dim: int = 256,
                 fusion_align_temperature: float = 0.07, xattn_use_emotion_prior: bool = False,
                 xattn_emotion_prior_dim: int = 8, xattn_emotion_prior_hidden_dim: int = 64,
                 xattn_emotion_prior_dropout: float = 0.1, precision: str = "auto") -> None:
        super().__init__()
        self.audio_model = audio_model
        self.video_model = video_model
        self.mode = mode
        self.d_model = d_model
        self.num_heads = num_heads
        self.audio_n_mels = audio_n_mels
        self.fusion_align_mode = fusion_align_mode
        self.alignment_loss = None
        self.semantic_alignment = None
        self.xattn_use_emotion_prior = xattn_use_emotion_prior
        if mode in {"concat", "gated"}:
            self._build_flat(audio_dim=audio_model.embedding_dim, video_dim=video_model.embedding_dim, num_classes=num_classes,
                             common_dim=common_dim, mode=mode, fusion_align_mode=fusion_align_mode,
                             fusion_align_dim=fusion_align_dim, fusion_align_temperature=fusion_align_temperature,
                             precision=precision)
        if mode in _XATTN_MODES:
            self._build_xattn(v_dim=getattr(video_model, "embedding_dim", 512),
                              audio_sequence_dim=getattr(audio_model, "sequence_dim", d_model), num_classes=num_classes,
                              common_dim=common_dim, xattn_head=xattn_head, d_model=d_model, num_heads=num_heads,
                              audio_n_mels=audio_n_mels, xattn_attn_dropout=xattn_attn_dropout,
                              xattn_stochastic_depth=xattn_stochastic_depth, temporal_pooling=temporal_pooling,
                              use_emotion_prior=xattn_use_emotion_prior, prior_dim=xattn_emotion_prior_dim,
                              prior_hidden=xattn_emotion_prior_hidden_dim, prior_dropout=xattn_emotion_prior_dropout,
                              precision=precision, temporal_dropout=temporal_dropout,
                              audio_frontend="seq" if hasattr(audio_model, "encode_sequence") else "mel",
                              temporal_num_heads=temporal_num_heads, temporal_num_layers=temporal_num_layers)

    def pop_alignment_loss(self):
        """``fusion.py:346-349``.  The reference computes the loss inside every concat/gated forward; here the B x B term is
        evaluated only when it is asked for, from the embeddings the last forward kept."""
        loss = self.alignment_loss
        pending = self.__dict__.pop("_align_pending", None)
        if loss is None and pending is not None and self.semantic_alignment is not None:
            a_emb, v_emb = pending
            loss = self.flat_alignment_loss(v_emb.unsqueeze(1), a_emb.unsqueeze(1))
        self.alignment_loss = None
        return loss

    def forward(self, video: torch.Tensor, audio: torch.Tensor):
        self.alignment_loss = None
        self.__dict__.pop("_align_pending", None)
        if self.mode == "late":  # fusion.py:358-363 — backbone-only, no fusion math
            a_probs = torch.softmax(self.audio_model(audio), dim=1)
            v_probs = torch.softmax(self.video_model(video), dim=1)
            return (a_probs + v_probs) / 2.0
        if self.mode in _XATTN_MODES:
            b, t, c, h, w = video.shape
            v_feat = self.video_model.backbone(video.view(b * t, c, h, w)).view(b, t, self.v_dim)   # fusion.py:368-370
            if self.audio_frontend == "mel":                                                         # fusion.py:379-384
                return self.xattn_forward_mel(v_feat, audio)
            a_seq = self.audio_model.encode_sequence(audio)                                          # fusion.py:377
            return self.xattn_forward_embeddings(v_feat, a_seq)
        if self.mode in {"concat", "gated"}:
            a_emb = self.audio_model.encode(audio)                                                   # fusion.py:414-415
            v_emb = self.video_model.encode(video)
            if self.semantic_alignment is not None:
                self._align_pending = (a_emb, v_emb)                                                 # fusion.py:417-418, evaluated lazily
            return self.flat_forward_pooled(a_emb, v_emb)
        raise ValueError(f"Unknown fusion mode: {self.mode}")
