"""CPU port (PyTorch eager, fp32) of the reference's xattn fusion path.

TEST INFRASTRUCTURE ONLY — see the header of ``oracle/xattn_oracle.py``.  Used by
(1) ``bench.py``'s ``cpu_baseline`` leg and ``--impl reference`` arm: the reference is pure Python on
stock torch ops, so this port dispatches the very same ATen CPU kernels (``addmm``, the
``need_weights`` branch of ``F.multi_head_attention_forward``, ``layer_norm``, ``mean``) in the same order
as ``/root/reference/src/models/fusion.py:372-411`` — ``/root/reference`` itself is not present on the
GPU box; (2) gradient parity tests: autograd through this port is the gradient oracle.

It is a functional restatement over a plain ``state_dict`` (no nn.Module tree), validated
against the live reference module by ``oracle/gen_golden.py`` / ``tests/test_oracle_golden.py``.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


def _prior_bias(sd: Dict[str, Tensor], v: Tensor, a: Tensor, p_drop: float, training: bool, relu_mask: Optional[Tensor] = None):
    """reference fusion.py:170-184 — including the expand+cat materialisation the reference does.  `relu_mask` (tests only):
    the ReLU's on/off pattern is taken from it instead of the sign of the pre-activation (see xattn_forward_tf32)."""
    pre = "emotion_prior_bias."
    pooled = torch.cat([v.mean(dim=1), a.mean(dim=1)], dim=-1)
    hid = F.linear(pooled, sd[pre + "prior_net.0.weight"], sd[pre + "prior_net.0.bias"])
    hid = F.relu(hid) if relu_mask is None else hid * relu_mask.to(hid.dtype)
    hid = F.dropout(hid, p_drop, training)
    prior = F.linear(hid, sd[pre + "prior_net.3.weight"], sd[pre + "prior_net.3.bias"])

    def token_bias(q: Tensor, k: Tensor, qh: str, kh: str) -> Tensor:
        qp = prior.unsqueeze(1).expand(-1, q.size(1), -1)
        kp = prior.unsqueeze(1).expand(-1, k.size(1), -1)
        qs = F.linear(torch.cat([q, qp], dim=-1), sd[pre + qh + ".weight"], sd[pre + qh + ".bias"]).squeeze(-1)
        ks = F.linear(torch.cat([k, kp], dim=-1), sd[pre + kh + ".weight"], sd[pre + kh + ".bias"]).squeeze(-1)
        return torch.tanh(qs.unsqueeze(-1) + ks.unsqueeze(-2)) * sd[pre + "bias_scale"]

    return (token_bias(v, a, "v_query_bias", "a_key_bias"), token_bias(a, v, "a_query_bias", "v_key_bias"))


def _mha(sd: Dict[str, Tensor], pre: str, q: Tensor, kv: Tensor, heads: int, bias: Optional[Tensor],
         p_attn: float, training: bool) -> Tensor:
    """nn.MultiheadAttention(batch_first=True) call as at fusion.py:394/:398 (need_weights default True)."""
    mask = None if bias is None else bias.repeat_interleave(heads, dim=0)        # fusion.py:351-354
    out, _ = F.multi_head_attention_forward(
        q.transpose(0, 1), kv.transpose(0, 1), kv.transpose(0, 1), q.shape[-1], heads,
        sd[pre + "in_proj_weight"], sd[pre + "in_proj_bias"], None, None, False, p_attn,
        sd[pre + "out_proj.weight"], sd[pre + "out_proj.bias"], training=training, key_padding_mask=None,
        need_weights=True, attn_mask=mask, average_attn_weights=True)
    return out.transpose(0, 1)


def _drop_path(x: Tensor, p: float, training: bool) -> Tensor:
    if p <= 0.0 or not training:
        return x
    keep = 1.0 - p
    mask = x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(keep)
    return x * mask / keep


def xattn_forward(sd: Dict[str, Tensor], video_emb: Tensor, audio_emb: Tensor, *, num_heads: int = 4,
                  head: str = "concat", use_prior: Optional[bool] = None, training: bool = False,
                  p_attn: float = 0.1, p_path: float = 0.1, p_head: float = 0.2, p_prior: float = 0.1) -> Tensor:
    if use_prior is None:
        use_prior = any(k.startswith("emotion_prior_bias.") for k in sd)
    v = F.linear(video_emb, sd["v_in_proj.weight"], sd["v_in_proj.bias"])
    a = F.linear(audio_emb, sd["audio_seq_proj.weight"], sd["audio_seq_proj.bias"])
    a = F.linear(a, sd["a_in_proj.weight"], sd["a_in_proj.bias"])
    b1 = b2 = None
    if use_prior:
        b1, b2 = _prior_bias(sd, v, a, p_prior, training)
    v2 = _mha(sd, "v2a_attn.", v, a, num_heads, b1, p_attn, training)
    v = F.layer_norm(v + _drop_path(v2, p_path, training), (v.shape[-1],), sd["v_norm.weight"], sd["v_norm.bias"])
    a2 = _mha(sd, "a2v_attn.", a, v, num_heads, b2, p_attn, training)
    a = F.layer_norm(a + _drop_path(a2, p_path, training), (a.shape[-1],), sd["a_norm.weight"], sd["a_norm.bias"])
    v_emb, a_emb = v.mean(dim=1), a.mean(dim=1)
    fused = torch.cat([v_emb, a_emb], dim=1)
    if head == "concat":
        hid = F.dropout(F.relu(F.linear(fused, sd["xattn_mlp.0.weight"], sd["xattn_mlp.0.bias"])), p_head, training)
        return F.linear(hid, sd["xattn_mlp.3.weight"], sd["xattn_mlp.3.bias"])
    if head == "gated":
        hid = F.dropout(F.relu(F.linear(fused, sd["xattn_gate.0.weight"], sd["xattn_gate.0.bias"])), p_head, training)
        g = torch.sigmoid(F.linear(hid, sd["xattn_gate.3.weight"], sd["xattn_gate.3.bias"]))
        return F.linear(g * v_emb + (1 - g) * a_emb, sd["xattn_classifier.weight"], sd["xattn_classifier.bias"])
    raise ValueError(f"unknown xattn head: {head}")


def flat_head_forward(sd: Dict[str, Tensor], audio_pooled: Tensor, video_pooled: Tensor, *, mode: str = "concat",
                      align: bool = True, return_align_loss: bool = False):
    """reference fusion.py:413-435 (+ :137-150), eval mode."""
    a, v = audio_pooled, video_pooled
    loss = None
    if align:
        a = F.linear(a, sd["semantic_alignment.audio_proj.weight"], sd["semantic_alignment.audio_proj.bias"])
        v = F.linear(v, sd["semantic_alignment.video_proj.weight"], sd["semantic_alignment.video_proj.bias"])
        if return_align_loss:
            scale = sd["semantic_alignment.logit_scale"].exp().clamp(max=100.0)
            lg = scale * (F.normalize(a, dim=-1) @ F.normalize(v, dim=-1).t())
            tgt = torch.arange(lg.size(0))
            loss = 0.5 * (F.cross_entropy(lg, tgt) + F.cross_entropy(lg.t(), tgt))
    a = F.linear(a, sd["audio_proj.weight"], sd["audio_proj.bias"])
    v = F.linear(v, sd["video_proj.weight"], sd["video_proj.bias"])
    cat = torch.cat([a, v], dim=1)
    if mode == "concat":
        out = F.linear(F.relu(F.linear(cat, sd["fusion.0.weight"], sd["fusion.0.bias"])), sd["fusion.3.weight"],
                       sd["fusion.3.bias"])
    elif mode == "gated":
        g = torch.sigmoid(F.linear(F.relu(F.linear(cat, sd["gate.0.weight"], sd["gate.0.bias"])),
                                   sd["gate.3.weight"], sd["gate.3.bias"]))
        out = F.linear(g * a + (1 - g) * v, sd["classifier.weight"], sd["classifier.bias"])
    else:
        raise ValueError(f"unknown flat head mode: {mode}")
    return (out, loss) if return_align_loss else out


def stage1_step(sd: Dict[str, Tensor], opt: torch.optim.Optimizer, video_emb: Tensor, audio_emb: Tensor,
                labels: Tensor, *, num_heads: int = 4, head: str = "concat", label_smoothing: float = 0.0,
                training: bool = True, **drop) -> Tensor:
    """One Stage-1 step as reference train.py:204-228 (zero_grad, forward, CE, backward, optimizer.step)."""
    opt.zero_grad()
    logits = xattn_forward(sd, video_emb, audio_emb, num_heads=num_heads, head=head, training=training, **drop)
    loss = F.cross_entropy(logits, labels, label_smoothing=label_smoothing)
    loss.backward()
    opt.step()
    return loss.detach()


# ---------------------------------------------------------------------------------------------------------------------
# Quantisation-matched oracle of the tensor-core TRAINING path (precision="tf32", csrc/gemm_tf32.cu).
# The reference computes every Linear in fp32 (it never enables allow_tf32); the B200 training path feeds the same
# fp32 tensors to tcgen05 kind::tf32, which keeps 10 mantissa bits of each operand by TRUNCATION and accumulates in fp32
# (with bf16 embeddings the two first-layer products run as kind::f16 on bf16-rounded weights / gradients).  This port
# applies exactly those operand roundings — and nothing else — to the reference's op sequence, so a comparison with it
# isolates kernel bugs from the (documented) precision choice.  Which products are rounded follows the shape predicates
# of csrc/gemm_tf32.cu (gemm_tf32_rows_ok / gemm_cols_ok); everything else (prior MLP, score heads, classifier,
# softmax, LayerNorm, and the attention contractions unless attn_tf32 is set) stays fp32.
# ---------------------------------------------------------------------------------------------------------------------
def trunc_tf32(x: Tensor) -> Tensor:
    """fp32 -> tf32 the way the tensor core reads it: the low 13 mantissa bits are dropped."""
    return (x.detach().contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def _rn_bf16(x: Tensor) -> Tensor:
    return x.detach().bfloat16().float()


class _QLinear(torch.autograd.Function):
    """y = x·Wᵀ + b with the operand roundings of the tensor-core path in forward, dX and dW.  `first_bf16`: the layer reads
    bf16 embeddings (x is exactly bf16-representable): W and dY are rounded to bf16 (round-to-nearest), no dX."""

    @staticmethod
    def forward(ctx, x, w, b, first_bf16):
        n, k = w.shape
        ctx.first_bf16 = first_bf16
        ctx.save_for_backward(x, w)
        x2 = x.reshape(-1, k)
        if first_bf16:
            y = x2 @ _rn_bf16(w).t()
        elif n % 128 == 0 and k % 32 == 0:          # gemm_tf32_rows_ok(M, N, K)
            y = trunc_tf32(x2) @ trunc_tf32(w).t()
        else:
            y = x2 @ w.t()
        return (y + b).reshape(*x.shape[:-1], n)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        n, k = w.shape
        dy2, x2 = dy.reshape(-1, n), x.reshape(-1, k)
        if ctx.first_bf16:
            dx = None
            dw = _rn_bf16(dy2).t() @ x2
        else:
            if k % 128 == 0 and n % 32 == 0:        # dX[M,K] = dY[M,N]·W[N,K]: gemm_tf32_rows_ok(M, K, N)
                dx = trunc_tf32(dy2) @ trunc_tf32(w)
            else:
                dx = dy2 @ w
            dx = dx.reshape(x.shape)
            if n % 128 == 0 and k % 128 == 0:       # dW[N,K] = dYᵀ·X: gemm_cols_ok(M, N, K)
                dw = trunc_tf32(dy2).t() @ trunc_tf32(x2)
            else:
                dw = dy2.t() @ x2
        return dx, dw, dy2.sum(dim=0), None


class _QMatmul(torch.autograd.Function):
    """a @ b (batched) with both operands truncated to tf32 in the product and in both gradient products (attn_tf32)."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return trunc_tf32(a) @ trunc_tf32(b)

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        return trunc_tf32(dc) @ trunc_tf32(b).transpose(-1, -2), trunc_tf32(a).transpose(-1, -2) @ trunc_tf32(dc)


def _mha_q(sd, pre, q_in, kv_in, heads, bias, attn_tf32):
    """The need_weights branch of F.multi_head_attention_forward (eval: no dropout) restated over _QLinear."""
    E = q_in.shape[-1]
    dh = E // heads
    w, b = sd[pre + "in_proj_weight"], sd[pre + "in_proj_bias"]
    q = _QLinear.apply(q_in, w[:E], b[:E], False)
    kv = _QLinear.apply(kv_in, w[E:], b[E:], False)
    B, L, S = q.shape[0], q.shape[1], kv.shape[1]
    qh = (q * (1.0 / dh ** 0.5)).view(B, L, heads, dh).transpose(1, 2)
    kh = kv[..., :E].reshape(B, S, heads, dh).transpose(1, 2)
    vh = kv[..., E:].reshape(B, S, heads, dh).transpose(1, 2)
    mm = _QMatmul.apply if attn_tf32 else torch.matmul
    sc = mm(qh, kh.transpose(-1, -2))
    if bias is not None:
        sc = sc + bias.unsqueeze(1)                 # head-independent (fusion.py:351-354)
    p = torch.softmax(sc, dim=-1)
    o = mm(p, vh).transpose(1, 2).reshape(B, L, E)
    return _QLinear.apply(o, sd[pre + "out_proj.weight"], sd[pre + "out_proj.bias"], False)


def xattn_forward_tf32(sd: Dict[str, Tensor], video_emb: Tensor, audio_emb: Tensor, *, num_heads: int = 4, head: str = "concat",
                       use_prior: Optional[bool] = None, bf16_io: bool = False, attn_tf32: bool = False,
                       taps: Optional[dict] = None, relu_masks: Optional[dict] = None) -> Tensor:
    """Eval-mode forward (dropouts identity) of fusion.py:372-411 with the tensor-core path's operand roundings.
    `relu_masks` = {"head": bool [B,HW], "prior": bool [B,Hd]}: pin the two ReLUs' on/off patterns (to the CUDA run's).  A
    tf32 operand that moves by one fp32 ulp can cross a truncation boundary and jump by 2^-10, so two correct
    implementations differ by ~1e-4 in the hidden pre-activations; a handful of the 24k ReLU units then sit on opposite
    sides of zero, and each such flip moves whole gradient rows by percents.  With the patterns pinned, what is compared is
    the smooth arithmetic of the kernels."""
    if use_prior is None:
        use_prior = any(k.startswith("emotion_prior_bias.") for k in sd)
    v = _QLinear.apply(video_emb, sd["v_in_proj.weight"], sd["v_in_proj.bias"], bf16_io)
    a = _QLinear.apply(audio_emb, sd["audio_seq_proj.weight"], sd["audio_seq_proj.bias"], bf16_io)
    if taps is not None:
        taps["v0"], taps["aseq"] = v.detach(), a.detach()
    a = _QLinear.apply(a, sd["a_in_proj.weight"], sd["a_in_proj.bias"], False)
    b1 = b2 = None
    if use_prior:
        b1, b2 = _prior_bias(sd, v, a, 0.0, False, None if relu_masks is None else relu_masks.get("prior"))
    v2 = _mha_q(sd, "v2a_attn.", v, a, num_heads, b1, attn_tf32)
    if taps is not None:
        taps["a0"], taps["s1_branch"] = a.detach(), v2.detach()
    v = F.layer_norm(v + v2, (v.shape[-1],), sd["v_norm.weight"], sd["v_norm.bias"])
    a2 = _mha_q(sd, "a2v_attn.", a, v, num_heads, b2, attn_tf32)
    a = F.layer_norm(a + a2, (a.shape[-1],), sd["a_norm.weight"], sd["a_norm.bias"])
    v_emb, a_emb = v.mean(dim=1), a.mean(dim=1)
    fused = torch.cat([v_emb, a_emb], dim=1)
    if taps is not None:
        taps["v1"], taps["a1"], taps["pooled"] = v.detach(), a.detach(), fused.detach()
    hmask = None if relu_masks is None else relu_masks.get("head")
    act = (lambda x: F.relu(x)) if hmask is None else (lambda x: x * hmask.to(x.dtype))
    if head == "concat":
        hid = act(_QLinear.apply(fused, sd["xattn_mlp.0.weight"], sd["xattn_mlp.0.bias"], False))
        if taps is not None:
            taps["hhid"] = hid.detach()
        return _QLinear.apply(hid, sd["xattn_mlp.3.weight"], sd["xattn_mlp.3.bias"], False)
    if head == "gated":
        hid = act(_QLinear.apply(fused, sd["xattn_gate.0.weight"], sd["xattn_gate.0.bias"], False))
        g = torch.sigmoid(_QLinear.apply(hid, sd["xattn_gate.3.weight"], sd["xattn_gate.3.bias"], False))
        return _QLinear.apply(g * v_emb + (1 - g) * a_emb, sd["xattn_classifier.weight"], sd["xattn_classifier.bias"], False)
    raise ValueError(f"unknown xattn head: {head}")
