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


def _prior_bias(sd: Dict[str, Tensor], v: Tensor, a: Tensor, p_drop: float, training: bool):
    """reference fusion.py:170-184 — including the expand+cat materialisation the reference does."""
    pre = "emotion_prior_bias."
    pooled = torch.cat([v.mean(dim=1), a.mean(dim=1)], dim=-1)
    hid = F.relu(F.linear(pooled, sd[pre + "prior_net.0.weight"], sd[pre + "prior_net.0.bias"]))
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
