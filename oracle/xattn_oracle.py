"""CPU oracle (numpy, float64) for the xattn fusion hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and there only as the checker.  The product path
(``multimodalemotionrecognition_b200``) never imports this module and raises when its CUDA
extension is missing.

This is an independent restatement of the arithmetic of the reference's
``FusionModel.forward`` for ``mode="xattn"`` written from the formulas, in float64:

* projections ............ reference ``src/models/fusion.py:372`` (``v_in_proj``), ``:378``
  (``audio_seq_proj``), ``:386`` (``a_in_proj``)
* emotion prior bias ..... ``src/models/fusion.py:170-184`` (``EmotionPriorBiasAdapter``)
* head-expansion of bias . ``src/models/fusion.py:351-354`` (bias is identical for every head)
* multi-head attention ... ``torch.nn.MultiheadAttention`` (third-party dependency: ``torch``, pinned
  2.10.0 in the reference's ``uv.lock:2659``; 2.11.0 in this image).  Published algorithm, the
  ``need_weights`` branch of ``torch/nn/functional.py`` ``multi_head_attention_forward``:
  ``q,k,v = in_proj``; ``q *= 1/sqrt(dh)``; ``S = q k^T + attn_mask`` (the additive mask is NOT scaled);
  ``P = softmax(S)``; ``O = P v``; heads concatenated; ``out_proj``.  Call sites:
  ``src/models/fusion.py:394`` and ``:398``.
* residual + LayerNorm ... ``src/models/fusion.py:395``, ``:399`` (eps 1e-5, affine; drop-path identity at eval)
* temporal mean pool ..... ``src/models/temporal.py:108-109``
* heads .................. concat ``src/models/fusion.py:405-407``; gated ``:408-411``
* flat concat/gated heads  ``src/models/fusion.py:413-435`` with ``ClipStyleAlignment`` ``:127-150``
* cross entropy .......... ``torch.nn.CrossEntropyLoss(label_smoothing)`` used at ``src/train.py:1033``
* Adam (coupled L2) ...... ``torch.optim.Adam(weight_decay=...)`` used at ``src/train.py:872``

Parity pinning: the reference's own tests hold NO golden values for this path (they assert
shapes only, ``tests/test_attention_integration.py:80-125``).  The oracle is therefore pinned against
outputs of the reference module itself, generated in the build container by
``oracle/gen_golden.py`` (which imports ``/root/reference/src/models/fusion.py``) and committed under
``tests/golden/``.  ``tests/test_oracle_golden.py`` checks this file against those vectors.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np

Array = np.ndarray


def _f64(x) -> Array:
    return np.asarray(x, dtype=np.float64)


def linear(x: Array, w: Array, b: Optional[Array] = None) -> Array:
    """y = x @ w.T + b  (torch.nn.Linear convention: w is [out, in])."""
    y = x @ _f64(w).T
    if b is not None:
        y = y + _f64(b)
    return y


def softmax(x: Array, axis: int = -1) -> Array:
    m = x.max(axis=axis, keepdims=True)
    e = np.exp(x - m)
    return e / e.sum(axis=axis, keepdims=True)


def layer_norm(x: Array, w: Array, b: Array, eps: float = 1e-5) -> Array:
    mu = x.mean(axis=-1, keepdims=True)
    var = ((x - mu) ** 2).mean(axis=-1, keepdims=True)  # biased, as torch LayerNorm
    return (x - mu) / np.sqrt(var + eps) * _f64(w) + _f64(b)


def emotion_prior_bias(sd: Dict[str, Array], v: Array, a: Array, prefix: str = "emotion_prior_bias."):
    """reference fusion.py:178-184.  Returns (prior [B,P], v2a_bias [B,T,Ta], a2v_bias [B,Ta,T])."""
    g = lambda k: _f64(sd[prefix + k])
    pooled = np.concatenate([v.mean(axis=1), a.mean(axis=1)], axis=-1)           # :179-181
    hid = np.maximum(linear(pooled, g("prior_net.0.weight"), g("prior_net.0.bias")), 0.0)
    prior = linear(hid, g("prior_net.3.weight"), g("prior_net.3.bias"))          # [B,P]

    def tok_score(tok: Array, name: str) -> Array:                              # :171-174
        w = g(name + ".weight")[0]
        c = g(name + ".bias")[0]
        e = tok.shape[-1]
        return tok @ w[:e] + (prior @ w[e:])[:, None] + c                        # [B,L]

    scale = float(g("bias_scale"))
    v2a = np.tanh(tok_score(v, "v_query_bias")[:, :, None] + tok_score(a, "a_key_bias")[:, None, :]) * scale
    a2v = np.tanh(tok_score(a, "a_query_bias")[:, :, None] + tok_score(v, "v_key_bias")[:, None, :]) * scale
    return prior, v2a, a2v


def multihead_attention(sd: Dict[str, Array], prefix: str, q_in: Array, kv_in: Array, num_heads: int,
                        bias: Optional[Array] = None, return_probs: bool = False):
    """torch.nn.MultiheadAttention(batch_first=True), eval mode, need_weights branch.

    q_in [B,L,E], kv_in [B,S,E], bias [B,L,S] (same for every head, fusion.py:351-354).
    """
    w = _f64(sd[prefix + "in_proj_weight"])
    b = _f64(sd[prefix + "in_proj_bias"])
    e = q_in.shape[-1]
    dh = e // num_heads
    bsz, L, _ = q_in.shape
    S = kv_in.shape[1]
    q = linear(q_in, w[:e], b[:e])
    k = linear(kv_in, w[e:2 * e], b[e:2 * e])
    v = linear(kv_in, w[2 * e:], b[2 * e:])
    q = q.reshape(bsz, L, num_heads, dh).transpose(0, 2, 1, 3) * (1.0 / math.sqrt(dh))
    k = k.reshape(bsz, S, num_heads, dh).transpose(0, 2, 1, 3)
    v = v.reshape(bsz, S, num_heads, dh).transpose(0, 2, 1, 3)
    s = q @ k.transpose(0, 1, 3, 2)                                             # [B,H,L,S]
    if bias is not None:
        s = s + bias[:, None, :, :]
    p = softmax(s, axis=-1)
    o = (p @ v).transpose(0, 2, 1, 3).reshape(bsz, L, e)
    out = linear(o, sd[prefix + "out_proj.weight"], sd[prefix + "out_proj.bias"])
    if return_probs:
        return out, p
    return out


def temporal_attention_pool(sd: Dict[str, Array], prefix: str, x: Array) -> Array:
    """TemporalAttentionPooling, reference temporal.py:9-26 (eval mode): ``score`` = LayerNorm -> Linear -> GELU (exact, erf) ->
    Dropout -> Linear; attn = softmax over time; out = sum_t attn_t * x_t (the un-normalised tokens).  x [B,L,D] -> [B,D]."""
    y = layer_norm(x, sd[prefix + "score.0.weight"], sd[prefix + "score.0.bias"])           # temporal.py:16
    h = linear(y, sd[prefix + "score.1.weight"], sd[prefix + "score.1.bias"])               # :17
    erf = np.vectorize(math.erf, otypes=[np.float64])
    g = 0.5 * h * (1.0 + erf(h / math.sqrt(2.0)))                                           # :18 nn.GELU()
    logits = linear(g, sd[prefix + "score.4.weight"], sd[prefix + "score.4.bias"])[..., 0]  # :20, :24
    attn = softmax(logits, axis=1)[..., None]                                               # :25
    return (x * attn).sum(axis=1)                                                           # :26


def sinusoidal_positions(length: int, dim: int) -> Array:
    """temporal.py:31-39 (fp32 buffer in the reference; float64 here)."""
    pos = np.arange(length, dtype=np.float64)[:, None]
    div = np.exp(np.arange(0, dim, 2, dtype=np.float64) * (-np.log(10000.0) / max(1, dim)))
    pe = np.zeros((length, dim))
    pe[:, 0::2] = np.sin(pos * div)
    if dim > 1:
        pe[:, 1::2] = np.cos(pos * div[: pe[:, 1::2].shape[1]])
    return pe


def temporal_transformer_pool(sd: Dict[str, Array], prefix: str, x: Array, num_heads: int) -> Array:
    """TemporalTransformerPooling, temporal.py:46-75, eval mode: pre-norm GELU nn.TransformerEncoderLayers (as many as ``sd``
    holds; norm_first: x = x + SA(LN1(x)); x = x + W2·GELU(W1·LN2(x))), then the attention pooling (``prefix + "pool."``)."""
    x = _f64(x) + sinusoidal_positions(x.shape[1], x.shape[2])[None]                          # :72
    erf = np.vectorize(math.erf, otypes=[np.float64])
    layer = 0
    while (prefix + f"encoder.layers.{layer}.norm1.weight") in sd:                            # :69 nn.TransformerEncoder(num_layers)
        lp = prefix + f"encoder.layers.{layer}."
        h = layer_norm(x, sd[lp + "norm1.weight"], sd[lp + "norm1.bias"])
        x = x + multihead_attention(sd, lp + "self_attn.", h, h, num_heads, None)
        h = layer_norm(x, sd[lp + "norm2.weight"], sd[lp + "norm2.bias"])
        f = linear(h, sd[lp + "linear1.weight"], sd[lp + "linear1.bias"])
        f = 0.5 * f * (1.0 + erf(f / math.sqrt(2.0)))                                         # activation="gelu" (exact)
        x = x + linear(f, sd[lp + "linear2.weight"], sd[lp + "linear2.bias"])                # :73
        layer += 1
    return temporal_attention_pool(sd, prefix + "pool.", x)                                  # :74


def audio_time_conv(sd: Dict[str, Array], mel: Array) -> Array:
    """Mel front end, reference fusion.py:379-384: Conv1d(n_mels -> a_dim, kernel_size=3, padding=1) over time, then
    permute to [B, Ta, a_dim].  ``mel`` [B,1,n_mels,Ta] or [B,n_mels,Ta].  Restated as the convolution sum itself
    (cross-correlation, zero padding), NOT as the unfold + Linear the CUDA path uses."""
    x = _f64(mel)
    if x.ndim == 4:
        x = x[:, 0]                                                                          # :382 squeeze(1)
    w = _f64(sd["audio_time_conv.weight"])                                                   # [a_dim, n_mels, 3]
    bsz, _, ta = x.shape
    xp = np.pad(x, ((0, 0), (0, 0), (1, 1)))                                                 # padding=1
    out = np.zeros((bsz, w.shape[0], ta))
    for k in range(3):                                                                       # out[b,o,t] += w[o,c,k] x[b,c,t+k-1]
        out += np.einsum("oc,bct->bot", w[:, :, k], xp[:, :, k:k + ta])
    out += _f64(sd["audio_time_conv.bias"])[None, :, None]                                   # :383
    return out.transpose(0, 2, 1)                                                            # :384


def xattn_forward(sd: Dict[str, Array], video_emb: Array, audio_emb: Array, *, num_heads: int = 4,
                  head: str = "concat", use_prior: Optional[bool] = None, taps: Optional[dict] = None,
                  audio_frontend: str = "seq", temporal_num_heads: int = 4) -> Array:
    """Forward of the xattn fusion head, reference fusion.py:372-411, eval mode.

    ``sd`` maps reference state_dict keys (head only) to arrays.  ``video_emb`` [B,T,Dv],
    ``audio_emb`` [B,Ta,Da].  Returns logits [B,C] float64.  ``taps`` (optional dict) receives
    intermediate tensors for stage-by-stage checks.
    """
    xv = _f64(video_emb)
    xa = _f64(audio_emb)
    if audio_frontend == "mel":
        a_seq = audio_time_conv(sd, xa)                                           # :379-384
    elif xv.ndim != 3 or xa.ndim != 3:
        raise ValueError("expected video_emb [B,T,Dv] and audio_emb [B,Ta,Da]")
    if use_prior is None:
        use_prior = any(k.startswith("emotion_prior_bias.") for k in sd)
    v = linear(xv, sd["v_in_proj.weight"], sd["v_in_proj.bias"])                 # :372
    if audio_frontend != "mel":
        a_seq = linear(xa, sd["audio_seq_proj.weight"], sd["audio_seq_proj.bias"])   # :378
    a = linear(a_seq, sd["a_in_proj.weight"], sd["a_in_proj.bias"])              # :386
    v2a_bias = a2v_bias = None
    prior = None
    if use_prior:
        prior, v2a_bias, a2v_bias = emotion_prior_bias(sd, v, a)                 # :390-391 (pre-update v, a)
    v2 = multihead_attention(sd, "v2a_attn.", v, a, num_heads, v2a_bias)         # :394
    v_new = layer_norm(v + v2, sd["v_norm.weight"], sd["v_norm.bias"])           # :395
    a2 = multihead_attention(sd, "a2v_attn.", a, v_new, num_heads, a2v_bias)     # :398 (uses UPDATED v)
    a_new = layer_norm(a + a2, sd["a_norm.weight"], sd["a_norm.bias"])           # :399
    if "v_temporal_pool.pool.encoder.layers.0.norm1.weight" in sd:              # temporal_pooling="transformer", temporal.py:95-102
        v_emb = temporal_transformer_pool(sd, "v_temporal_pool.pool.", v_new, temporal_num_heads)
        a_emb = temporal_transformer_pool(sd, "a_temporal_pool.pool.", a_new, temporal_num_heads)
    elif "v_temporal_pool.pool.score.0.weight" in sd:                           # temporal_pooling="attn", temporal.py:93-94
        v_emb = temporal_attention_pool(sd, "v_temporal_pool.pool.", v_new)
        a_emb = temporal_attention_pool(sd, "a_temporal_pool.pool.", a_new)
    else:
        v_emb = v_new.mean(axis=1)                                              # temporal.py:108-109
        a_emb = a_new.mean(axis=1)
    fused = np.concatenate([v_emb, a_emb], axis=-1)
    if head == "concat":                                                        # :405-407
        hid = np.maximum(linear(fused, sd["xattn_mlp.0.weight"], sd["xattn_mlp.0.bias"]), 0.0)
        logits = linear(hid, sd["xattn_mlp.3.weight"], sd["xattn_mlp.3.bias"])
    elif head == "gated":                                                       # :408-411
        hid = np.maximum(linear(fused, sd["xattn_gate.0.weight"], sd["xattn_gate.0.bias"]), 0.0)
        z = linear(hid, sd["xattn_gate.3.weight"], sd["xattn_gate.3.bias"])
        gate = 1.0 / (1.0 + np.exp(-z))
        mix = gate * v_emb + (1.0 - gate) * a_emb
        logits = linear(mix, sd["xattn_classifier.weight"], sd["xattn_classifier.bias"])
    else:
        raise ValueError(f"unknown xattn head: {head}")
    if taps is not None:
        taps.update(v=v, a_seq=a_seq, a=a, prior=prior, v2a_bias=v2a_bias, a2v_bias=a2v_bias, v2=v2,
                    v_new=v_new, a2=a2, a_new=a_new, v_emb=v_emb, a_emb=a_emb)
    return logits


def flat_head_forward(sd: Dict[str, Array], audio_pooled: Array, video_pooled: Array, *, mode: str = "concat",
                      align: bool = True, return_align_loss: bool = False):
    """Flat concat/gated heads (BASELINE config 5), reference fusion.py:413-435, eval mode.

    ``audio_pooled`` [B,Da], ``video_pooled`` [B,Dv] are the temporally mean-pooled embeddings
    (what ``audio_model.encode`` / ``video_model.encode`` return).  With ``align`` the
    ``ClipStyleAlignment`` projections (fusion.py:138-139) come first; its loss (:141-149) is
    returned on request.
    """
    a = _f64(audio_pooled)
    v = _f64(video_pooled)
    loss = None
    if align:
        a = linear(a, sd["semantic_alignment.audio_proj.weight"], sd["semantic_alignment.audio_proj.bias"])
        v = linear(v, sd["semantic_alignment.video_proj.weight"], sd["semantic_alignment.video_proj.bias"])
        if return_align_loss:
            an = a / np.maximum(np.linalg.norm(a, axis=-1, keepdims=True), 1e-12)
            vn = v / np.maximum(np.linalg.norm(v, axis=-1, keepdims=True), 1e-12)
            scale = min(math.exp(float(_f64(sd["semantic_alignment.logit_scale"]))), 100.0)
            lg = scale * (an @ vn.T)
            tgt = np.arange(lg.shape[0])
            loss = 0.5 * (cross_entropy(lg, tgt)[0] + cross_entropy(lg.T, tgt)[0])
    a = linear(a, sd["audio_proj.weight"], sd["audio_proj.bias"])                # :421
    v = linear(v, sd["video_proj.weight"], sd["video_proj.bias"])                # :422
    cat = np.concatenate([a, v], axis=-1)                                        # audio first (:425, :432)
    if mode == "concat":
        hid = np.maximum(linear(cat, sd["fusion.0.weight"], sd["fusion.0.bias"]), 0.0)
        logits = linear(hid, sd["fusion.3.weight"], sd["fusion.3.bias"])
    elif mode == "gated":
        hid = np.maximum(linear(cat, sd["gate.0.weight"], sd["gate.0.bias"]), 0.0)
        g = 1.0 / (1.0 + np.exp(-linear(hid, sd["gate.3.weight"], sd["gate.3.bias"])))
        logits = linear(g * a + (1.0 - g) * v, sd["classifier.weight"], sd["classifier.bias"])   # :434
    else:
        raise ValueError(f"unknown flat head mode: {mode}")
    if return_align_loss:
        return logits, loss
    return logits


def cross_entropy(logits: Array, labels: Array, label_smoothing: float = 0.0):
    """Mean-reduced CE with label smoothing as torch.nn.CrossEntropyLoss.  Returns (loss, dlogits)."""
    z = _f64(logits)
    n, c = z.shape
    logp = z - z.max(axis=1, keepdims=True)
    logp = logp - np.log(np.exp(logp).sum(axis=1, keepdims=True))
    tgt = np.full((n, c), label_smoothing / c)
    tgt[np.arange(n), np.asarray(labels)] += 1.0 - label_smoothing
    loss = float(-(tgt * logp).sum(axis=1).mean())
    dlogits = (np.exp(logp) - tgt) / n
    return loss, dlogits


def adam_step(p: Array, g: Array, m: Array, v: Array, *, step: int, lr: float, beta1: float = 0.9,
              beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.0, decoupled: bool = False):
    """One Adam step.  decoupled=False: torch.optim.Adam (L2 added to the gradient, train.py:872);
    decoupled=True: torch.optim.AdamW.  Returns (p, m, v) updated, float64."""
    p, g, m, v = _f64(p).copy(), _f64(g).copy(), _f64(m).copy(), _f64(v).copy()
    if decoupled:
        p *= 1.0 - lr * weight_decay
    else:
        g = g + weight_decay * p
    m = beta1 * m + (1.0 - beta1) * g
    v = beta2 * v + (1.0 - beta2) * g * g
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = np.sqrt(v) / math.sqrt(bc2) + eps
    p = p - (lr / bc1) * m / denom
    return p, m, v
