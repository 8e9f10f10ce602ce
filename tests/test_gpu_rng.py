"""GPU tests (``-m gpu``) of the training-mode randomness (SURVEY 7.2 "Training-mode RNG").

Bitwise parity with torch's generator is impossible; what the reference's five dropout sites DO is testable:
``StochasticDepth`` draws ONE Bernoulli(keep) per sample, shape [B,1,1], and divides by keep (fusion.py:18-26); attention
dropout acts on the softmax probabilities (fusion.py:277/:280 -> F.multi_head_attention_forward); ``Dropout(0.2)`` sits
behind the head's ReLU (:315 / :322); ``Dropout`` behind prior_net's ReLU (:161); the attention pooler drops units of its
score MLP (temporal.py:19).  Every site here is a Philox mask ``drop_factor(p, seed, offset, stream, index)``; the test
library exposes that function (``xattn_debug_dropout_factors``), so each site is checked EXACTLY — the training-mode
activations in the kernel workspace equal the eval-mode ones with the mask applied where the reference applies it — and
the mask itself statistically: values in {0, 1/keep}, keep-rate within 4 sigma, independent across streams and steps."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

from helpers import make_head, ws_f32_views

pytestmark = pytest.mark.gpu
RS_ATTN1, RS_ATTN2, RS_PATH1, RS_PATH2, RS_HEAD, RS_PRIOR, RS_POOLV, RS_POOLA = range(1, 9)


def _factors(dev, p, seed, offset, stream, n, idx0=0):
    from multimodalemotionrecognition_b200 import _lib

    out = torch.empty(n, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        _lib.debug_check(_lib.load_debug().xattn_debug_dropout_factors(C.c_float(p), seed, offset, stream, idx0, n,
                                                                        C.c_void_p(out.data_ptr()),
                                                                        C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    torch.cuda.synchronize(dev)
    return out


def _check_mask_statistics(f, p, what):
    keep = 1.0 - p
    kept = f > 0
    assert torch.all((f == 0) | ((f - 1.0 / keep).abs() <= 1e-6 / keep)), f"{what}: factors must be 0 or 1/keep"
    n = f.numel()
    rate = kept.float().mean().item()
    sigma = math.sqrt(p * keep / n)
    assert abs(rate - keep) <= 4 * sigma + 1e-12, (what, rate, keep, sigma)
    return rate


@pytest.mark.parametrize("p", [0.1, 0.2, 0.5])
def test_dropout_mask_keep_rate_scaling_and_independence(cuda_device, p):
    n = 1 << 20
    seed, off = 1234567, 5
    masks = [_factors(cuda_device, p, seed, off, s, n) for s in range(1, 9)]
    for s, f in enumerate(masks, start=1):
        _check_mask_statistics(f, p, f"stream {s}")
        assert abs(f.mean().item() - 1.0) <= 4 * math.sqrt(p / (1 - p) / n)      # E[mask / keep] = 1: expectation preserved
    keep = 1 - p
    for a, b in [(0, 1), (2, 3), (4, 5), (0, 7)]:                                   # streams are independent draws
        both = ((masks[a] > 0) & (masks[b] > 0)).float().mean().item()
        assert abs(both - keep * keep) <= 5 * math.sqrt(keep * keep * (1 - keep * keep) / n), (a, b, both)
    nxt = _factors(cuda_device, p, seed, off + 1, 1, n)                            # the next step draws a new mask
    both = ((masks[0] > 0) & (nxt > 0)).float().mean().item()
    assert abs(both - keep * keep) <= 5 * math.sqrt(keep * keep * (1 - keep * keep) / n)
    again = _factors(cuda_device, p, seed, off, 1, n)                              # and the same step the same one (backward)
    assert torch.equal(again, masks[0])
    part = _factors(cuda_device, p, seed, off, 1, 1000, idx0=4096 + 3)             # index-addressable: any sub-range agrees
    assert torch.equal(part, masks[0][4099:5099])


def _run(m, xv, xa, training):
    with torch.no_grad():
        logits, (cfg, _, _, ws) = m._launch_forward(xv, xa, save=True, training=training)
    torch.cuda.synchronize()
    return logits, cfg, ws_f32_views(ws, cfg)


def _inputs(dev, B, seed=5):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, 16, 512, generator=g).to(dev), torch.randn(B, 149, 768, generator=g).to(dev)


def _head(dev, precision="fp32", **kw):
    kw.setdefault("xattn_attn_dropout", 0.0)
    kw.setdefault("xattn_stochastic_depth", 0.0)
    kw.setdefault("xattn_emotion_prior_dropout", 0.0)
    m = make_head("concat", True, seed=0, precision=precision, **kw).to(dev)
    m.xattn_head_dropout = 0.0   # (hard-coded 0.2 in the reference, fusion.py:315; switched off unless a test sets it)
    return m


def test_stochastic_depth_is_one_bernoulli_per_sample_scaled_by_keep(cuda_device):
    """fusion.py:18-26: mask shape [B,1,1], x * mask / keep; identity in eval."""
    B, p = 512, 0.3
    m = _head(cuda_device, xattn_stochastic_depth=p)
    xv, xa = _inputs(cuda_device, B)
    _, cfg_e, we = _run(m, xv, xa, training=False)
    s1_eval, s2_eval, v0, a0 = we["s1"].clone(), we["s2"].clone(), we["v0"].clone(), we["a0"].clone()
    assert torch.all(we["dp1"] == 1) and torch.all(we["dp2"] == 1)
    _, cfg, wt = _run(m, xv, xa, training=True)
    for stream, name in ((RS_PATH1, "dp1"), (RS_PATH2, "dp2")):
        f = _factors(cuda_device, p, cfg.seed, cfg.offset, stream, B)
        assert torch.equal(wt[name], f), name                                      # one draw per SAMPLE (index = clip)
        _check_mask_statistics(wt[name], p, name)
    # video block: s1 = v0 + f_b * branch with the eval branch (the branch's inputs do not depend on the mask)
    f1 = wt["dp1"].view(B, 1, 1)
    assert torch.allclose(wt["s1"], v0 + f1 * (s1_eval - v0), atol=2e-5, rtol=1e-5)
    dropped = (wt["dp1"] == 0)
    assert dropped.any() and torch.equal(wt["s1"][dropped], v0[dropped])           # a dropped sample keeps only the residual
    # audio block: its branch sees the masked video tokens, so only the per-sample structure is checked
    f2 = wt["dp2"]
    d2 = (wt["s2"] - a0).abs().amax(dim=(1, 2))
    assert torch.all(d2[f2 == 0] == 0) and torch.all(d2[f2 > 0] > 0)


def test_head_and_prior_dropout_sit_behind_the_relu(cuda_device):
    """fusion.py:315 (Dropout(0.2) after the head's ReLU) and :161 (prior_net Dropout): hidden = relu(...) * mask / keep."""
    B = 256
    xv, xa = _inputs(cuda_device, B)
    m = _head(cuda_device)
    m.xattn_head_dropout = 0.2
    _, _, we = _run(m, xv, xa, training=False)
    hh_eval = we["hhid"].clone()
    _, cfg, wt = _run(m, xv, xa, training=True)
    f = _factors(cuda_device, 0.2, cfg.seed, cfg.offset, RS_HEAD, hh_eval.numel()).view_as(hh_eval)
    _check_mask_statistics(f, 0.2, "head")
    assert torch.allclose(wt["hhid"], hh_eval * f, atol=1e-6, rtol=1e-6)
    m2 = _head(cuda_device, xattn_emotion_prior_dropout=0.25)
    m2.xattn_head_dropout = 0.0
    _, _, we = _run(m2, xv, xa, training=False)
    ph_eval = we["phid"].clone()
    _, cfg, wt = _run(m2, xv, xa, training=True)
    f = _factors(cuda_device, 0.25, cfg.seed, cfg.offset, RS_PRIOR, ph_eval.numel()).view_as(ph_eval)
    assert torch.allclose(wt["phid"], ph_eval * f, atol=1e-6, rtol=1e-6)
    assert (f == 0).float().mean().item() == pytest.approx(0.25, abs=4 * math.sqrt(0.25 * 0.75 / f.numel()))


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("tf32", 4e-3)])
def test_attention_dropout_acts_on_the_probabilities(cuda_device, precision, tol):
    """F.multi_head_attention_forward: attn = dropout(softmax(scores)); out = attn @ v — per (clip, head, query, key) element.
    fp32: the SIMT attention kernels; tf32: the mma.sync tensor-core kernels (csrc/attn_mma.cu), same Philox indexing."""
    B, p, H, E = 64, 0.1, 4, 128
    m = _head(cuda_device, precision=precision, xattn_attn_dropout=p)
    xv, xa = _inputs(cuda_device, B)
    _, cfg, w = _run(m, xv, xa, training=True)
    for stream, pn, kvn, on in ((RS_ATTN1, "p1", "kv1", "o1"), (RS_ATTN2, "p2", "kv2", "o2")):
        S = w[kvn].shape[1]          # keys; under tf32 the rows of P (and with them the Philox index space) are padded to 4
        f = _factors(cuda_device, p, cfg.seed, cfg.offset, stream, w[pn].numel()).view_as(w[pn]).double()[..., :S]
        Pw = w[pn][..., :S]
        if precision == "tf32":   # the tensor-core forward saves the dropout decision in the sign of P (csrc/attn_mma.cu)
            assert torch.equal(torch.signbit(Pw), f == 0), pn
        P = Pw.double().abs()                                                       # [B,H,L,S], pre-dropout
        assert torch.allclose(P.sum(-1), torch.ones_like(P.sum(-1)), atol=1e-5)
        _check_mask_statistics(f.float(), p, pn)
        V = w[kvn][..., E:].double().view(B, -1, H, E // H).transpose(1, 2)         # [B,H,S,dh]
        O = ((P * f) @ V).transpose(1, 2).reshape(B, -1, E)
        assert (O - w[on].double()).abs().max().item() <= tol, on


def test_attention_pooler_dropout_drops_score_mlp_units(cuda_device):
    """temporal.py:15-25: score = Linear(Dropout(GELU(Linear(LN(x))))); weights = softmax over time; pooled = sum w_t x_t."""
    B, p, E = 32, 0.3, 128
    m = make_head("concat", True, seed=0, precision="fp32", temporal_pooling="attn", temporal_dropout=p, xattn_attn_dropout=0.0,
                  xattn_stochastic_depth=0.0, xattn_emotion_prior_dropout=0.0).to(cuda_device)
    m.xattn_head_dropout = 0.0
    xv, xa = _inputs(cuda_device, B)
    _, cfg, w = _run(m, xv, xa, training=True)
    sd = {k: v.detach().double() for k, v in m.state_dict().items()}
    for stream, side, tok, out in ((RS_POOLV, "v", "v1", "poolv"), (RS_POOLA, "a", "a1", "poola")):
        x = w[tok].double()
        L = x.shape[1]
        pre = f"{side}_temporal_pool.pool.score."
        h = torch.nn.functional.layer_norm(x, (E,), sd[pre + "0.weight"], sd[pre + "0.bias"])
        h = torch.nn.functional.gelu(h @ sd[pre + "1.weight"].T + sd[pre + "1.bias"])
        f = _factors(cuda_device, p, cfg.seed, cfg.offset, stream, h.numel()).view_as(h).double()
        _check_mask_statistics(f.float(), p, out)
        sc = (h * f) @ sd[pre + "4.weight"].T + sd[pre + "4.bias"]
        wts = torch.softmax(sc.squeeze(-1), dim=1)
        pooled = (wts.unsqueeze(-1) * x).sum(1)
        assert (pooled - w[out].double()).abs().max().item() <= 2e-5, out
