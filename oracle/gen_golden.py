"""Generates tests/golden/*.npz by running the UNMODIFIED reference module in the build container.

TEST INFRASTRUCTURE ONLY.  Imports ``/root/reference/src/models/fusion.py`` (pure torch; nothing is copied into the
repo), runs it on seeded synthetic inputs and stores inputs / weights (small cases) or SHA-256 digests of them
(BASELINE config 1, whose weights come from ``torch.manual_seed(0)`` in reference constructor order and inputs from
``torch.Generator().manual_seed(1)``) together with the reference's logits, loss, gradients and
post-Adam parameters.  ``/root/reference`` does not exist on the GPU box, hence the committed fixtures.

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py
"""
from __future__ import annotations

import hashlib
import os
import sys
from pathlib import Path

import numpy as np
import torch
from torch import nn

REF = os.environ.get("XATTN_REFERENCE_ROOT", "/root/reference")
OUT = Path(__file__).resolve().parents[1] / "tests" / "golden"


def sha(t) -> str:
    a = t.detach().cpu().contiguous().numpy() if isinstance(t, torch.Tensor) else np.ascontiguousarray(t)
    return hashlib.sha256(a.tobytes()).hexdigest()


class StubVideo(nn.Module):
    """Backbone stand-in: the 'frames' already are the embeddings ([B*T, Dv, 1, 1] -> same)."""

    def __init__(self, dim: int) -> None:
        super().__init__()
        self.embedding_dim = dim
        self.backbone = nn.Identity()

    def encode(self, x):
        return x


class StubAudio(nn.Module):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.sequence_dim = dim
        self.embedding_dim = dim

    def encode_sequence(self, x):
        return x

    def encode(self, x):
        return x


def build_ref(FusionModel, *, dv, da, d_model, heads, head, prior, common=256, prior_dim=8, prior_hidden=64, classes=8,
              mode="xattn", align="none", pooling="mean"):
    return FusionModel(StubAudio(da), StubVideo(dv), classes, mode=mode, common_dim=common, xattn_head=head, d_model=d_model,
                       temporal_pooling=pooling,
                       num_heads=heads, audio_n_mels=da, xattn_use_emotion_prior=prior, xattn_emotion_prior_dim=prior_dim,
                       xattn_emotion_prior_hidden_dim=prior_hidden, fusion_align_mode=align, fusion_align_dim=256)


def fusion_state(model) -> dict:
    return {k: v.detach().clone() for k, v in model.state_dict().items()
            if not (k.startswith("audio_model.") or k.startswith("video_model."))}


def grads_of(model) -> dict:
    return {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}


def digest(t: torch.Tensor) -> np.ndarray:
    t = t.detach().double().flatten()
    idx = torch.linspace(0, t.numel() - 1, steps=8).long()   # 8 evenly spaced samples (repeats for tiny tensors)
    return np.concatenate([[t.sum().item(), t.abs().sum().item(), t.norm().item()], t[idx].numpy()])


def run_case(model, xv, xa, labels, lr=1e-3, wd=1e-4):
    """eval-mode forward, CE, backward, one torch.optim.Adam step (train.py:204-228 with dropouts identity)."""
    model.eval()
    with torch.no_grad():
        logits = model(xv[..., None, None], xa)
    params = [p for n, p in model.named_parameters()]
    opt = torch.optim.Adam(params, lr=lr, weight_decay=wd)
    opt.zero_grad()
    out = model(xv[..., None, None], xa)
    loss = nn.functional.cross_entropy(out, labels)
    loss.backward()
    grads = grads_of(model)
    opt.step()
    after = fusion_state(model)
    return logits, loss.detach(), grads, after


def gen_attnpool(FusionModel) -> None:
    """temporal_pooling="attn" (temporal.py:9-26), eval forward: a tiny case stored in full and the
    BASELINE config-1 dimensions with seed-0 weights (digest of the weights, full logits)."""
    blob = {}
    torch.manual_seed(12)
    m = build_ref(FusionModel, dv=16, da=8, d_model=8, heads=2, head="concat", prior=True, common=16, prior_dim=4, prior_hidden=8,
                  pooling="attn").eval()
    g = torch.Generator().manual_seed(13)
    xv = torch.randn(3, 4, 16, generator=g)
    xa = torch.randn(3, 12, 8, generator=g)
    with torch.no_grad():
        logits = m(xv[..., None, None], xa)
    blob.update({"tiny:xv": xv.numpy(), "tiny:xa": xa.numpy(), "tiny:logits": logits.numpy()})
    blob.update({"tiny:w:" + k: v.numpy() for k, v in fusion_state(m).items()})
    labels = torch.randint(0, 8, (3,), generator=g)
    loss = nn.functional.cross_entropy(m(xv[..., None, None], xa), labels)   # eval mode: dropouts identity, autograd on
    loss.backward()
    blob.update({"tiny:labels": labels.numpy(), "tiny:loss": loss.detach().numpy()})
    blob.update({"tiny:g:" + k: v.numpy() for k, v in grads_of(m).items()})
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    for head in ("concat", "gated"):
        torch.manual_seed(0)
        m = build_ref(FusionModel, dv=512, da=768, d_model=128, heads=4, head=head, prior=True, pooling="attn").eval()
        with torch.no_grad():
            logits = m(xv[..., None, None], xa)
        sd = fusion_state(m)
        labels = torch.randint(0, 8, (8,), generator=torch.Generator().manual_seed(2))
        loss = nn.functional.cross_entropy(m(xv[..., None, None], xa), labels)
        loss.backward()
        blob[f"{head}:labels"] = labels.numpy()
        blob[f"{head}:loss"] = loss.detach().numpy()
        for k, v in grads_of(m).items():
            blob[f"{head}:gd:{k}"] = digest(v)
            if "temporal_pool" in k:
                blob[f"{head}:g:{k}"] = v.numpy()
        blob[f"{head}:logits"] = logits.numpy()
        blob[f"{head}:keys"] = np.array(sorted(sd))
        blob[f"{head}:pool_digest"] = digest(torch.cat([v.flatten() for k, v in sorted(sd.items()) if "temporal_pool" in k]))
    np.savez_compressed(OUT / "xattn_attnpool.npz", **blob)


def gen_txpool(FusionModel) -> None:
    """temporal_pooling="transformer" (temporal.py:46-75), eval mode: a tiny case stored in full (2 pooler heads,
    one layer) with its CE gradients and post-Adam parameters, the same with TWO encoder layers (``tiny2:``), and the BASELINE
    config-1 dimensions with seed-0 weights (keys, digest of the pooler weights, full logits, loss, gradient digests and the pooler
    gradients in full)."""
    blob = {}
    for tag, layers, seed in (("tiny", 1, 31), ("tiny2", 2, 33)):
        torch.manual_seed(seed)
        m = FusionModel(StubAudio(8), StubVideo(16), 8, mode="xattn", common_dim=16, xattn_head="concat", d_model=8, num_heads=2,
                        audio_n_mels=8, temporal_pooling="transformer", temporal_num_heads=2, temporal_num_layers=layers,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8).eval()
        g = torch.Generator().manual_seed(seed + 1)
        xv = torch.randn(3, 4, 16, generator=g)
        xa = torch.randn(3, 12, 8, generator=g)
        labels = torch.randint(0, 8, (3,), generator=g)
        before = fusion_state(m)
        logits, loss, grads, after = run_case(m, xv, xa, labels)
        blob.update({f"{tag}:xv": xv.numpy(), f"{tag}:xa": xa.numpy(), f"{tag}:logits": logits.numpy(), f"{tag}:labels": labels.numpy(),
                     f"{tag}:loss": loss.numpy()})
        blob.update({f"{tag}:w:" + k: v.numpy() for k, v in before.items()})
        blob.update({f"{tag}:g:" + k: v.numpy() for k, v in grads.items()})
        blob.update({f"{tag}:after:" + k: v.numpy() for k, v in after.items()})
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    labels = torch.randint(0, 8, (8,), generator=torch.Generator().manual_seed(2))
    for head in ("concat", "gated"):
        torch.manual_seed(0)
        m = build_ref(FusionModel, dv=512, da=768, d_model=128, heads=4, head=head, prior=True, pooling="transformer").eval()
        with torch.no_grad():
            logits = m(xv[..., None, None], xa)
        sd = fusion_state(m)
        blob[f"{head}:logits"] = logits.numpy()
        blob[f"{head}:keys"] = np.array(sorted(sd))
        blob[f"{head}:pool_digest"] = digest(torch.cat([v.flatten() for k, v in sorted(sd.items()) if "temporal_pool" in k]))
        loss = nn.functional.cross_entropy(m(xv[..., None, None], xa), labels)   # eval mode: dropouts identity, autograd on
        loss.backward()
        blob[f"{head}:labels"] = labels.numpy()
        blob[f"{head}:loss"] = loss.detach().numpy()
        for k, v in grads_of(m).items():
            blob[f"{head}:gd:{k}"] = digest(v)
            if "temporal_pool" in k and v.numel() <= 8192:
                blob[f"{head}:g:{k}"] = v.numpy()
    np.savez_compressed(OUT / "xattn_txpool.npz", **blob)


class StubAudioMel(nn.Module):
    """An encoder without encode_sequence(): the reference takes the mel fallback (fusion.py:379-384)."""

    def __init__(self, dim: int) -> None:
        super().__init__()
        self.embedding_dim = dim


def gen_mel(FusionModel) -> None:
    """Mel front end (audio_time_conv, fusion.py:379-384): eval forward, CE, backward and one Adam step.
    A tiny case stored in full; n_mels=64 / Ta=149 / d_model=128 with seed-0 weights stored as logits + gradient digests."""
    blob = {}
    torch.manual_seed(21)
    m = FusionModel(StubAudioMel(8), StubVideo(16), 8, mode="xattn", common_dim=16, xattn_head="concat", d_model=8, num_heads=2,
                    audio_n_mels=5, xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8)
    g = torch.Generator().manual_seed(22)
    xv = torch.randn(3, 4, 16, generator=g)
    mel = torch.randn(3, 1, 5, 37, generator=g)
    labels = torch.randint(0, 8, (3,), generator=g)
    before = fusion_state(m)
    logits, loss, grads, after = run_case(m, xv, mel, labels)
    blob.update({"tiny:xv": xv.numpy(), "tiny:mel": mel.numpy(), "tiny:labels": labels.numpy(), "tiny:logits": logits.numpy(),
                 "tiny:loss": loss.numpy()})
    blob.update({"tiny:w:" + k: v.numpy() for k, v in before.items()})
    blob.update({"tiny:g:" + k: v.numpy() for k, v in grads.items()})
    blob.update({"tiny:after:" + k: v.numpy() for k, v in after.items()})
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    mel = torch.randn(8, 1, 64, 149, generator=g)
    labels = torch.randint(0, 8, (8,), generator=g)
    for head in ("concat", "gated"):
        torch.manual_seed(0)
        m = FusionModel(StubAudioMel(128), StubVideo(512), 8, mode="xattn", xattn_head=head, d_model=128, num_heads=4,
                        audio_n_mels=64, xattn_use_emotion_prior=True)
        logits, loss, grads, after = run_case(m, xv, mel, labels)
        blob[f"{head}:logits"] = logits.numpy()
        blob[f"{head}:loss"] = loss.numpy()
        blob[f"{head}:keys"] = np.array(sorted(fusion_state(m)))
        blob[f"{head}:g:audio_time_conv.weight"] = grads["audio_time_conv.weight"].numpy()
        blob[f"{head}:g:audio_time_conv.bias"] = grads["audio_time_conv.bias"].numpy()
        for k, v in grads.items():
            blob[f"{head}:gd:{k}"] = digest(v)
        assert "audio_seq_proj.weight" not in grads
    np.savez_compressed(OUT / "xattn_mel.npz", **blob)


def gen_flat_train(FusionModel) -> None:
    """Trainable flat concat / gated heads (fusion.py:226-260, :413-435, :29-55, :127-150; train.py:218-227): logits, CE loss,
    alignment loss and the gradients of ``CE + 0.1 * alignment`` w.r.t. every parameter and both (sequence) inputs — eval mode, and
    training mode with the hidden Dropout's p set to 0 on the instance (torch's mask cannot be reproduced) for torch seeds whose two
    ``torch.rand(1)`` draws make ModalityDropout drop the audio, the video, both and neither.  Tiny dimensions stored in full;
    BASELINE config-5 dimensions by digest (+ the small tensors in full)."""
    blob = {}

    def run(m, xv, xa, labels, tag):
        m.zero_grad(set_to_none=True)
        xv = xv.clone().requires_grad_(True)
        xa = xa.clone().requires_grad_(True)
        logits = m(xv.mean(dim=1), xa.mean(dim=1))
        cls = nn.functional.cross_entropy(logits, labels)
        align = m.pop_alignment_loss()
        total = cls + 0.1 * align
        total.backward()
        blob[f"{tag}:logits"] = logits.detach().numpy()
        blob[f"{tag}:cls"] = cls.detach().numpy()
        blob[f"{tag}:align"] = align.detach().numpy()
        return {k: p.grad.detach().clone() for k, p in m.named_parameters() if p.grad is not None}, xv.grad.clone(), xa.grad.clone()

    for mode in ("concat", "gated"):
        torch.manual_seed(40)
        m = build_ref(FusionModel, dv=12, da=10, d_model=8, heads=2, head="concat", prior=False, mode=mode, align="clip", common=16)
        g = torch.Generator().manual_seed(41)
        xa = torch.randn(5, 5, 10, generator=g)
        xv = torch.randn(5, 4, 12, generator=g)
        labels = torch.randint(0, 8, (5,), generator=g)
        blob.update({f"tiny_{mode}:xa": xa.numpy(), f"tiny_{mode}:xv": xv.numpy(), f"tiny_{mode}:labels": labels.numpy()})
        blob.update({f"tiny_{mode}:w:" + k: v.numpy() for k, v in fusion_state(m).items()})
        m.eval()
        grads, dxv, dxa = run(m, xv, xa, labels, f"tiny_{mode}:eval")
        blob.update({f"tiny_{mode}:eval:g:" + k: v.numpy() for k, v in grads.items()})
        blob[f"tiny_{mode}:eval:dxv"], blob[f"tiny_{mode}:eval:dxa"] = dxv.numpy(), dxa.numpy()
        if mode == "gated":
            m.gate[2].p = 0.0      # instance attribute only: the hidden Dropout's mask is torch RNG, not reproducible elsewhere
            m.train()
            want = {(1, 0): None, (0, 1): None, (1, 1): None, (0, 0): None}
            for seed in range(400):
                torch.manual_seed(seed)
                d = (int(torch.rand(1).item() < 0.2), int(torch.rand(1).item() < 0.2))
                if want[d] is None:
                    want[d] = seed
                if all(v is not None for v in want.values()):
                    break
            for (da_, dv_), seed in want.items():
                tag = f"tiny_gated:train_a{da_}v{dv_}"
                torch.manual_seed(seed)
                grads, dxv, dxa = run(m, xv, xa, labels, tag)
                blob[f"{tag}:seed"] = np.array(seed)
                blob.update({f"{tag}:g:" + k: v.numpy() for k, v in grads.items()})
                blob[f"{tag}:dxv"], blob[f"{tag}:dxa"] = dxv.numpy(), dxa.numpy()
    g = torch.Generator().manual_seed(3)
    ap = torch.randn(6, 1, 768, generator=g)
    vp = torch.randn(6, 1, 512, generator=g)
    labels = torch.randint(0, 8, (6,), generator=torch.Generator().manual_seed(4))
    blob["c5:labels"] = labels.numpy()
    for mode in ("concat", "gated"):
        torch.manual_seed(0)
        m = build_ref(FusionModel, dv=512, da=768, d_model=128, heads=4, head="concat", prior=False, mode=mode, align="clip").eval()
        grads, dxv, dxa = run(m, vp, ap, labels, f"c5_{mode}:eval")
        for k, v in grads.items():
            blob[f"c5_{mode}:eval:gd:{k}"] = digest(v)
            if v.numel() <= 2048:
                blob[f"c5_{mode}:eval:g:{k}"] = v.numpy()
        blob[f"c5_{mode}:eval:dxv"], blob[f"c5_{mode}:eval:dxa"] = digest(dxv), digest(dxa)
    np.savez_compressed(OUT / "flat_train.npz", **blob)


def main() -> None:
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    from src.models.fusion import FusionModel  # the reference, unmodified

    OUT.mkdir(parents=True, exist_ok=True)
    torch.set_num_threads(1)
    if "--only-flat-train" in sys.argv:
        gen_flat_train(FusionModel)
        return
    if "--only-txpool" in sys.argv:
        gen_txpool(FusionModel)
        return
    if "--only-mel" in sys.argv:
        gen_mel(FusionModel)
        return
    gen_attnpool(FusionModel)
    if "--only-attnpool" in sys.argv:
        return
    gen_mel(FusionModel)
    gen_txpool(FusionModel)
    gen_flat_train(FusionModel)

    # ---- tiny case: everything stored in full (d_model=8, heads=2 as the reference's own shape tests) -------------
    for head in ("concat", "gated"):
        for prior in (True, False):
            torch.manual_seed(10)
            m = build_ref(FusionModel, dv=16, da=8, d_model=8, heads=2, head=head, prior=prior, common=16, prior_dim=4,
                          prior_hidden=8)
            g = torch.Generator().manual_seed(11)
            xv = torch.randn(3, 4, 16, generator=g)
            xa = torch.randn(3, 12, 8, generator=g)
            labels = torch.tensor([1, 7, 3])
            sd0 = fusion_state(m)
            logits, loss, grads, after = run_case(m, xv, xa, labels)
            blob = {"xv": xv.numpy(), "xa": xa.numpy(), "labels": labels.numpy(), "logits": logits.numpy(),
                    "loss": loss.numpy()}
            blob.update({"w:" + k: v.numpy() for k, v in sd0.items()})
            blob.update({"g:" + k: v.numpy() for k, v in grads.items()})
            blob.update({"a:" + k: v.numpy() for k, v in after.items()})
            np.savez_compressed(OUT / f"xattn_tiny_{head}_{'prior' if prior else 'noprior'}.npz", **blob)

    # ---- BASELINE config 1: B=8, [8,16,512] + [8,149,768]; digests of weights/inputs, full logits ------------------
    blob = {}
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    labels = torch.randint(0, 8, (8,), generator=g)
    blob["labels"] = labels.numpy()
    blob["sha_xv"] = np.array(sha(xv))
    blob["sha_xa"] = np.array(sha(xa))
    for head in ("concat", "gated"):
        for prior in (True, False):
            tag = f"{head}_{'prior' if prior else 'noprior'}"
            torch.manual_seed(0)
            m = build_ref(FusionModel, dv=512, da=768, d_model=128, heads=4, head=head, prior=prior)
            sd0 = fusion_state(m)
            logits, loss, grads, after = run_case(m, xv, xa, labels)
            blob[f"{tag}:logits"] = logits.numpy()
            blob[f"{tag}:loss"] = loss.numpy()
            blob[f"{tag}:keys"] = np.array(sorted(sd0))
            blob[f"{tag}:shapes"] = np.array([str(tuple(sd0[k].shape)) for k in sorted(sd0)])
            blob[f"{tag}:sha"] = np.array([sha(sd0[k]) for k in sorted(sd0)])
            blob[f"{tag}:gkeys"] = np.array(sorted(grads))
            blob[f"{tag}:gdigest"] = np.stack([digest(grads[k]) for k in sorted(grads)])
            blob[f"{tag}:adigest"] = np.stack([digest(after[k]) for k in sorted(grads)])
    np.savez_compressed(OUT / "xattn_config1.npz", **blob)

    # ---- flat concat/gated heads with CLIP-style alignment (config 5), pooled inputs ----------------------------------
    blob = {}
    g = torch.Generator().manual_seed(3)
    ap = torch.randn(6, 768, generator=g)
    vp = torch.randn(6, 512, generator=g)
    blob["sha_ap"] = np.array(sha(ap))
    blob["sha_vp"] = np.array(sha(vp))
    for mode in ("concat", "gated"):
        torch.manual_seed(0)
        m = build_ref(FusionModel, dv=512, da=768, d_model=128, heads=4, head="concat", prior=False, mode=mode, align="clip")
        m.eval()
        sd0 = fusion_state(m)
        with torch.no_grad():
            logits = m(vp, ap)
            loss = m.pop_alignment_loss()
        blob[f"{mode}:logits"] = logits.numpy()
        blob[f"{mode}:align_loss"] = loss.numpy()
        blob[f"{mode}:keys"] = np.array(sorted(sd0))
        blob[f"{mode}:sha"] = np.array([sha(sd0[k]) for k in sorted(sd0)])
    # small fully-stored flat case
    for mode in ("concat", "gated"):
        torch.manual_seed(20)
        m = build_ref(FusionModel, dv=12, da=10, d_model=8, heads=2, head="concat", prior=False, mode=mode, align="clip", common=16)
        m.eval()
        g2 = torch.Generator().manual_seed(21)
        xa = torch.randn(3, 5, 10, generator=g2)
        xv = torch.randn(3, 4, 12, generator=g2)
        with torch.no_grad():
            logits = m(xv.mean(dim=1), xa.mean(dim=1))
            loss = m.pop_alignment_loss()
        sd0 = fusion_state(m)
        small = {"xa": xa.numpy(), "xv": xv.numpy(), "logits": logits.numpy(), "align_loss": loss.numpy()}
        small.update({"w:" + k: v.numpy() for k, v in sd0.items()})
        np.savez_compressed(OUT / f"flat_tiny_{mode}.npz", **small)
    np.savez_compressed(OUT / "flat_config5.npz", **blob)
    print("golden fixtures written to", OUT)
    for f in sorted(OUT.glob("*.npz")):
        print(f"  {f.name}: {f.stat().st_size} bytes")


if __name__ == "__main__":
    main()
