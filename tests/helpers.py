"""Shared test helpers: fixture loading, seeded inputs, state_dict conversion.  Nothing here reads /root/reference."""
from __future__ import annotations

import hashlib
from pathlib import Path

import numpy as np
import torch

GOLDEN = Path(__file__).resolve().parent / "golden"


def sha(t) -> str:
    a = t.detach().cpu().contiguous().numpy() if isinstance(t, torch.Tensor) else np.ascontiguousarray(t)
    return hashlib.sha256(a.tobytes()).hexdigest()


def load_npz(name: str) -> dict:
    with np.load(GOLDEN / name, allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def fusion_state_np(module) -> dict:
    return {k: v.detach().cpu().numpy() for k, v in module.state_dict().items()
            if not (k.startswith("audio_model.") or k.startswith("video_model."))}


def config1_inputs():
    """BASELINE config 1 inputs: torch.Generator().manual_seed(1), N(0,1)."""
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    labels = torch.randint(0, 8, (8,), generator=g)
    return xv, xa, labels


def make_head(head: str, prior: bool, seed: int = 0, **kw):
    """Canonical-dimension head with weights from torch.manual_seed(seed) in reference constructor order."""
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    torch.manual_seed(seed)
    kw.setdefault("audio_n_mels", 768)
    return XAttnFusionHead(num_classes=8, xattn_head=head, xattn_use_emotion_prior=prior, **kw)


def tiny_head(blob: dict, head: str, prior: bool):
    """The tiny golden case (d_model=8, heads=2) with the stored reference weights loaded."""
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, xattn_head=head, d_model=8, num_heads=2,
                        audio_n_mels=8, xattn_use_emotion_prior=prior, xattn_emotion_prior_dim=4,
                        xattn_emotion_prior_hidden_dim=8)
    sd = {k[2:]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("w:")}
    m.load_state_dict(sd, strict=True)
    return m


def digest(t) -> np.ndarray:
    t = torch.as_tensor(t).detach().double().flatten().cpu()
    idx = torch.linspace(0, t.numel() - 1, steps=8).long()
    return np.concatenate([[t.sum().item(), t.abs().sum().item(), t.norm().item()], t[idx].numpy()])
