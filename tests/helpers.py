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


def ws_f32_views(ws: torch.Tensor, cfg) -> dict:
    """Typed views of the fp32 / tf32 workspace (mirror of compute_ws_f32, csrc/common.cu): name -> float32 tensor (device).
    Under tf32 the rows of p1 / p2 are padded to a multiple of 4 keys (only [..., :S] is written)."""
    B, T, Ta, E, H = cfg.B, cfg.T, cfg.Ta, cfg.E, cfg.H
    P = max(cfg.prior_dim, 1)
    Hd = max(cfg.prior_hidden, 1)
    HW = E if cfg.head_kind == 1 else cfg.common_dim
    bwd = bool(cfg.training or cfg.save_for_backward)
    pad = (lambda n: (n + 3) & ~3) if cfg.precision == 2 else (lambda n: n)   # P row stride under tf32 (p_row_stride, common.cuh)
    spec = [("v0", (B, T, E)), ("aseq", (B, Ta, E)), ("a0", (B, Ta, E)), ("q1", (B, T, E)), ("kv1", (B, Ta, 2 * E)),
            ("p1", (B, H, T, pad(Ta)) if bwd else None), ("o1", (B, T, E)), ("s1", (B, T, E)), ("mean1", (B, T)), ("rstd1", (B, T)),
            ("v1", (B, T, E)), ("q2", (B, Ta, E)), ("kv2", (B, T, 2 * E)), ("p2", (B, H, Ta, pad(T)) if bwd else None),
            ("o2", (B, Ta, E)), ("s2", (B, Ta, E)), ("mean2", (B, Ta)), ("rstd2", (B, Ta)), ("a1", (B, Ta, E)),
            ("pooled0", (B, 2 * E)), ("phid", (B, Hd)), ("prior", (B, P)), ("qv", (B, T)), ("ka", (B, Ta)), ("qa", (B, Ta)),
            ("kvs", (B, T)), ("pooled", (B, 2 * E)), ("hhid", (B, HW)), ("gate", (B,)), ("fusedg", (B, E)), ("dp1", (B,)),
            ("dp2", (B,))]
    if bwd:
        spec += [("d_pooled", (B, 2 * E)), ("d_hh", (B, HW)), ("d_z", (B,)), ("d_fused", (B, E)), ("d_a0", (B, Ta, E)),
                 ("d_br2", (B, Ta, E)), ("d_o2", (B, Ta, E)), ("d_q2", (B, Ta, E)), ("d_kv2", (B, T, 2 * E)), ("d_v1", (B, T, E)),
                 ("d_v0", (B, T, E)), ("d_br1", (B, T, E)), ("d_o1", (B, T, E)), ("d_q1", (B, T, E)), ("d_kv1", (B, Ta, 2 * E)),
                 ("d_aseq", (B, Ta, E)), ("d_qv", (B, T)), ("d_ka", (B, Ta)), ("d_qa", (B, Ta)), ("d_kvs", (B, T)),
                 ("d_prior", (B, P)), ("d_phid", (B, Hd))]
    spec += [("wbf", ((E * cfg.Da + E * cfg.Dv + 1) // 2 + 8,))]
    pooled_tok = cfg.pool_mode != 0
    spec += [("poolv", (B, E) if pooled_tok else None), ("poola", (B, E) if pooled_tok else None)]
    f = ws.view(torch.float32) if ws.dtype != torch.float32 else ws
    out, off = {}, 0
    for name, shape in spec:
        if shape is None:
            continue
        n = int(np.prod(shape))
        out[name] = f[off:off + n].view(shape)
        off += (n + 3) & ~3
    return out
