"""Builds the reference's own ``FusionModel`` (from ``oracle/_ref``, staged by ``oracle/stage_ref.py``) with identity-stub
backbones, for the timed CPU baseline.  TEST INFRASTRUCTURE ONLY: used by ``bench.py``'s ``cpu_baseline`` leg and
``--impl reference`` arm; falls back to nothing — callers use ``oracle/torch_port.py`` (``kind: "port"``) when the staged
sources are absent.  The stubs satisfy reference fusion.py:368-370 (``video_model.backbone`` on ``[B*T, Dv, 1, 1]``) and
:374-378 (``audio_model.encode_sequence``), as SURVEY 8b describes."""
from __future__ import annotations

import sys
from pathlib import Path

import torch
from torch import nn

REF_DIR = Path(__file__).resolve().parent / "_ref"


def available() -> bool:
    return (REF_DIR / "src" / "models" / "fusion.py").is_file() and (REF_DIR / "src" / "models" / "temporal.py").is_file()


class _StubVideo(nn.Module):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.embedding_dim = dim
        self.backbone = nn.Identity()

    def encode(self, x):
        return x


class _StubAudio(nn.Module):
    def __init__(self, dim: int) -> None:
        super().__init__()
        self.sequence_dim = dim
        self.embedding_dim = dim

    def encode_sequence(self, x):
        return x

    def encode(self, x):
        return x


def build_reference_module(state_dict=None, *, dv: int = 512, da: int = 768, classes: int = 8, head: str = "concat",
                           prior: bool = True) -> nn.Module:
    """The unmodified reference ``FusionModel(mode="xattn")``; call it as ``m(video_emb[..., None, None], audio_emb)``."""
    if str(REF_DIR) not in sys.path:
        sys.path.insert(0, str(REF_DIR))
    from src.models.fusion import FusionModel   # noqa: E402  (the staged reference file)

    m = FusionModel(_StubAudio(da), _StubVideo(dv), classes, mode="xattn", xattn_head=head, d_model=128, num_heads=4,
                    audio_n_mels=da, xattn_use_emotion_prior=prior)
    if state_dict is not None:
        missing, unexpected = m.load_state_dict(state_dict, strict=False)
        assert not unexpected, unexpected
    return m
