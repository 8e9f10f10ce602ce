"""B200-native (sm_100a) cross-attention fusion head — drop-in for the xattn path of
Wionerlol/MultimodalEmotionRecognition (``src/models/fusion.py``)."""
from .fusion import FlatFusionHead, FusionModel, XAttnFusionHead, softmax_top1  # noqa: F401

__all__ = ["XAttnFusionHead", "FusionModel", "FlatFusionHead", "softmax_top1"]
