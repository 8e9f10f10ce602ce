"""Minimal driver for ncu: a few bf16 forwards at B clips (no timing, no CPU baseline)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import XAttnFusionHead

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="bf16").to(dev).eval()
g = torch.Generator(device=dev).manual_seed(1)
xv = torch.randn(B, 16, 512, device=dev, generator=g).bfloat16()
xa = torch.randn(B, 149, 768, device=dev, generator=g).bfloat16()
with torch.no_grad():
    for _ in range(n):
        out = head(xv, xa)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()))
