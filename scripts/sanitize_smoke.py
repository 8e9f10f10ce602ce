"""Smallest program that launches every kernel family once, for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck|racecheck|synccheck python scripts/sanitize_smoke.py
bf16 tcgen05 forward (B=8), tf32 Stage-1 step with the mma.sync attention (B=8, dropouts on), fp32 forward, flat heads."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import FlatFusionHead, XAttnFusionHead
from multimodalemotionrecognition_b200.training import Stage1Trainer

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(1)
B = 8
xv = torch.randn(B, 16, 512, generator=g).bfloat16().to(dev)
xa = torch.randn(B, 149, 768, generator=g).bfloat16().to(dev)
lab = torch.randint(0, 8, (B,), generator=g).to(dev)
torch.manual_seed(0)
with torch.no_grad():
    out = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="bf16").to(dev).eval()(xv, xa)
    out32 = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="fp32").to(dev).eval()(xv.float(), xa.float())
    flat = FlatFusionHead(num_classes=8, mode="gated", fusion_align_mode="clip").to(dev).eval()(xv, xa)
head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="tf32").to(dev).train()
tr = Stage1Trainer(head)
for _ in range(2):
    loss = tr.step(xv, xa, lab, training=True)
torch.cuda.synchronize()
print("ok", float(out.abs().sum()), float(out32.abs().sum()), float(flat.abs().sum()), float(loss))
