"""Minimal driver for ncu: a few Stage-1 training steps at B clips (no timing)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import XAttnFusionHead
from multimodalemotionrecognition_b200.training import Stage1Trainer

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
torch.manual_seed(0)
prec = sys.argv[3] if len(sys.argv) > 3 else "fp32"
head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision=prec).to(dev).train()
tr = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4)
g = torch.Generator(device=dev).manual_seed(1)
xv = torch.randn(B, 16, 512, device=dev, generator=g).bfloat16()
xa = torch.randn(B, 149, 768, device=dev, generator=g).bfloat16()
lab = torch.randint(0, 8, (B,), device=dev, generator=g)
for _ in range(n):
    loss = tr.step(xv, xa, lab, training=True)
torch.cuda.synchronize()
print("ok", float(loss))
