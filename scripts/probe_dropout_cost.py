#!/usr/bin/env python
"""How much of the Stage-1 step is the attention-dropout RNG: the tf32 step at B=1024 with p_attn 0.1 and 0.0."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from multimodalemotionrecognition_b200 import XAttnFusionHead
from multimodalemotionrecognition_b200.training import Stage1Trainer

dev = torch.device("cuda", 0)
B = 1024
g = torch.Generator(device=dev).manual_seed(1)
sets = [(torch.randn(B, 16, 512, device=dev, generator=g).bfloat16(), torch.randn(B, 149, 768, device=dev, generator=g).bfloat16(),
         torch.randint(0, 8, (B,), device=dev, generator=g)) for _ in range(2)]
for p in (0.1, 0.0):
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="tf32", xattn_attn_dropout=p).to(dev).train()
    tr = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4)
    for i in range(5):
        tr.step(*sets[i % 2], training=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        tr.step(*sets[i % 2], training=True)
    e1.record()
    torch.cuda.synchronize()
    print(f"p_attn={p}: {e0.elapsed_time(e1) / 20:.3f} ms/step", flush=True)
