"""Experiment: how much of the Stage-1 step is launch gaps?  One step captured in a CUDA graph (FIXED Philox offset and Adam step
count — timing only, not a valid training loop) replayed against the eager step."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import XAttnFusionHead
from multimodalemotionrecognition_b200.training import Stage1Trainer

B = 1024
dev = torch.device("cuda:0")
torch.manual_seed(0)
head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="tf32").to(dev).train()
tr = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4)
g = torch.Generator(device=dev).manual_seed(1)
xv = torch.randn(B, 16, 512, device=dev, generator=g).bfloat16()
xa = torch.randn(B, 149, 768, device=dev, generator=g).bfloat16()
lab = torch.randint(0, 8, (B,), device=dev, generator=g)
for _ in range(5):
    tr.step(xv, xa, lab, training=True)
torch.cuda.synchronize()

def timed(fn, n=30):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

eager = timed(lambda: tr.step(xv, xa, lab, training=True))
s = torch.cuda.Stream(dev)
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(2):
        tr.step(xv, xa, lab, training=True)
torch.cuda.current_stream().wait_stream(s)
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    loss = tr.step(xv, xa, lab, training=True)
graphed = timed(graph.replay)
print(f"eager {eager:.4f} ms/step   graph replay {graphed:.4f} ms/step   loss {float(loss):.4f}")
