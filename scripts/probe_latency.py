"""Where does a small-batch forward spend its time?  host issue time per call vs per-kernel device time."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import XAttnFusionHead, _lib
dev = torch.device("cuda:0")
torch.manual_seed(0)
head = XAttnFusionHead(num_classes=8, audio_n_mels=768, xattn_use_emotion_prior=True, precision="bf16").to(dev).eval()
for B in (1, 64, 1024):
    xv = torch.randn(B, 16, 512, device=dev).bfloat16(); xa = torch.randn(B, 149, 768, device=dev).bfloat16()
    with torch.no_grad():
        for _ in range(20): head(xv, xa)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(200): head(xv, xa)
        t_host = (time.perf_counter() - t0) / 200
        torch.cuda.synchronize()
        t_all = (time.perf_counter() - t0) / 200
        _lib.stage_timing(True)
        for _ in range(50): head(xv, xa)
        torch.cuda.synchronize()
        rec = _lib.stage_timing_read(); _lib.stage_timing(False)
    print(f"B={B}: host issue {t_host*1e6:.1f} us/call, wall {t_all*1e6:.1f} us/call; device us per kernel:",
          {k: round(ms / max(n, 1) * 1e3, 1) for k, (ms, n) in rec.items()}, "sum", round(sum(ms / max(n, 1) for ms, n in rec.values()) * 1e3, 1))
