"""Experiment: does any kernel of the training path read workspace it has not written?  Forward + CE + backward through the C-ABI
with the workspace pre-filled with 0xFF bytes (NaN) against a zero-filled one; and eager-vs-eager spread of a 3-step trajectory."""
import ctypes as C, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import torch
from helpers import make_head
from multimodalemotionrecognition_b200 import _lib
from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr, _io_dtype
from multimodalemotionrecognition_b200.training import Stage1Trainer

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(5)
B = 48
xv = torch.randn(B, 16, 512, generator=g).bfloat16().to(dev)
xa = torch.randn(B, 149, 768, generator=g).bfloat16().to(dev)
lab = torch.randint(0, 8, (B,), generator=g).to(dev)
lib = _lib.load()
for prec in ("fp32", "tf32"):
    torch.manual_seed(1234)
    head = make_head("concat", True, seed=0, precision=prec).to(dev).train()
    flat = head.flat_parameters()
    out = {}
    for fill in (0x00, 0xFF):
        head._rng_offset = 7
        cfg = head._make_cfg(B, 16, 149, _io_dtype(xv), head._pick_precision(_io_dtype(xv), True, 16, 149), True, True)
        ws = torch.full((_lib.workspace_bytes(cfg),), fill, dtype=torch.uint8, device=dev)
        logits = torch.empty((B, 8), device=dev); loss = torch.empty((), device=dev); dl = torch.empty((B, 8), device=dev)
        grads = torch.full_like(flat, float("nan"))
        _lib.check(lib.xattn_forward(C.byref(cfg), _ptr(flat), None, _ptr(xv), _ptr(xa), _ptr(logits), _ptr(ws), ws.numel(), _stream_ptr(dev)))
        _lib.check(lib.xattn_ce_loss_fwd_bwd(_ptr(logits), _ptr(lab), B, 8, 0.05, 1.0, _ptr(loss), _ptr(dl), _stream_ptr(dev)))
        _lib.check(lib.xattn_backward(C.byref(cfg), _ptr(flat), _ptr(xv), _ptr(xa), _ptr(dl), _ptr(grads), _ptr(ws), ws.numel(), _stream_ptr(dev)))
        torch.cuda.synchronize()
        out[fill] = (logits.clone(), grads.clone())
    lg0, g0 = out[0x00]; lg1, g1 = out[0xFF]
    print(prec, "finite with poisoned ws:", bool(torch.isfinite(lg1).all()), bool(torch.isfinite(g1).all()),
          " max |dlogits|", float((lg0 - lg1).abs().max()), " max |dgrad|", float((g0 - g1).abs().max()), " grad max", float(g0.abs().max()))
    # eager vs eager spread
    res = []
    for rep in range(2):
        torch.manual_seed(1234)
        h = make_head("concat", True, seed=0, precision=prec).to(dev).train()
        tr = Stage1Trainer(h, lr=1e-3, weight_decay=1e-4, eps=1e-3, label_smoothing=0.05, graph=False)
        ls = [float(tr.step(xv, xa, lab, training=True)) for _ in range(4)]
        res.append(ls)
    print(prec, "eager vs eager losses:", res[0], res[1], [abs(a - b) for a, b in zip(*res)])
