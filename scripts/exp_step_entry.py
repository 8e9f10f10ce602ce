"""Experiment: separate calls vs xattn_stage1_step(use_graph=0) vs (use_graph=1) from identical state: max differences."""
import ctypes as C, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import torch
from helpers import make_head
from multimodalemotionrecognition_b200 import _lib
from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr, _io_dtype

dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(5)
B = 48
xv = torch.randn(B, 16, 512, generator=g).bfloat16().to(dev)
xa = torch.randn(B, 149, 768, generator=g).bfloat16().to(dev)
lab = torch.randint(0, 8, (B,), generator=g).to(dev)
lib = _lib.load()
for prec in ("fp32", "tf32"):
    res = {}
    for mode in ("separate", "entry_eager", "entry_graph", "entry_graph2"):
        torch.manual_seed(1234)
        head = make_head("concat", True, seed=0, precision=prec).to(dev).train()
        flat = head.flat_parameters()
        head._rng_offset = 7
        cfg = head._make_cfg(B, 16, 149, _io_dtype(xv), head._pick_precision(_io_dtype(xv), True, 16, 149), True, True)
        ws = torch.zeros((_lib.workspace_bytes(cfg),), dtype=torch.uint8, device=dev)
        logits = torch.empty((B, 8), device=dev); loss = torch.empty((), device=dev); dl = torch.empty((B, 8), device=dev)
        grads = torch.zeros_like(flat); m = torch.zeros_like(flat); v = torch.zeros_like(flat)
        if mode == "separate":
            _lib.check(lib.xattn_forward(C.byref(cfg), _ptr(flat), None, _ptr(xv), _ptr(xa), _ptr(logits), _ptr(ws), ws.numel(), _stream_ptr(dev)))
            _lib.check(lib.xattn_ce_loss_fwd_bwd(_ptr(logits), _ptr(lab), B, 8, 0.05, 1.0, _ptr(loss), _ptr(dl), _stream_ptr(dev)))
            _lib.check(lib.xattn_backward(C.byref(cfg), _ptr(flat), _ptr(xv), _ptr(xa), _ptr(dl), _ptr(grads), _ptr(ws), ws.numel(), _stream_ptr(dev)))
            _lib.check(lib.xattn_adam_step(_ptr(flat), _ptr(grads), _ptr(m), _ptr(v), flat.numel(), 1e-3, 0.9, 0.999, 1e-3, 1e-4, 2, 0, _stream_ptr(dev)))
        else:
            a = _lib.XAttnStepArgs()
            a.cfg = C.pointer(cfg)
            a.params, a.video_emb, a.audio_emb, a.labels = _ptr(flat), _ptr(xv), _ptr(xa), _ptr(lab)
            a.logits, a.loss, a.dlogits, a.grads = _ptr(logits), _ptr(loss), _ptr(dl), _ptr(grads)
            a.exp_avg, a.exp_avg_sq = _ptr(m), _ptr(v)
            a.workspace, a.workspace_bytes = _ptr(ws), ws.numel()
            a.lr, a.beta1, a.beta2, a.eps, a.weight_decay, a.label_smoothing = 1e-3, 0.9, 0.999, 1e-3, 1e-4, 0.05
            a.step, a.decoupled, a.use_graph = 2, 0, 0 if mode == "entry_eager" else 1
            _lib.check(lib.xattn_stage1_step(C.byref(a), _stream_ptr(dev)))
        torch.cuda.synchronize()
        res[mode] = (float(loss), logits.clone(), grads.clone(), flat.clone(), m.clone())
    ref = res["separate"]
    names = head._xattn_param_names()
    offs, sizes, _ = _lib.xattn_param_layout(cfg)
    for n, o, sz in zip(names, offs, sizes):
        if sz:
            d = float((ref[2][o:o + sz] - res["entry_graph"][2][o:o + sz]).abs().max())
            if d > 1e-6:
                print("   ", prec, n, "max |dgrad|", d, " |grad| max", float(ref[2][o:o + sz].abs().max()))
    for mode in ("entry_eager", "entry_graph", "entry_graph2"):
        r = res[mode]
        print(prec, mode, "loss", ref[0], r[0], " dlogits", float((ref[1] - r[1]).abs().max()), " dgrad", float((ref[2] - r[2]).abs().max()),
              " dparam", float((ref[3] - r[3]).abs().max()), " dm", float((ref[4] - r[4]).abs().max()))
