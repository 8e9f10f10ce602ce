import sys; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch
from helpers import make_head
from multimodalemotionrecognition_b200.training import Stage1Trainer
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
B = 96
xv = torch.randn(B, 16, 512, generator=g); xa = torch.randn(B, 149, 768, generator=g); labels = torch.randint(0, 8, (B,), generator=g)
for io in ("bf16", "fp32"):
    a, b = (xv.bfloat16(), xa.bfloat16()) if io == "bf16" else (xv, xa)
    res = {}
    for prec in ("fp32", "tf32"):
        m = make_head("concat", True, seed=0, precision=prec).to(dev).train()
        tr = Stage1Trainer(m)
        loss = tr.forward_backward(a.to(dev), b.to(dev), labels.to(dev), training=False)
        res[prec] = (float(loss), {k: p.grad.detach().cpu().clone() for k, p in m.named_parameters() if p.grad is not None})
    print(io, res["fp32"][0], res["tf32"][0])
    for k in res["fp32"][1]:
        g0, g1 = res["fp32"][1][k], res["tf32"][1][k]
        print(f"  {k:45s} max|g|={g0.abs().max().item():.3e} rel={(g0-g1).abs().max().item()/max(g0.abs().max().item(),1e-12):.3e}")
