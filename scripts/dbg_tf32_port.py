"""Diagnostic: per-parameter gradient agreement of the CUDA tf32 / fp32 paths with the matched / plain CPU ports."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import torch
from helpers import make_head
from oracle import torch_port
from multimodalemotionrecognition_b200.training import Stage1Trainer

io = sys.argv[1] if len(sys.argv) > 1 else "fp32"
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(3)
B = 96
xv = torch.randn(B, 16, 512, generator=g); xa = torch.randn(B, 149, 768, generator=g); labels = torch.randint(0, 8, (B,), generator=g)
if io == "bf16":
    xv, xa = xv.bfloat16(), xa.bfloat16()
from helpers import ws_f32_views
m = make_head("concat", True, seed=0, precision="tf32").to(dev).train()
with torch.no_grad():
    lg, (cfg, _, _, ws) = m._launch_forward(xv.to(dev), xa.to(dev), save=True, training=False)
torch.cuda.synchronize()
w = ws_f32_views(ws, cfg)
taps = {}
sd0 = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
ref = torch_port.xattn_forward_tf32(sd0, xv.float(), xa.float(), num_heads=4, bf16_io=(io == "bf16"), taps=taps)
for k in ("v0", "aseq", "a0", "v1", "a1", "pooled", "hhid"):
    d = (w[k].cpu() - taps[k]).abs().max().item()
    print(f"fwd stage {k:8s} max|diff| {d:.3e}  (scale {taps[k].abs().max().item():.2f})")
print("fwd s1-v0 (branch)", ((w["s1"].cpu() - w["v0"].cpu()) - taps["s1_branch"]).abs().max().item())
print("logits", (lg.cpu() - ref).abs().max().item())
res = {}
for prec in ("fp32", "tf32"):
    m = make_head("concat", True, seed=0, precision=prec).to(dev).train()
    tr = Stage1Trainer(m)
    tr.forward_backward(xv.to(dev), xa.to(dev), labels.to(dev), training=False)
    res["cuda_" + prec] = {k: p.grad.detach().cpu().clone() for k, p in m.named_parameters() if p.grad is not None}
for name, fn in (("port_fp32", lambda sd: torch_port.xattn_forward(sd, xv.float(), xa.float(), num_heads=4)),
                 ("port_tf32", lambda sd: torch_port.xattn_forward_tf32(sd, xv.float(), xa.float(), num_heads=4, bf16_io=(io == "bf16"),
                                                                          relu_masks={"head": w["hhid"].cpu() > 0, "prior": w["phid"].cpu() > 0}))):
    sd = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    torch.nn.functional.cross_entropy(fn(sd), labels).backward()
    res[name] = {k: v.grad for k, v in sd.items() if v.grad is not None}
def rel(a, b):
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-20)
def l2(a, b):
    return ((a - b).double().norm() / b.double().norm().clamp_min(1e-30)).item()
print(f"{'param':48s} {'L2 cudaTF-portTF':>18s} {'L2 cudaTF-cudaFP':>18s} {'L2 portTF-portFP':>18s}")
for k in res["cuda_tf32"]:
    print(f"{k:48s} {l2(res['cuda_tf32'][k], res['port_tf32'][k]):18.2e} {l2(res['cuda_tf32'][k], res['cuda_fp32'][k]):18.2e} "
          f"{l2(res['port_tf32'][k], res['port_fp32'][k]):18.2e}")
print(f"{'param':48s} {'cudaTF-portTF':>14s} {'cudaFP-portFP':>14s} {'portTF-portFP':>14s} {'cudaTF-cudaFP':>14s}")
for k in res["cuda_tf32"]:
    print(f"{k:48s} {rel(res['cuda_tf32'][k], res['port_tf32'][k]):14.2e} {rel(res['cuda_fp32'][k], res['port_fp32'][k]):14.2e} "
          f"{rel(res['port_tf32'][k], res['port_fp32'][k]):14.2e} {rel(res['cuda_tf32'][k], res['cuda_fp32'][k]):14.2e}")
