"""GPU tests (``-m gpu``) of the bf16 tensor-core path: the tcgen05/TMA projection GEMM in isolation, then the whole
bf16 forward against the quantisation-matched oracle.  Tolerance: BASELINE.json's 1e-2 max-abs in bf16."""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import config1_inputs, fusion_state_np, make_head
from oracle import torch_port, xattn_oracle

pytestmark = pytest.mark.gpu
BF16_TOL = 1e-2


def _gemm_tc(x, w, bias, out_bf16, pitch=None):
    from multimodalemotionrecognition_b200 import _lib
    from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr

    M, K = x.shape
    N = w.shape[0]
    pitch = N if pitch is None else pitch
    out = torch.zeros((M, pitch), dtype=torch.bfloat16 if out_bf16 else torch.float32, device=x.device)
    _lib.debug_check(_lib.load_debug().xattn_debug_gemm_tc(_ptr(x), _ptr(w), _ptr(bias), _ptr(out), int(out_bf16), pitch, M, N, K,
                                               _stream_ptr(x.device)))
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("M,N,K,out_bf16", [(128, 128, 64, False), (128, 128, 768, False), (149 * 3, 128, 768, True),
                                            (16 * 5, 128, 512, False), (1000, 256, 128, True), (4768 * 128 + 5, 128, 768, True)])
def test_tcgen05_projection_gemm_vs_fp32(cuda_device, M, N, K, out_bf16):
    g = torch.Generator(device=cuda_device).manual_seed(M + N + K)
    x = torch.randn(M, K, device=cuda_device, generator=g).bfloat16()
    w = (torch.randn(N, K, device=cuda_device, generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device=cuda_device, generator=g)
    pitch = N + 8 if out_bf16 else N
    out = _gemm_tc(x, w, bias, out_bf16, pitch)
    rows = slice(0, M) if M <= 4096 else torch.cat([torch.arange(0, 300), torch.arange(M - 300, M)]).to(cuda_device)
    ref = x[rows].float() @ w.float().t() + bias     # fp32 reference of the same op on the same bf16-rounded operands
    got = out[rows][:, :N].float()
    tol = 2e-2 if out_bf16 else 2e-4                 # bf16 output rounding of O(1) values vs fp32 accumulation order
    err = (got - ref).abs().max().item()
    assert err <= tol, err
    if out_bf16:
        assert (out[rows][:, N:] == 0).all()         # the pad columns of the pitched layout are never written


# ---------------------------------------------------------------------------------------------------------------------
# whole bf16 forward
# ---------------------------------------------------------------------------------------------------------------------
def _al256(n):
    return (n + 255) & ~255


def _ws_views(ws, B, T, Ta):
    """Mirror of compute_ws_bf16 (csrc/bf16_pack.cu): typed views of the workspace intermediates."""
    E, EP = 128, 128
    spec = [("v0", torch.float32, (B, T, E)), ("a0", torch.bfloat16, (B * Ta, EP)), ("qt", torch.bfloat16, (B, 64, EP)),
            ("meanv", torch.float32, (B, E)), ("tqv", torch.float32, (B, 16)), ("tkv", torch.float32, (B, 16)),
            ("u", torch.bfloat16, (B, 64, E)), ("tok", torch.float32, (2, B, 160)), ("acol", torch.float32, (B, 6, E)),
            ("pc", torch.float32, (B, 4)),
            ("kt", torch.bfloat16, (B, 64, EP)), ("vt", torch.bfloat16, (B, 64, EP)), ("cv", torch.float32, (B, 64)),
            ("pooled", torch.float32, (B, 2 * E)), ("poolp", torch.float32, (B, 8, E)), ("v1", torch.bfloat16, (B * 16, E))]
    out, off = {}, 0
    for name, dt, shape in spec:
        n = int(np.prod(shape)) * (4 if dt == torch.float32 else 2)
        out[name] = ws[off:off + n].view(dt).view(shape).float().cpu().numpy()
        off += _al256(n)
    return out


def _stage_report(m, xvb, xab, head, prior, dev):
    """Runs the bf16 path and returns {stage: max-abs error vs the fp64 oracle on the same bf16-rounded inputs}."""
    sd = fusion_state_np(m)
    taps = {}
    ref = xattn_oracle.xattn_forward(sd, xvb.float().numpy(), xab.float().numpy(), num_heads=4, head=head, use_prior=prior,
                                     taps=taps)
    with torch.no_grad():
        logits, (cfg, _, _, ws) = m._launch_forward(xvb.to(dev), xab.to(dev), save=False, training=False)
    torch.cuda.synchronize()
    B, T, Ta = xvb.shape[0], xvb.shape[1], xab.shape[1]
    w = _ws_views(ws, B, T, Ta)
    rep = {}
    rep["v0"] = np.abs(w["v0"] - taps["v"]).max()
    rep["a0"] = np.abs(w["a0"][:, :128].reshape(B, Ta, 128) - taps["a"]).max()
    _, p1 = xattn_oracle.multihead_attention(sd, "v2a_attn.", taps["v"], taps["a"], 4, taps["v2a_bias"], return_probs=True)
    u_ref = np.einsum("bhls,bse->bhle", p1, taps["a"])                       # U_h = P_h · a0
    # U is stored row-major as [16*clip + i][128*h + e] (the A operand of the video_mix GEMM)
    u_got = w["u"].reshape(B, 16, 4, 128).transpose(0, 2, 1, 3)
    rep["u"] = np.abs(u_got[:, :, :T] - u_ref).max()
    rep["v1"] = np.abs(w["v1"].reshape(B, 16, 128)[:, :T] - taps["v_new"]).max()
    rep["pooled_v"] = np.abs(w["pooled"][:, :128] - taps["v_emb"]).max()
    rep["pooled_a"] = np.abs(w["pooled"][:, 128:] - taps["a_emb"]).max()
    rep["logits"] = np.abs(logits.cpu().numpy() - ref).max()
    return rep, logits.cpu().numpy(), ref


@pytest.mark.parametrize("head,prior", [("concat", True), ("gated", True), ("concat", False), ("gated", False)])
def test_bf16_forward_config1_stage_by_stage(cuda_device, head, prior):
    m = make_head(head, prior, seed=0, precision="bf16").to(cuda_device).eval()
    xv, xa, _ = config1_inputs()
    rep, out, ref = _stage_report(m, xv.bfloat16(), xa.bfloat16(), head, prior, cuda_device)
    print("bf16 stage errors:", {k: float(f"{v:.3e}") for k, v in rep.items()})
    for k, v in rep.items():
        assert np.isfinite(v) and v <= 3e-2, (k, rep)     # intermediates are bf16-rounded O(1) values
    assert rep["logits"] <= BF16_TOL, rep


@pytest.mark.parametrize("B,T,Ta", [(1, 16, 149), (3, 8, 100), (5, 1, 1), (2, 16, 160), (9, 7, 33)])
def test_bf16_forward_ragged_shapes(cuda_device, B, T, Ta):
    m = make_head("concat", True, seed=3, precision="bf16").to(cuda_device).eval()
    g = torch.Generator().manual_seed(200 + B)
    xv = torch.randn(B, T, 512, generator=g).bfloat16()
    xa = torch.randn(B, Ta, 768, generator=g).bfloat16()
    rep, out, ref = _stage_report(m, xv, xa, "concat", True, cuda_device)
    assert rep["logits"] <= BF16_TOL, rep


def test_bf16_unsupported_shapes_raise_or_route(cuda_device):
    m = make_head("concat", True, precision="bf16").to(cuda_device).eval()
    with pytest.raises(ValueError), torch.no_grad():   # forced bf16 outside the specialisation: clean error, no fallback
        m(torch.zeros(2, 17, 512, device=cuda_device, dtype=torch.bfloat16), torch.zeros(2, 149, 768, device=cuda_device, dtype=torch.bfloat16))
    auto = make_head("concat", True, precision="auto").to(cuda_device).eval()
    with torch.no_grad():               # auto: the fp32 kernels take shapes the bf16 kernels do not
        out = auto(torch.zeros(2, 17, 512, device=cuda_device, dtype=torch.bfloat16), torch.zeros(2, 149, 768, device=cuda_device, dtype=torch.bfloat16))
    assert out.shape == (2, 8)


def structured_inputs(B, seed=11):
    """SURVEY 8d "structured" distribution: class c ~ U{0..7} per clip, x = N(0,1) + 2*mu_c with fixed per-modality unit
    vectors mu_c (seed 2) — inputs that carry a class signal, so top-1 is meaningful under random-init weights."""
    g2 = torch.Generator().manual_seed(2)
    mu_v = torch.nn.functional.normalize(torch.randn(8, 512, generator=g2), dim=1)
    mu_a = torch.nn.functional.normalize(torch.randn(8, 768, generator=g2), dim=1)
    g = torch.Generator().manual_seed(seed)
    cls = torch.randint(0, 8, (B,), generator=g)
    xv = torch.randn(B, 16, 512, generator=g) + 2.0 * mu_v[cls][:, None, :]
    xa = torch.randn(B, 149, 768, generator=g) + 2.0 * mu_a[cls][:, None, :]
    return xv.bfloat16(), xa.bfloat16(), cls


def matched_reference(m, xv, xa):
    """The quantisation-matched oracle of SURVEY 7.2: the fp32 port of the reference fed the same bf16-rounded inputs AND
    bf16-rounded weights (the kernels read every matrix as bf16; vectors — biases, LayerNorm affine, score heads — stay fp32)."""
    sd = {k: (v.detach().cpu().bfloat16().float() if v.ndim >= 2 else v.detach().cpu().clone()) for k, v in m.state_dict().items()}
    with torch.inference_mode():
        return torch_port.xattn_forward(sd, xv.float(), xa.float(), num_heads=4)


def top1_report(out, ref):
    err = (out - ref).abs().max().item()
    top2 = ref.topk(2, dim=1).values
    margin = top2[:, 0] - top2[:, 1]
    agree = (out.argmax(1) == ref.argmax(1)).float()
    return {"max_abs_err": err, "raw_top1_agreement": agree.mean().item(), "disagreeing_clips": int((1 - agree).sum().item()),
            "near_tie_fraction_margin_lt_2err": (margin < 2 * err).float().mean().item(),
            "largest_margin_among_disagreements": (margin[agree == 0].max().item() if (agree == 0).any() else 0.0),
            "median_margin": margin.median().item()}


def test_bf16_full_batch_top1_and_tolerance(cuda_device):
    """BASELINE configs[1] at B=4096 against the quantisation-matched oracle.  Gates, all on RAW top-1 agreement (no filtering):
      * trained weights (the head fitted for 150 Stage-1 steps to the class-structured inputs of SURVEY 8d, so its margins are
        those of a model that has learnt something): >= 99.9 % — north_star's bar, on a fixture where top-1 means something;
      * random-init weights, iid N(0,1) and structured inputs: the top1-top2 margin is below 1e-3 for ~1-2 % of the clips by
        construction (torch's own bf16 run of the reference agrees with fp32 on 99.7 %): >= 99.5 %,
        and every disagreeing clip must be a near-tie (reference margin below twice the measured max error);
      * max-abs logit error <= 1e-2 everywhere.
    The numbers are written to gpurun_out/bf16_top1_report.json (kept under profiles/).  Also: determinism, clip independence."""
    import json
    from pathlib import Path

    m = make_head("concat", True, seed=0, precision="bf16").to(cuda_device).eval()
    B = 4096
    g = torch.Generator().manual_seed(11)
    xv = torch.randn(B, 16, 512, generator=g).bfloat16()
    xa = torch.randn(B, 149, 768, generator=g).bfloat16()
    with torch.no_grad():
        out = m(xv.to(cuda_device), xa.to(cuda_device))
        again = m(xv.to(cuda_device), xa.to(cuda_device))
        part = m(xv[777:900].to(cuda_device), xa[777:900].to(cuda_device))
    assert torch.equal(out, again)                       # deterministic: fixed-order reductions, no atomics
    # Clips are independent.  A clip's position in the batch decides which row tiles of the a2v kernel it lands in, hence
    # the order of its fp32 partial sums: equal up to rounding, not bitwise.
    shift = (out[777:900] - part).abs().max().item()
    assert shift <= 2e-5, shift
    report = {"batch_position_sensitivity": shift}
    report["iid"] = top1_report(out.cpu(), matched_reference(m, xv, xa))
    sxv, sxa, _ = structured_inputs(B)
    with torch.no_grad():
        sout = m(sxv.to(cuda_device), sxa.to(cuda_device))
    report["structured"] = top1_report(sout.cpu(), matched_reference(m, sxv, sxa))
    # trained fixture: fit the head to the structured data with the Stage-1 trainer, then compare on fresh clips
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    mt = make_head("concat", True, seed=0, precision="tf32").to(cuda_device).train()
    tr = Stage1Trainer(mt, lr=1e-3, weight_decay=1e-4)
    for step in range(150):
        bxv, bxa, bcls = structured_inputs(256, seed=1000 + step)
        tr.step(bxv.to(cuda_device), bxa.to(cuda_device), bcls.to(cuda_device), training=True)
    mt.eval()
    mt.precision = "bf16"
    txv, txa, tcls = structured_inputs(B, seed=7)
    with torch.no_grad():
        tout = mt(txv.to(cuda_device), txa.to(cuda_device)).cpu()
    tref = matched_reference(mt, txv, txa)
    report["trained"] = top1_report(tout, tref)
    report["trained"]["accuracy_b200"] = (tout.argmax(1) == tcls).float().mean().item()
    report["trained"]["accuracy_oracle"] = (tref.argmax(1) == tcls).float().mean().item()
    outdir = Path(__file__).resolve().parents[1] / "gpurun_out"
    if outdir.is_dir():
        (outdir / "bf16_top1_report.json").write_text(json.dumps(report, indent=1))
    print("bf16 B=4096 top-1 report:", json.dumps(report))
    report["trained"]["max_abs_logit"] = tref.abs().max().item()
    for name, r in report.items():
        if name == "batch_position_sensitivity":
            continue
        # 1e-2 absolute is north_star's bar at random-init scale (|logit| < 1); the trained head's logits are ~20x larger
        assert r["max_abs_err"] <= BF16_TOL * max(1.0, r.get("max_abs_logit", 0.0) / 4), (name, r)
        assert r["largest_margin_among_disagreements"] <= 2 * r["max_abs_err"], (name, r)   # only near-ties may flip
    assert report["trained"]["raw_top1_agreement"] >= 0.999, report
    assert report["trained"]["accuracy_oracle"] >= 0.9, report          # the fixture really learnt the classes
    assert report["structured"]["raw_top1_agreement"] >= 0.995, report
    assert report["iid"]["raw_top1_agreement"] >= 0.995, report


@pytest.mark.parametrize("mode,N,K", [(0, 160, 128), (1, 128, 160), (0, 64, 128), (1, 128, 64), (2, 160, 128), (3, 128, 160),
                                      (3, 128, 64)])
def test_umma_probe_m64_halves_and_mn_major(cuda_device, mode, N, K):
    """tcgen05 features of the attention kernels, pinned in isolation: two M=64 products land in lanes 0-15 / 16-31 of each
    TMEM sub-partition; mode 1 feeds B as an MN-major (rows = K) swizzled tile; 16x256b loads return mma C-fragment order."""
    from multimodalemotionrecognition_b200 import _lib
    lib = _lib.load_debug()
    g = torch.Generator().manual_seed(mode * 7 + N + K)
    A = torch.randn(128, K, generator=g).bfloat16()
    ts, mode_b = mode >= 2, mode & 1          # modes 2/3: A operand in tensor memory
    B = (torch.randn(2, N, K, generator=g) if mode_b == 0 else torch.randn(2, K, N, generator=g)).bfloat16()
    Ad, Bd = A.to(cuda_device), B.to(cuda_device)
    out32 = torch.full((128, N), float("nan"), device=cuda_device)
    out16 = torch.full((2, 4, 32, 64), float("nan"), device=cuda_device)
    _lib.debug_check(lib.xattn_debug_umma_probe(Ad.data_ptr(), Bd.data_ptr(), out32.data_ptr(), out16.data_ptr(), N, K, mode,
                                          torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = []
    for p in range(2):
        a = A[64 * p:64 * p + 64].double()
        b = B[p].double()
        ref.append(a @ (b.T if mode_b == 0 else b))
    got = out32.cpu().double().reshape(4, 2, 16, N)          # [sub-partition][half][lane in half][col]
    for p in range(2):
        want = ref[p].reshape(4, 16, N)                      # row m -> sub-partition m // 16, lane m % 16 (+16 for the upper half)
        err = (got[:, p] - want).abs().max().item()
        assert err < 1e-3 * max(1.0, want.abs().max().item()), (mode, ts, p, err)
    if N >= 128:
        o = out16.cpu().double()                             # [half][warp][lane][4*i + r]
        for p in range(2):
            want = ref[p].reshape(4, 16, N)
            for lane in range(32):
                g_, t_ = lane >> 2, lane & 3
                for i in range(16):
                    for r in range(4):
                        w = want[:, g_ + 8 * (r >> 1), 8 * i + 2 * t_ + (r & 1)]
                        assert (o[p, :, lane, 4 * i + r] - w).abs().max().item() < 1e-3 * max(1.0, w.abs().max().item()), (p, lane, i, r)


@pytest.mark.parametrize("B", [1, 8, 37])
def test_graphed_forward_and_small_batch_host_path(cuda_device, B):
    """SURVEY 8f rank 2 (worker batch path, optimized_runtime.py:99-108): one CUDA graph per batch size.  The graph replays the
    same deterministic kernels, so its logits are bit-identical to the eager launch; predict_host routes worker-sized
    batches through it; an in-place parameter update re-captures (the graph holds the packed-weight pointers)."""
    m = make_head("concat", True, seed=0, precision="bf16").to(cuda_device).eval()
    g = torch.Generator().manual_seed(5)
    xv = torch.randn(B, 16, 512, generator=g).bfloat16()
    xa = torch.randn(B, 149, 768, generator=g).bfloat16()
    with torch.no_grad():
        eager = m(xv.to(cuda_device), xa.to(cuda_device))
        gf = m.graphed_forward(B)
        gf.video.copy_(xv)
        gf.audio.copy_(xa)
        assert torch.equal(gf(), eager)
        assert m.graphed_forward(B) is gf                      # cached
        host = m.predict_host(xv.pin_memory(), xa.pin_memory())
        assert torch.equal(host, eager.cpu())
        probs = m.predict_host(xv.pin_memory(), xa.pin_memory(), probs=True)
        assert torch.allclose(probs, torch.softmax(eager, dim=1).cpu(), atol=1e-6)
        big = m.predict_host(xv.pin_memory(), xa.pin_memory(), graph_max=0)   # the chunked pipeline gives the same answer
        assert torch.equal(big, eager.cpu())
        m.v_norm.weight.mul_(1.5)                              # parameters changed in place: stale graph must not be replayed
        eager2 = m(xv.to(cuda_device), xa.to(cuda_device))
        assert not torch.equal(eager2, eager)
        host2 = m.predict_host(xv.pin_memory(), xa.pin_memory())
        assert torch.equal(host2, eager2.cpu())
        assert m.graphed_forward(B) is not gf


def test_softmax_top1_epilogue_and_predict_host_top1(cuda_device):
    """Serving epilogue: xattn_softmax_top1 against torch.softmax / argmax (first maximum on ties), and
    predict_host(top1=True) — probabilities, top-1 class and its probability on the host — on the CUDA-graph path (B=32) and
    the chunked path (B=600), as inference_worker.py:136-143 derives them from predict_probs."""
    from multimodalemotionrecognition_b200 import softmax_top1

    g = torch.Generator(device=cuda_device).manual_seed(5)
    x = torch.randn(1000, 8, device=cuda_device, generator=g) * 4
    x[3] = 1.5                      # all equal: first index wins, as torch.argmax
    x[4, 2] = x[4, 6] = 9.0
    probs, top1, top1p = softmax_top1(x)
    ref = torch.softmax(x, dim=1)
    assert (probs - ref).abs().max() < 1e-6
    assert torch.equal(top1.long(), ref.argmax(1)) and int(top1[3]) == 0 and int(top1[4]) == 2
    assert (top1p - ref.max(1).values).abs().max() < 1e-6
    y = x.clone()
    softmax_top1(y, y)              # in place
    assert torch.equal(y, probs)
    m = make_head("concat", True, seed=0, precision="bf16").to(cuda_device).eval()
    for B in (32, 600):
        xv = torch.randn(B, 16, 512, generator=torch.Generator().manual_seed(B)).bfloat16().pin_memory()
        xa = torch.randn(B, 149, 768, generator=torch.Generator().manual_seed(B + 1)).bfloat16().pin_memory()
        logits = m.predict_host(xv, xa).clone()
        p, t1, t1p = m.predict_host(xv, xa, top1=True)
        assert not p.is_cuda and t1.dtype == torch.int32 and p.shape == (B, 8)
        refp = torch.softmax(logits, dim=1)
        assert (p - refp).abs().max() < 1e-6
        assert torch.equal(t1.long(), refp.argmax(1))
        assert (t1p - refp.max(1).values).abs().max() < 1e-6
