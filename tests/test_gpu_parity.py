"""GPU parity tests (``-m gpu``): the CUDA path, called through the C ABI (ctypes, ``_lib``), against the oracle and
the committed reference fixtures.  Tolerances are BASELINE.json's: fp32 <= 1e-4 max-abs logit error, bf16 <= 1e-2."""
import numpy as np
import pytest
import torch

from helpers import config1_inputs, digest, fusion_state_np, load_npz, make_head, tiny_head
from multimodalemotionrecognition_b200 import _lib
from oracle import torch_port, xattn_oracle

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4   # north_star: within 1e-4 max-abs logit error in fp32
BF16_TOL = 1e-2   # north_star: 1e-2 in bf16
VARIANTS = [(h, p) for h in ("concat", "gated") for p in (True, False)]


def _tag(head, prior):
    return f"{head}_{'prior' if prior else 'noprior'}"


def test_extension_is_loaded_from_tree():
    from multimodalemotionrecognition_b200 import _lib

    lib = _lib.load()
    assert lib.xattn_abi_version() == _lib.ABI_VERSION
    assert _lib.LIB_PATH.exists()


@pytest.mark.parametrize("head,prior", VARIANTS)
def test_fp32_forward_tiny_matches_reference(cuda_device, head, prior):
    blob = load_npz(f"xattn_tiny_{_tag(head, prior)}.npz")
    m = tiny_head(blob, head, prior).to(cuda_device).eval()
    with torch.no_grad():
        out = m(torch.from_numpy(blob["xv"]).to(cuda_device), torch.from_numpy(blob["xa"]).to(cuda_device))
    assert out.shape == (3, 8) and out.dtype == torch.float32
    np.testing.assert_allclose(out.cpu().numpy(), blob["logits"], atol=1e-5, rtol=0)


@pytest.mark.parametrize("head,prior", VARIANTS)
def test_fp32_forward_config1_matches_reference(cuda_device, head, prior):
    """BASELINE config 1 (B=8, [8,16,512] + [8,149,768]) against the reference module's own logits."""
    blob = load_npz("xattn_config1.npz")
    xv, xa, _ = config1_inputs()
    m = make_head(head, prior, seed=0).to(cuda_device).eval()
    with torch.no_grad():
        out = m(xv.to(cuda_device), xa.to(cuda_device)).cpu().numpy()
    err = np.abs(out - blob[f"{_tag(head, prior)}:logits"]).max()
    assert err <= FP32_TOL, err
    assert err <= 5e-6, f"fp32 path drifted from the reference more than expected: {err}"


@pytest.mark.parametrize("B,T,Ta", [(1, 16, 149), (3, 8, 100), (5, 1, 1), (2, 33, 7), (257, 16, 149)])
def test_fp32_forward_ragged_shapes_vs_oracle(cuda_device, B, T, Ta):
    m = make_head("concat", True, seed=3).to(cuda_device).eval()
    g = torch.Generator().manual_seed(100 + B)
    xv = torch.randn(B, T, 512, generator=g)
    xa = torch.randn(B, Ta, 768, generator=g)
    with torch.no_grad():
        out = m(xv.to(cuda_device), xa.to(cuda_device)).cpu().numpy()
    n = min(B, 16)   # the fp64 oracle is slow; check a prefix and a suffix
    sd = fusion_state_np(m)
    ref_a = xattn_oracle.xattn_forward(sd, xv[:n].numpy(), xa[:n].numpy(), num_heads=4, head="concat", use_prior=True)
    ref_b = xattn_oracle.xattn_forward(sd, xv[-n:].numpy(), xa[-n:].numpy(), num_heads=4, head="concat", use_prior=True)
    assert np.abs(out[:n] - ref_a).max() <= FP32_TOL
    assert np.abs(out[-n:] - ref_b).max() <= FP32_TOL


def test_empty_batch(cuda_device):
    m = make_head("concat", True).to(cuda_device).eval()
    with torch.no_grad():
        out = m(torch.zeros(0, 16, 512, device=cuda_device), torch.zeros(0, 149, 768, device=cuda_device))
    assert out.shape == (0, 8)


def test_error_conventions(cuda_device):
    m = make_head("concat", False).to(cuda_device).eval()
    with pytest.raises(ValueError):   # reference: TemporalPooler raises ValueError for ndim != 3 (temporal.py:106-107)
        m(torch.zeros(4, 512, device=cuda_device), torch.zeros(4, 149, 768, device=cuda_device))
    with pytest.raises(ValueError):
        m(torch.zeros(4, 16, 500, device=cuda_device), torch.zeros(4, 149, 768, device=cuda_device))
    with pytest.raises(ValueError):
        m(torch.zeros(4, 16, 512, device=cuda_device, dtype=torch.float16), torch.zeros(4, 149, 768, device=cuda_device, dtype=torch.float16))
    with pytest.raises(RuntimeError):  # no CPU path
        m.cpu()(torch.zeros(4, 16, 512), torch.zeros(4, 149, 768))


def test_fp32_kernels_with_bf16_inputs(cuda_device):
    """io_dtype=bf16 with fp32 arithmetic equals the oracle fed the same bf16-rounded inputs."""
    m = make_head("gated", True, seed=5, precision="fp32").to(cuda_device).eval()
    xv, xa, _ = config1_inputs()
    xvb, xab = xv.bfloat16(), xa.bfloat16()
    with torch.no_grad():
        out = m(xvb.to(cuda_device), xab.to(cuda_device)).cpu().numpy()
    ref = xattn_oracle.xattn_forward(fusion_state_np(m), xvb.float().numpy(), xab.float().numpy(), num_heads=4, head="gated")
    assert np.abs(out - ref).max() <= FP32_TOL


def test_clip_independence_and_determinism_full_size(cuda_device):
    """Size-independent property at BASELINE's full B=4096: every clip is independent, so any shard of the
    batch gives the same rows, and repeated launches are bit-identical."""
    m = make_head("concat", True, seed=1).to(cuda_device).eval()
    g = torch.Generator(device=cuda_device).manual_seed(7)
    B = 4096
    xv = torch.randn(B, 16, 512, device=cuda_device, generator=g)
    xa = torch.randn(B, 149, 768, device=cuda_device, generator=g)
    with torch.no_grad():
        full = m(xv, xa)
        again = m(xv, xa)
        part = m(xv[1000:1100], xa[1000:1100])
        tail = m(xv[4090:], xa[4090:])
    assert torch.equal(full, again)
    assert torch.equal(full[1000:1100], part)
    assert torch.equal(full[4090:], tail)
    # spot-check against the fp32 CPU port of the reference on a slice
    ref = torch_port.xattn_forward({k: v.cpu() for k, v in m.state_dict().items()}, xv[:32].cpu(), xa[:32].cpu(), num_heads=4)
    assert (full[:32].cpu() - ref).abs().max() <= FP32_TOL
    assert (full[:32].cpu().argmax(1) == ref.argmax(1)).all()


@pytest.mark.parametrize("head,prior", VARIANTS)
def test_backward_tiny_matches_reference_gradients(cuda_device, head, prior):
    blob = load_npz(f"xattn_tiny_{_tag(head, prior)}.npz")
    m = tiny_head(blob, head, prior).to(cuda_device).eval()   # eval: dropouts identity, autograd on
    out = m(torch.from_numpy(blob["xv"]).to(cuda_device), torch.from_numpy(blob["xa"]).to(cuda_device))
    loss = torch.nn.functional.cross_entropy(out, torch.from_numpy(blob["labels"]).to(cuda_device))
    assert abs(float(loss) - float(blob["loss"])) < 1e-5
    loss.backward()
    named = dict(m.named_parameters())
    n = 0
    for k, v in blob.items():
        if k.startswith("g:"):
            g = named[k[2:]].grad
            assert g is not None, k
            np.testing.assert_allclose(g.cpu().numpy(), v, atol=2e-6, rtol=1e-4, err_msg=k)
            n += 1
    assert n >= 20
    assert named["audio_time_conv.weight"].grad is None


@pytest.mark.parametrize("head,prior", VARIANTS)
def test_backward_config1_matches_port_autograd_and_reference_digests(cuda_device, head, prior):
    blob = load_npz("xattn_config1.npz")
    tag = _tag(head, prior)
    xv, xa, labels = config1_inputs()
    m = make_head(head, prior, seed=0).to(cuda_device).eval()
    out = m(xv.to(cuda_device), xa.to(cuda_device))
    loss = torch.nn.functional.cross_entropy(out, labels.to(cuda_device))
    loss.backward()
    assert abs(float(loss) - float(blob[f"{tag}:loss"])) < 1e-5
    named = dict(m.named_parameters())
    # (1) digests of the reference's own gradients
    for k, d in zip(blob[f"{tag}:gkeys"], blob[f"{tag}:gdigest"]):
        np.testing.assert_allclose(digest(named[str(k)].grad), d, atol=5e-6, rtol=2e-4, err_msg=str(k))
    # (2) element-wise against autograd through the CPU port
    sdt = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    ref = torch_port.xattn_forward(sdt, xv, xa, num_heads=4, head=head, use_prior=prior)
    torch.nn.functional.cross_entropy(ref, labels).backward()
    for k, p in named.items():
        if sdt[k].grad is None:
            assert p.grad is None, k
            continue
        np.testing.assert_allclose(p.grad.cpu().numpy(), sdt[k].grad.numpy(), atol=3e-6, rtol=2e-4, err_msg=k)


def test_ce_loss_kernel_vs_oracle(cuda_device):
    import ctypes as C

    from multimodalemotionrecognition_b200 import _lib
    from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr

    g = torch.Generator().manual_seed(9)
    for B, Cn, eps in [(1, 8, 0.0), (37, 8, 0.1), (1024, 8, 0.05), (5, 3, 0.0)]:
        logits = torch.randn(B, Cn, generator=g) * 3
        labels = torch.randint(0, Cn, (B,), generator=g)
        ref_loss, ref_d = xattn_oracle.cross_entropy(logits.numpy(), labels.numpy(), eps)
        lg, lb = logits.to(cuda_device), labels.to(cuda_device)
        loss = torch.empty((), device=cuda_device)
        d = torch.empty_like(lg)
        _lib.check(_lib.load().xattn_ce_loss_fwd_bwd(_ptr(lg), _ptr(lb), B, Cn, eps, 1.0, _ptr(loss), _ptr(d), _stream_ptr(cuda_device)))
        assert abs(float(loss) - ref_loss) < 1e-5
        np.testing.assert_allclose(d.cpu().numpy(), ref_d, atol=1e-7, rtol=1e-5)
        tl = torch.nn.functional.cross_entropy(logits, labels, label_smoothing=eps)
        assert abs(float(loss) - float(tl)) < 1e-5


@pytest.mark.parametrize("decoupled", [False, True])
def test_adam_kernel_vs_oracle_and_torch(cuda_device, decoupled):
    from multimodalemotionrecognition_b200 import _lib
    from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr

    n = 10007
    g = torch.Generator().manual_seed(4)
    p0 = torch.randn(n, generator=g)
    p = p0.clone().to(cuda_device)
    m = torch.zeros(n, device=cuda_device)
    v = torch.zeros(n, device=cuda_device)
    pt = torch.nn.Parameter(p0.clone())
    opt = (torch.optim.AdamW if decoupled else torch.optim.Adam)([pt], lr=1e-3, weight_decay=1e-2)
    po, mo, vo = p0.double().numpy(), np.zeros(n), np.zeros(n)
    for step in range(1, 6):
        grad = torch.randn(n, generator=g)
        gd = grad.to(cuda_device)
        _lib.check(_lib.load().xattn_adam_step(_ptr(p), _ptr(gd), _ptr(m), _ptr(v), n, 1e-3, 0.9, 0.999, 1e-8, 1e-2, step,
                                               int(decoupled), _stream_ptr(cuda_device)))
        pt.grad = grad.clone()
        opt.step()
        po, mo, vo = xattn_oracle.adam_step(po, grad.numpy(), mo, vo, step=step, lr=1e-3, weight_decay=1e-2, decoupled=decoupled)
    np.testing.assert_allclose(p.cpu().numpy(), po, atol=2e-6, rtol=0)
    np.testing.assert_allclose(p.cpu().numpy(), pt.detach().numpy(), atol=2e-6, rtol=0)


@pytest.mark.parametrize("head,prior", [("concat", True), ("gated", False)])
def test_stage1_step_matches_reference_adam_step(cuda_device, head, prior):
    """Fused step (fwd + CE + bwd + Adam, eval-mode dropouts) against the reference's post-step parameters."""
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    blob = load_npz(f"xattn_tiny_{_tag(head, prior)}.npz")
    m = tiny_head(blob, head, prior).to(cuda_device)
    tr = Stage1Trainer(m, lr=1e-3, weight_decay=1e-4, decoupled=False)
    loss = tr.step(torch.from_numpy(blob["xv"]).to(cuda_device), torch.from_numpy(blob["xa"]).to(cuda_device),
                   torch.from_numpy(blob["labels"]).to(cuda_device), training=False)
    assert abs(float(loss) - float(blob["loss"])) < 1e-5
    sd = m.state_dict()
    for k, v in blob.items():
        if k.startswith("a:") and ("g:" + k[2:]) in blob:
            got, g, w0 = sd[k[2:]].cpu().numpy(), blob["g:" + k[2:]], blob["w:" + k[2:]]
            # Adam's first step is lr * g/(|g|+eps): where the true gradient is ~0 (e.g. the key bias of an attention,
            # to which softmax is invariant) the update is the sign of rounding noise, in torch as much as here —
            # compare those entries only by the |update| <= lr bound.
            solid = np.abs(g) > 1e-5
            np.testing.assert_allclose(got[solid], v[solid], atol=3e-6, rtol=0, err_msg=k)
            assert np.all(np.abs(got - w0) <= 1e-3 * 1.001 + 1e-7), k
    # audio_time_conv receives no gradient -> untouched (no weight decay either), as in torch
    np.testing.assert_array_equal(sd["audio_time_conv.weight"].cpu().numpy(), blob["w:audio_time_conv.weight"])


def test_training_mode_dropout_statistics(cuda_device):
    """Training-mode RNG cannot match torch bit-for-bit; check determinism under a fixed seed/offset,
    that masks differ between steps, and that the expectation stays close to the eval output."""
    m = make_head("concat", True, seed=2).to(cuda_device)
    xv, xa, _ = config1_inputs()
    xv, xa = xv.to(cuda_device).repeat(64, 1, 1), xa.to(cuda_device).repeat(64, 1, 1)
    m.eval()
    with torch.no_grad():
        ref = m(xv, xa)
    m.train()
    outs = []
    with torch.no_grad():
        for _ in range(3):
            logits, _ = m._launch_forward(xv, xa, save=False, training=True)
            outs.append(logits)
    assert not torch.equal(outs[0], outs[1])
    assert torch.isfinite(outs[0]).all()
    # same clip repeated 64x with independent masks: the mean over repeats approaches the eval logits
    mean_train = torch.stack(outs).mean(0).view(64, 8, 8).mean(0)
    assert (mean_train - ref[:8]).abs().max() < 0.1
    m._rng_offset = 0
    a, _ = m._launch_forward(xv, xa, save=False, training=True)
    m._rng_offset = 0
    b, _ = m._launch_forward(xv, xa, save=False, training=True)
    assert torch.equal(a, b)


def test_fusion_model_drop_in_with_stub_backbones(cuda_device):
    """FusionModel keeps the reference's forward(video, audio) contract (fusion.py:356-411)."""
    from multimodalemotionrecognition_b200 import FusionModel

    class StubVideo(torch.nn.Module):
        embedding_dim = 512

        def __init__(self):
            super().__init__()
            self.backbone = torch.nn.Identity()

    class StubAudio(torch.nn.Module):
        sequence_dim = 768
        embedding_dim = 768
        called = False

        def encode_sequence(self, x):
            self.called = True
            return x

    torch.manual_seed(0)
    audio = StubAudio()
    fm = FusionModel(audio, StubVideo(), 8, mode="xattn", xattn_head="concat", d_model=128, num_heads=4, audio_n_mels=768,
                     xattn_use_emotion_prior=True).to(cuda_device).eval()
    blob = load_npz("xattn_config1.npz")
    xv, xa, _ = config1_inputs()
    with torch.no_grad():
        out = fm(xv.to(cuda_device)[..., None, None], xa.to(cuda_device))
    assert audio.called
    assert np.abs(out.cpu().numpy() - blob["concat_prior:logits"]).max() <= FP32_TOL
    assert fm.pop_alignment_loss() is None
    assert sorted(k for k in fm.state_dict()) == list(blob["concat_prior:keys"])


@pytest.mark.parametrize("mode", ["concat", "gated"])
def test_flat_heads_match_reference(cuda_device, mode):
    from multimodalemotionrecognition_b200 import FlatFusionHead

    blob = load_npz(f"flat_tiny_{mode}.npz")
    m = FlatFusionHead(num_classes=8, video_dim=12, audio_dim=10, common_dim=16, mode=mode, fusion_align_mode="clip",
                       fusion_align_dim=256)
    m.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("w:")}, strict=True)
    m = m.to(cuda_device).eval()
    with torch.no_grad():
        out = m(torch.from_numpy(blob["xv"]).to(cuda_device), torch.from_numpy(blob["xa"]).to(cuda_device))
    np.testing.assert_allclose(out.cpu().numpy(), blob["logits"], atol=1e-5, rtol=0)
    # config-5 dimensions, seed-0 init identical to the reference (digests), pooled inputs
    big = load_npz("flat_config5.npz")
    torch.manual_seed(0)
    mm = FlatFusionHead(num_classes=8, video_dim=512, audio_dim=768, common_dim=256, mode=mode).to(cuda_device).eval()
    g = torch.Generator().manual_seed(3)
    ap = torch.randn(6, 768, generator=g)
    vp = torch.randn(6, 512, generator=g)
    with torch.no_grad():
        out = mm.flat_forward_pooled(ap.to(cuda_device), vp.to(cuda_device))
    assert np.abs(out.cpu().numpy() - big[f"{mode}:logits"]).max() <= FP32_TOL
    # bf16 sequence inputs with the fused temporal mean vs the oracle on the same rounded inputs
    xa = torch.randn(9, 149, 768, generator=g).bfloat16()
    xv = torch.randn(9, 16, 512, generator=g).bfloat16()
    mm._flat_precision = "fp32"
    with torch.no_grad():
        out = mm(xv.to(cuda_device), xa.to(cuda_device)).cpu().numpy()
    sd = {k: v.cpu().numpy() for k, v in mm.state_dict().items()}
    ref = xattn_oracle.flat_head_forward(sd, xa.float().numpy().mean(1), xv.float().numpy().mean(1), mode=mode)
    assert np.abs(out - ref).max() <= FP32_TOL
    # the same on the tensor-core path (precision="auto" picks it for bf16 embeddings): bf16 operands, fp32 accumulate
    mm._flat_precision = "auto"
    n0 = _lib.launch_count()
    with torch.no_grad():
        out_tc = mm(xv.to(cuda_device), xa.to(cuda_device)).cpu().numpy()
    assert _lib.launch_count() - n0 == 9     # cvt + 2 means + 5 tcgen05 GEMMs + output kernel
    assert np.abs(out_tc - ref).max() <= BF16_TOL
    assert np.abs(out_tc - ref).max() > 0    # (a different code path really ran)


@pytest.mark.parametrize("mode", ["concat", "gated"])
def test_flat_alignment_loss_matches_reference(cuda_device, mode):
    """ClipStyleAlignment loss (fusion.py:141-149), SURVEY 8a row a14: the tiny golden case against the reference's own value
    (fp32 kernels), then config-5 dimensions against the oracle on fp32 SIMT and on the tensor cores (tf32 products), and the
    drop-in FusionModel.pop_alignment_loss() (train.py:221-222)."""
    from multimodalemotionrecognition_b200 import FlatFusionHead, FusionModel

    blob = load_npz(f"flat_tiny_{mode}.npz")
    m = FlatFusionHead(num_classes=8, video_dim=12, audio_dim=10, common_dim=16, mode=mode, fusion_align_mode="clip",
                       fusion_align_dim=256)
    m.load_state_dict({k[2:]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("w:")}, strict=True)
    m = m.to(cuda_device).eval()
    loss = m.alignment_loss(torch.from_numpy(blob["xv"]).to(cuda_device), torch.from_numpy(blob["xa"]).to(cuda_device))
    assert loss.shape == () and abs(float(loss) - float(blob["align_loss"])) < 1e-5
    torch.manual_seed(0)
    mm = FlatFusionHead(num_classes=8, video_dim=512, audio_dim=768, common_dim=256, mode=mode).to(cuda_device).eval()
    with torch.no_grad():
        mm.semantic_alignment.logit_scale.fill_(2.0)
    g = torch.Generator().manual_seed(4)
    B = 256
    xa = torch.randn(B, 5, 768, generator=g)
    xv = torch.randn(B, 3, 512, generator=g)
    sd = {k: v.cpu().numpy() for k, v in mm.state_dict().items()}
    _, ref = xattn_oracle.flat_head_forward(sd, xa.numpy().mean(1), xv.numpy().mean(1), mode=mode, return_align_loss=True)
    mm._flat_precision = "fp32"
    l32 = float(mm.alignment_loss(xv.to(cuda_device), xa.to(cuda_device)))
    assert abs(l32 - ref) < 1e-4, (l32, ref)
    mm._flat_precision = "auto"
    n0 = _lib.launch_count()
    ltc = float(mm.alignment_loss(xv.to(cuda_device), xa.to(cuda_device)))
    assert _lib.launch_count() - n0 == 11     # 2 means, 4 tcgen05 GEMMs, 2 normalisations, 2 row passes, final sum
    assert abs(ltc - ref) < 1e-2, (ltc, ref)
    # drop-in: the loss of the last concat/gated forward, evaluated when popped
    class Enc(torch.nn.Module):
        def __init__(self, dim):
            super().__init__()
            self.embedding_dim = dim

        def encode(self, x):
            return x
    torch.manual_seed(0)
    fm = FusionModel(Enc(768), Enc(512), 8, mode=mode, fusion_align_mode="clip", precision="fp32").to(cuda_device).eval()
    with torch.no_grad():
        fm.semantic_alignment.logit_scale.fill_(2.0)
        fm(xv.mean(1).to(cuda_device), xa.mean(1).to(cuda_device))
    sdf = {k: v.cpu().numpy() for k, v in fm.state_dict().items()}
    _, reff = xattn_oracle.flat_head_forward(sdf, xa.numpy().mean(1), xv.numpy().mean(1), mode=mode, return_align_loss=True)
    popped = fm.pop_alignment_loss()
    assert popped is not None and abs(float(popped) - reff) < 1e-4
    assert fm.pop_alignment_loss() is None


def test_attention_pooling_matches_reference(cuda_device):
    """temporal_pooling="attn" (TemporalAttentionPooling, temporal.py:9-26) in the fp32 kernel family: the tiny
    golden case and the BASELINE config-1 dimensions against the reference's logits (1e-4), also with bf16 embeddings; with those
    the tensor-core path runs too (its a2v kernel emits the normalised audio tokens, the pooling kernel reads them in bf16): 1e-2."""
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    blob = load_npz("xattn_attnpool.npz")
    sd = {k[len("tiny:w:"):]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("tiny:w:")}
    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=8,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                        temporal_pooling="attn")
    m.load_state_dict(sd, strict=True)
    m = m.to(cuda_device).eval()
    with torch.no_grad():
        out = m(torch.from_numpy(blob["tiny:xv"]).to(cuda_device), torch.from_numpy(blob["tiny:xa"]).to(cuda_device))
    assert np.abs(out.cpu().numpy() - blob["tiny:logits"]).max() <= 1e-5
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    for head in ("concat", "gated"):
        mm = make_head(head, True, seed=0, temporal_pooling="attn").to(cuda_device).eval()
        with torch.no_grad():
            out = mm(xv.to(cuda_device), xa.to(cuda_device))
            mm.precision = "fp32"
            outb = mm(xv.bfloat16().to(cuda_device), xa.bfloat16().to(cuda_device))      # fp32 kernels on bf16 embeddings
            mm.precision = "auto"
            mm(xv.bfloat16().to(cuda_device), xa.bfloat16().to(cuda_device))             # (packs the bf16 weights once)
            n0 = _lib.launch_count()
            outtc = mm(xv.bfloat16().to(cuda_device), xa.bfloat16().to(cuda_device))     # tcgen05 path + the pooling kernel
            n_tc = _lib.launch_count() - n0
        assert np.abs(out.cpu().numpy() - blob[f"{head}:logits"]).max() <= FP32_TOL
        sdn = {k: v.cpu().numpy() for k, v in mm.state_dict().items()}
        refb = xattn_oracle.xattn_forward(sdn, xv.bfloat16().float().numpy(), xa.bfloat16().float().numpy(), num_heads=4, head=head)
        assert np.abs(outb.cpu().numpy() - refb).max() <= FP32_TOL
        assert mm._pick_precision(_lib.DTYPE_BF16, False, 16, 149) == _lib.DTYPE_BF16 and n_tc == 11   # 9 launches + 2 poolings
        assert np.abs(outtc.cpu().numpy() - refb).max() <= BF16_TOL
        assert np.abs(outtc.cpu().numpy() - outb.cpu().numpy()).max() > 0


def test_attention_pooling_backward_matches_reference(cuda_device):
    """Backward through TemporalAttentionPooling (attn_pool_bwd_kernel): every gradient of the tiny case against the reference's
    autograd (eval mode, dropouts identity); at the BASELINE config-1 dimensions the pooler's own gradients element-wise and all
    others by digest, in fp32 and (cosine) tf32; training mode is deterministic for a given Philox offset and differs from eval."""
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    blob = load_npz("xattn_attnpool.npz")
    sd = {k[len("tiny:w:"):]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("tiny:w:")}
    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=8,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                        temporal_pooling="attn")
    m.load_state_dict(sd, strict=True)
    m = m.to(cuda_device).eval()
    out = m(torch.from_numpy(blob["tiny:xv"]).to(cuda_device), torch.from_numpy(blob["tiny:xa"]).to(cuda_device))
    loss = torch.nn.functional.cross_entropy(out, torch.from_numpy(blob["tiny:labels"]).to(cuda_device))
    assert abs(float(loss.detach()) - float(blob["tiny:loss"])) < 1e-5
    loss.backward()
    named = dict(m.named_parameters())
    n = 0
    for k, v in blob.items():
        if k.startswith("tiny:g:"):
            g = named[k[len("tiny:g:"):]].grad
            assert g is not None, k
            np.testing.assert_allclose(g.cpu().numpy(), v, atol=2e-6, rtol=1e-4, err_msg=k)
            n += 1
    assert n >= 30 and "tiny:g:v_temporal_pool.pool.score.1.weight" in blob
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g).to(cuda_device)
    xa = torch.randn(8, 149, 768, generator=g).to(cuda_device)
    for head in ("concat", "gated"):
        labels = torch.from_numpy(blob[f"{head}:labels"]).to(cuda_device)
        ref_pool = {}
        for prec in ("fp32", "tf32"):
            mm = make_head(head, True, seed=0, temporal_pooling="attn", precision=prec).to(cuda_device).eval()
            loss = torch.nn.functional.cross_entropy(mm(xv, xa), labels)
            loss.backward()
            named = dict(mm.named_parameters())
            if prec == "fp32":
                assert abs(float(loss.detach()) - float(blob[f"{head}:loss"])) < 1e-5
                for k, v in blob.items():
                    if k.startswith(f"{head}:g:"):
                        np.testing.assert_allclose(named[k[len(head) + 3:]].grad.cpu().numpy(), v, atol=2e-6, rtol=2e-4, err_msg=k)
                    elif k.startswith(f"{head}:gd:"):
                        np.testing.assert_allclose(digest(named[k[len(head) + 4:]].grad), v, atol=5e-5, rtol=2e-4, err_msg=k)
            else:
                for k, v in blob.items():
                    if k.startswith(f"{head}:g:"):
                        got = named[k[len(head) + 3:]].grad.flatten().double().cpu()
                        want = torch.from_numpy(v).flatten().double()
                        if want.numel() > 1:
                            assert float(torch.dot(got, want) / (got.norm() * want.norm() + 1e-30)) >= 0.99, k
    mm = make_head("concat", True, seed=0, temporal_pooling="attn").to(cuda_device).train()
    mm._rng_offset = 7
    a = mm(xv, xa).detach().clone()
    mm._rng_offset = 7
    b = mm(xv, xa).detach().clone()
    assert torch.equal(a, b) and torch.isfinite(a).all()
    mm.eval()
    assert not torch.equal(a, mm(xv, xa).detach())


def test_mel_front_end_matches_reference(cuda_device):
    """audio_time_conv (fusion.py:379-384) as unfold (csrc/mel_frontend.cu) + the head's first audio Linear:
    the unfold is exact; tiny case logits / loss / every gradient / post-Adam parameters against the reference; n_mels=64,
    Ta=149, d_model=128 through FusionModel.forward(video, mel) in fp32, tf32 and bf16, gradients in fp32."""
    from multimodalemotionrecognition_b200 import FusionModel, XAttnFusionHead
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    blob = load_npz("xattn_mel.npz")
    sd = {k[len("tiny:w:"):]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("tiny:w:")}
    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=5,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                        audio_frontend="mel")
    m.load_state_dict(sd, strict=True)
    m = m.to(cuda_device).eval()
    xv = torch.from_numpy(blob["tiny:xv"]).to(cuda_device)
    mel = torch.from_numpy(blob["tiny:mel"]).to(cuda_device)
    labels = torch.from_numpy(blob["tiny:labels"]).to(cuda_device)
    unf = m.mel_unfold(mel)
    pad = torch.nn.functional.pad(mel[:, 0], (1, 1))
    want = torch.stack([pad[:, :, k:k + mel.shape[3]] for k in range(3)], dim=-1).permute(0, 2, 1, 3).reshape(3, mel.shape[3], 15)
    assert torch.equal(unf, want)
    assert torch.equal(m.mel_unfold(mel.bfloat16()), want.bfloat16())
    out = m(xv, mel)
    assert np.abs(out.detach().cpu().numpy() - blob["tiny:logits"]).max() <= 1e-5
    loss = torch.nn.functional.cross_entropy(out, labels)
    assert abs(float(loss) - float(blob["tiny:loss"])) < 1e-5
    loss.backward()
    named = dict(m.named_parameters())
    n = 0
    for k, v in blob.items():
        if k.startswith("tiny:g:"):
            g = named[k[len("tiny:g:"):]].grad
            assert g is not None, k
            np.testing.assert_allclose(g.cpu().numpy(), v, atol=2e-6, rtol=1e-4, err_msg=k)
            n += 1
    assert n >= 20 and "tiny:g:audio_time_conv.weight" in blob
    assert named["audio_seq_proj.weight"].grad is None
    # one Stage-1 step (CE + backward + Adam lr 1e-3 wd 1e-4) against the reference's post-step parameters
    m2 = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=5,
                         xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                         audio_frontend="mel")
    m2.load_state_dict(sd, strict=True)
    m2 = m2.to(cuda_device)
    tr = Stage1Trainer(m2, lr=1e-3, weight_decay=1e-4)
    tr.step(xv, m2.mel_unfold(mel), labels, training=False)
    sd2 = m2.state_dict()
    for k, v in blob.items():
        name = k[len("tiny:after:"):]
        if k.startswith("tiny:after:") and ("tiny:g:" + name) in blob:
            got, gr, w0 = sd2[name].cpu().numpy(), blob["tiny:g:" + name], blob["tiny:w:" + name]
            solid = np.abs(gr) > 1e-5     # ~0 gradients (attention key bias): Adam's first step is the sign of rounding noise
            np.testing.assert_allclose(got[solid], v[solid], atol=3e-6, rtol=0, err_msg=k)
            assert np.all(np.abs(got - w0) <= 1e-3 * 1.001 + 1e-7), k
    np.testing.assert_array_equal(sd2["audio_seq_proj.weight"].cpu().numpy(), blob["tiny:w:audio_seq_proj.weight"])

    class StubVideo(torch.nn.Module):
        embedding_dim = 512

        def __init__(self):
            super().__init__()
            self.backbone = torch.nn.Identity()

    class StubAudioMel(torch.nn.Module):
        embedding_dim = 128

    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g).to(cuda_device)
    mel = torch.randn(8, 1, 64, 149, generator=g).to(cuda_device)
    labels = torch.randint(0, 8, (8,), generator=g).to(cuda_device)
    for head in ("concat", "gated"):
        torch.manual_seed(0)
        fm = FusionModel(StubAudioMel(), StubVideo(), 8, mode="xattn", xattn_head=head, d_model=128, num_heads=4, audio_n_mels=64,
                         xattn_use_emotion_prior=True).to(cuda_device).eval()
        out = fm(xv[..., None, None], mel)
        assert np.abs(out.detach().cpu().numpy() - blob[f"{head}:logits"]).max() <= FP32_TOL
        loss = torch.nn.functional.cross_entropy(out, labels)
        assert abs(float(loss) - float(blob[f"{head}:loss"])) < 1e-5
        loss.backward()
        for name in ("audio_time_conv.weight", "audio_time_conv.bias"):
            np.testing.assert_allclose(dict(fm.named_parameters())[name].grad.cpu().numpy(), blob[f"{head}:g:{name}"],
                                       atol=2e-6, rtol=1e-4, err_msg=name)
        for k, v in blob.items():
            if k.startswith(f"{head}:gd:"):
                gr = dict(fm.named_parameters())[k[len(head) + 4:]].grad
                np.testing.assert_allclose(digest(gr), v, atol=5e-5, rtol=1e-4, err_msg=k)
        with torch.no_grad():
            fm.precision = "tf32"
            assert np.abs(fm(xv[..., None, None], mel).cpu().numpy() - blob[f"{head}:logits"]).max() <= BF16_TOL
            fm.precision = "bf16"
            outb = fm(xv.bfloat16()[..., None, None], mel.bfloat16())
            assert np.abs(outb.cpu().numpy() - blob[f"{head}:logits"]).max() <= 2 * BF16_TOL
            fm.precision = "auto"


def test_transformer_pooling_matches_reference(cuda_device):
    """temporal_pooling="transformer" (TemporalTransformerPooling, temporal.py:46-75), forward: the tiny case
    (generic attention kernel, 2 pooler heads) and the BASELINE config-1 dimensions against the reference's logits in fp32 (1e-4),
    tf32 and with bf16 embeddings; two encoder layers (temporal_num_layers=2) on the second tiny case."""
    import warnings

    from multimodalemotionrecognition_b200 import XAttnFusionHead

    warnings.simplefilter("ignore")
    blob = load_npz("xattn_txpool.npz")
    sd = {k[len("tiny:w:"):]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("tiny:w:")}
    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=8,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                        temporal_pooling="transformer", temporal_num_heads=2)
    m.load_state_dict(sd, strict=True)
    m = m.to(cuda_device).eval()
    with torch.no_grad():
        out = m(torch.from_numpy(blob["tiny:xv"]).to(cuda_device), torch.from_numpy(blob["tiny:xa"]).to(cuda_device))
    assert np.abs(out.cpu().numpy() - blob["tiny:logits"]).max() <= 1e-5
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g)
    xa = torch.randn(8, 149, 768, generator=g)
    for head in ("concat", "gated"):
        mm = make_head(head, True, seed=0, temporal_pooling="transformer").to(cuda_device).eval()
        with torch.no_grad():
            out = mm(xv.to(cuda_device), xa.to(cuda_device))
            outb = mm(xv.bfloat16().to(cuda_device), xa.bfloat16().to(cuda_device))
            mm.precision = "tf32"
            outt = mm(xv.to(cuda_device), xa.to(cuda_device))
            mm.precision = "auto"
        assert np.abs(out.cpu().numpy() - blob[f"{head}:logits"]).max() <= FP32_TOL
        assert np.abs(outt.cpu().numpy() - blob[f"{head}:logits"]).max() <= BF16_TOL
        sdn = {k: v.cpu().numpy() for k, v in mm.state_dict().items()}
        refb = xattn_oracle.xattn_forward(sdn, xv.bfloat16().float().numpy(), xa.bfloat16().float().numpy(), num_heads=4, head=head)
        assert np.abs(outb.cpu().numpy() - refb).max() <= FP32_TOL
    sd2 = {k[len("tiny2:w:"):]: torch.from_numpy(v) for k, v in blob.items() if k.startswith("tiny2:w:")}
    m2 = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=8,
                         xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                         temporal_pooling="transformer", temporal_num_heads=2, temporal_num_layers=2)
    m2.load_state_dict(sd2, strict=True)
    m2 = m2.to(cuda_device).eval()
    with torch.no_grad():
        out2 = m2(torch.from_numpy(blob["tiny2:xv"]).to(cuda_device), torch.from_numpy(blob["tiny2:xa"]).to(cuda_device))
    assert np.abs(out2.cpu().numpy() - blob["tiny2:logits"]).max() <= 1e-5
    ref2 = xattn_oracle.xattn_forward({k: v.numpy() for k, v in sd2.items()}, blob["tiny2:xv"], blob["tiny2:xa"], num_heads=2,
                                      temporal_num_heads=2)
    assert np.abs(ref2 - blob["tiny2:logits"]).max() <= 1e-5


def _tiny_tx_head(blob, tag, layers, cuda_device, **kw):
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    sd = {k[len(tag) + 3:]: torch.from_numpy(v) for k, v in blob.items() if k.startswith(f"{tag}:w:")}
    m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=8,
                        xattn_use_emotion_prior=True, xattn_emotion_prior_dim=4, xattn_emotion_prior_hidden_dim=8,
                        temporal_pooling="transformer", temporal_num_heads=2, temporal_num_layers=layers, **kw)
    m.load_state_dict(sd, strict=True)
    return m.to(cuda_device)


def test_transformer_pooling_backward_matches_reference(cuda_device):
    """Backward through TemporalTransformerPooling (temporal.py:46-75: encoder layers + learned pooling): every gradient and the
    post-Adam parameters of the tiny cases (one and TWO encoder layers) against the reference's autograd (eval mode, dropouts
    identity); at the BASELINE config-1 dimensions (audio self-attention 149 x 149: the row-tiled backward kernel) the loss, the
    pooler's small gradients element-wise and every gradient by digest in fp32, cosine in tf32."""
    import warnings

    from multimodalemotionrecognition_b200.training import Stage1Trainer

    warnings.simplefilter("ignore")
    blob = load_npz("xattn_txpool.npz")
    for tag, layers in (("tiny", 1), ("tiny2", 2)):
        m = _tiny_tx_head(blob, tag, layers, cuda_device).eval()
        xv = torch.from_numpy(blob[f"{tag}:xv"]).to(cuda_device)
        xa = torch.from_numpy(blob[f"{tag}:xa"]).to(cuda_device)
        labels = torch.from_numpy(blob[f"{tag}:labels"]).to(cuda_device)
        loss = torch.nn.functional.cross_entropy(m(xv, xa), labels)
        assert abs(float(loss.detach()) - float(blob[f"{tag}:loss"])) < 1e-5
        loss.backward()
        named = dict(m.named_parameters())
        n = 0
        for k, v in blob.items():
            if k.startswith(f"{tag}:g:"):
                g = named[k[len(tag) + 3:]].grad
                assert g is not None, k
                np.testing.assert_allclose(g.cpu().numpy(), v, atol=2e-6, rtol=1e-4, err_msg=k)
                n += 1
        assert n >= 30 + 24 * layers and f"{tag}:g:a_temporal_pool.pool.encoder.layers.{layers - 1}.linear1.weight" in blob
        # one Stage-1 step (CE + backward + Adam lr 1e-3 wd 1e-4) against the reference's post-step parameters
        m2 = _tiny_tx_head(blob, tag, layers, cuda_device)
        Stage1Trainer(m2, lr=1e-3, weight_decay=1e-4).step(xv, xa, labels, training=False)
        sd2 = m2.state_dict()
        for k, v in blob.items():
            name = k[len(tag) + 7:]
            if k.startswith(f"{tag}:after:") and (f"{tag}:g:" + name) in blob:
                got, gr, w0 = sd2[name].cpu().numpy(), blob[f"{tag}:g:" + name], blob[f"{tag}:w:" + name]
                solid = np.abs(gr) > 1e-5     # ~0 gradients (attention key bias): Adam's first step is the sign of rounding noise
                np.testing.assert_allclose(got[solid], v[solid], atol=3e-6, rtol=0, err_msg=k)
                assert np.all(np.abs(got - w0) <= 1e-3 * 1.001 + 1e-7), k
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(8, 16, 512, generator=g).to(cuda_device)
    xa = torch.randn(8, 149, 768, generator=g).to(cuda_device)
    for head in ("concat", "gated"):
        labels = torch.from_numpy(blob[f"{head}:labels"]).to(cuda_device)
        for prec in ("fp32", "tf32"):
            mm = make_head(head, True, seed=0, temporal_pooling="transformer", precision=prec).to(cuda_device).eval()
            loss = torch.nn.functional.cross_entropy(mm(xv, xa), labels)
            loss.backward()
            named = dict(mm.named_parameters())
            n = 0
            for k, v in blob.items():
                if prec == "fp32" and k.startswith(f"{head}:g:"):
                    np.testing.assert_allclose(named[k[len(head) + 3:]].grad.cpu().numpy(), v, atol=2e-6, rtol=2e-4, err_msg=k)
                    n += 1
                elif prec == "fp32" and k.startswith(f"{head}:gd:"):
                    np.testing.assert_allclose(digest(named[k[len(head) + 4:]].grad), v, atol=5e-5, rtol=2e-4, err_msg=k)
                    n += 1
                elif prec == "tf32" and k.startswith(f"{head}:g:") and v.size > 1:
                    got = named[k[len(head) + 3:]].grad.flatten().double().cpu()
                    want = torch.from_numpy(v).flatten().double()
                    if float(want.norm()) > 1e-7:   # (the attention key bias has a zero gradient)
                        assert float(torch.dot(got, want) / (got.norm() * want.norm() + 1e-30)) >= 0.99, k
                    n += 1
            assert n >= 26
            if prec == "fp32":
                assert abs(float(loss.detach()) - float(blob[f"{head}:loss"])) < 1e-5


def test_transformer_pooling_training_mode_gradients_are_consistent(cuda_device):
    """Training mode of the transformer pooler (four dropouts per encoder layer + the pooler's, temporal.py:60-68, :19): for a fixed
    Philox offset the loss is a deterministic function of the parameters, so the analytic gradient must match central differences
    along random directions — this ties every backward mask to its forward mask.  With temporal_dropout=0 (and the other dropouts
    off) training mode reproduces the eval-mode gradients exactly."""
    import warnings

    warnings.simplefilter("ignore")
    blob = load_npz("xattn_txpool.npz")
    xv = torch.from_numpy(blob["tiny2:xv"]).to(cuda_device)
    xa = torch.from_numpy(blob["tiny2:xa"]).to(cuda_device)
    labels = torch.from_numpy(blob["tiny2:labels"]).to(cuda_device)
    off = dict(xattn_attn_dropout=0.0, xattn_stochastic_depth=0.0, xattn_emotion_prior_dropout=0.0)
    m = _tiny_tx_head(blob, "tiny2", 2, cuda_device, temporal_dropout=0.25, **off).train()
    m.xattn_head_dropout = 0.0

    def loss_at():
        m._rng_offset = 40          # the forward advances it by one: every evaluation draws the masks of offset 41
        return torch.nn.functional.cross_entropy(m(xv, xa), labels)

    base = loss_at()
    base.backward()
    m.eval()
    with torch.no_grad():
        assert abs(float(torch.nn.functional.cross_entropy(m(xv, xa), labels)) - float(base)) > 1e-4   # the masks are active
    m.train()
    params = [(n, p) for n, p in m.named_parameters() if p.grad is not None and "temporal_pool" in n]
    assert len(params) >= 2 * (24 + 6)
    grads = {n: p.grad.detach().clone() for n, p in params}
    gen = torch.Generator().manual_seed(5)
    eps = 2e-2
    for trial in range(6):
        dirs = {n: torch.randn(p.shape, generator=gen).to(cuda_device) for n, p in params}
        scale = 1.0 / float(sum(float(d.norm()) ** 2 for d in dirs.values())) ** 0.5
        want = sum(float((grads[n] * dirs[n]).sum()) for n, _ in params) * scale
        vals = []
        for sign in (+1.0, -1.0):
            with torch.no_grad():
                for n, p in params:
                    p.add_(dirs[n], alpha=sign * eps * scale)
                vals.append(float(loss_at()))
                for n, p in params:
                    p.add_(dirs[n], alpha=-sign * eps * scale)
        got = (vals[0] - vals[1]) / (2 * eps)
        assert abs(got - want) <= 0.03 * abs(want) + 2e-4, (trial, got, want)
    # dropout-free training mode == eval mode
    m0 = _tiny_tx_head(blob, "tiny2", 2, cuda_device, temporal_dropout=0.0, **off).train()
    m0.xattn_head_dropout = 0.0
    torch.nn.functional.cross_entropy(m0(xv, xa), labels).backward()
    named = dict(m0.named_parameters())
    for k, v in blob.items():
        if k.startswith("tiny2:g:"):
            np.testing.assert_allclose(named[k[len("tiny2:g:"):]].grad.cpu().numpy(), v, atol=2e-6, rtol=1e-4, err_msg=k)


def test_mel_unfold_and_serving_epilogue_edge_cases(cuda_device):
    """Ragged / empty inputs of the two small boundary kernels: mel_unfold at Ta = 1, 31, 33, 149 and n_mels = 1, 5, 64 (exact against
    pad + stack), an empty batch; softmax_top1 on an empty batch and a single class; predict_host(top1=True) on an empty batch."""
    from multimodalemotionrecognition_b200 import XAttnFusionHead, softmax_top1

    g = torch.Generator().manual_seed(3)
    for n_mels, Ta, B in [(1, 1, 2), (5, 31, 3), (5, 33, 1), (64, 149, 4), (64, 1, 1)]:
        m = XAttnFusionHead(num_classes=8, video_dim=16, audio_dim=8, common_dim=16, d_model=8, num_heads=2, audio_n_mels=n_mels,
                            audio_frontend="mel").to(cuda_device).eval()
        mel = torch.randn(B, 1, n_mels, Ta, generator=g).to(cuda_device)
        pad = torch.nn.functional.pad(mel[:, 0], (1, 1))
        want = torch.stack([pad[:, :, k:k + Ta] for k in range(3)], dim=-1).permute(0, 2, 1, 3).reshape(B, Ta, 3 * n_mels)
        assert torch.equal(m.mel_unfold(mel), want), (n_mels, Ta)
        assert torch.equal(m.mel_unfold(mel[:, 0].bfloat16()), want.bfloat16())
        with torch.no_grad():
            assert m(torch.randn(B, 4, 16, generator=g).to(cuda_device), mel).shape == (B, 8)
    assert m.mel_unfold(mel[:0]).shape == (0, 1, 192)
    with pytest.raises(ValueError):
        m.mel_unfold(torch.zeros(2, 63, 5, device=cuda_device))
    p, t1, t1p = softmax_top1(torch.empty(0, 8, device=cuda_device))
    assert p.shape == (0, 8) and t1.shape == (0,) and t1p.shape == (0,)
    p, t1, t1p = softmax_top1(torch.tensor([[3.0], [-2.0]], device=cuda_device))
    assert torch.equal(p, torch.ones(2, 1, device=cuda_device)) and t1.tolist() == [0, 0] and t1p.tolist() == [1.0, 1.0]
    with pytest.raises(ValueError):
        softmax_top1(torch.zeros(4, 8))
    head = make_head("concat", True, seed=0).to(cuda_device).eval()
    p, t1, t1p = head.predict_host(torch.empty(0, 16, 512, dtype=torch.bfloat16), torch.empty(0, 149, 768, dtype=torch.bfloat16), top1=True)
    assert p.shape == (0, 8) and t1.shape == (0,) and t1p.shape == (0,)


@pytest.mark.parametrize("io", ["fp32", "bf16"])
def test_input_gradients_reach_trainable_backbones(cuda_device, io):
    """Stage 2 of the reference unfreezes backbone layers (train.py:815-830): loss.backward() at :227 must reach v_feat /
    a_seq.  The head returns dL/d(video_emb), dL/d(audio_emb) (xattn_backward_ex) — checked against autograd of the CPU port —
    while the parameter gradients stay what they are without it."""
    from oracle import torch_port

    m = make_head("concat", True, seed=0, precision="fp32").to(cuda_device).eval()
    xv, xa, labels = config1_inputs()
    if io == "bf16":
        xv, xa = xv.bfloat16(), xa.bfloat16()
    xvd = xv.to(cuda_device).requires_grad_(True)
    xad = xa.to(cuda_device).requires_grad_(True)
    torch.nn.functional.cross_entropy(m(xvd, xad), labels.to(cuda_device)).backward()
    sd = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    rv, ra = xv.float().clone().requires_grad_(True), xa.float().clone().requires_grad_(True)
    torch.nn.functional.cross_entropy(torch_port.xattn_forward(sd, rv, ra, num_heads=4), labels).backward()
    assert xvd.grad is not None and xvd.grad.dtype == xv.dtype and xvd.grad.shape == xv.shape
    tol = 1e-6 if io == "fp32" else 1e-2   # bf16 embeddings receive bf16-rounded gradients (relative to the largest entry)
    for got, ref in ((xvd.grad, rv.grad), (xad.grad, ra.grad)):
        scale = ref.abs().max().item()
        assert (got.float().cpu() - ref).abs().max().item() <= tol * scale + 1e-9, (io, scale)
    gerr = max(float((p.grad.cpu() - sd[k].grad).abs().max()) for k, p in m.named_parameters() if p.grad is not None)
    assert gerr <= 1e-5, gerr
    # frozen head, trainable embeddings: still differentiable w.r.t. the inputs
    for p in m.parameters():
        p.requires_grad_(False)
    xv2 = xv.detach().to(cuda_device).requires_grad_(True)
    torch.nn.functional.cross_entropy(m(xv2, xa.detach().to(cuda_device)), labels.to(cuda_device)).backward()
    assert (xv2.grad.float().cpu() - rv.grad).abs().max().item() <= tol * rv.grad.abs().max().item() + 1e-9


def test_frozen_embedding_cache_emits_bf16_from_the_backbone_tail(cuda_device):
    """SURVEY 8f rank 1 on the GPU: the cache fills its bf16 table through xattn_emit_embeddings_bf16 (spatial mean of the
    ResNet-style feature map + bf16 rounding + scatter by dataset index, csrc/embed_ops.cu), the backbones run once per
    sample, and a cached Stage-1 step equals the step on embeddings computed the reference's way (fusion.py:368-370, :377)."""
    from multimodalemotionrecognition_b200 import FusionModel
    from multimodalemotionrecognition_b200.training import FrozenEmbeddingCache, Stage1Trainer

    class Video(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.embedding_dim = 512
            self.backbone = torch.nn.Sequential(torch.nn.Conv2d(3, 512, 3, stride=2), torch.nn.ReLU(), torch.nn.AdaptiveAvgPool2d((1, 1)))
            self.calls = 0

        def encode(self, x):
            raise NotImplementedError

    class Audio(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.sequence_dim = self.embedding_dim = 768
            self.proj = torch.nn.Linear(40, 768)
            self.calls = 0

        def encode_sequence(self, x):          # [B, Ta, 40] "waveform frames" -> [B, Ta, 768]
            self.calls += 1
            return self.proj(x)

    torch.manual_seed(0)
    model = FusionModel(Audio(), Video(), 8, mode="xattn", audio_n_mels=768, xattn_use_emotion_prior=True, precision="tf32").to(cuda_device)
    N, B, T, Ta = 24, 8, 16, 149
    g = torch.Generator().manual_seed(4)
    video = torch.randn(N, T, 3, 9, 9, generator=g).to(cuda_device)
    audio = torch.randn(N, Ta, 40, generator=g).to(cuda_device)
    labels = torch.randint(0, 8, (N,), generator=g).to(cuda_device)
    cache = FrozenEmbeddingCache(N, T, Ta, device=cuda_device)
    idx = torch.tensor([3, 20, 7, 11, 0, 5, 23, 9])
    v_emb, a_emb = cache.fetch_or_compute(model, idx, video[idx], audio[idx])
    assert cache.misses == B and cache.hits == 0 and model.audio_model.calls == 1
    with torch.no_grad():   # the reference's way: full backbone (with its pooling), fp32, then the cast
        v_ref = model.video_model.backbone(video[idx].view(B * T, 3, 9, 9)).view(B, T, 512)
        a_ref = model.audio_model.proj(audio[idx])
    assert v_emb.dtype == torch.bfloat16 and v_emb.shape == (B, T, 512) and a_emb.shape == (B, Ta, 768)
    assert torch.equal(a_emb, a_ref.bfloat16())                               # a pure rounding: bit-identical
    assert (v_emb.float() - v_ref).abs().max().item() <= 2.0 ** -8 * max(1.0, v_ref.abs().max().item())   # mean order + rounding
    assert torch.equal(cache.video[idx], v_emb) and bool(cache.valid[idx].all()) and int(cache.valid.sum()) == B
    again_v, again_a = cache.fetch_or_compute(model, idx, video[idx], audio[idx])
    assert cache.hits == B and model.audio_model.calls == 1                   # second epoch: no backbone call
    assert torch.equal(again_v, v_emb) and torch.equal(again_a, a_emb)
    # a cached step == a step on the same embeddings
    model.train()
    tr = Stage1Trainer(model, lr=1e-3)
    loss_c = float(tr.step_cached(cache, idx, video[idx], audio[idx], labels[idx], training=False))
    torch.manual_seed(0)
    model2 = FusionModel(Audio(), Video(), 8, mode="xattn", audio_n_mels=768, xattn_use_emotion_prior=True, precision="tf32").to(cuda_device).train()
    loss_d = float(Stage1Trainer(model2, lr=1e-3).step(v_emb, a_emb, labels[idx], training=False))
    assert abs(loss_c - loss_d) <= 1e-6
