"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol include/xattn_fusion.h declares,
host-only layout queries work without a GPU, the module mirrors the reference's state_dict, and there is no CPU
fallback.  No compute entry point is called here."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest
import torch

from helpers import load_npz, make_head, sha

ROOT = Path(__file__).resolve().parents[1]


def _declared_symbols(header="xattn_fusion.h"):
    text = (ROOT / "include" / header).read_text()
    return sorted(set(re.findall(r"XATTN_API\s+[\w\s\*]+?\b((?:xattn|flat_head)_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from multimodalemotionrecognition_b200 import _lib

    lib = _lib.load()
    declared = _declared_symbols()
    assert len(declared) >= 13, declared
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/xattn_fusion.h but not exported"
        assert name in _lib.EXPORTED_SYMBOLS, f"{name} has no ctypes prototype"
    assert lib.xattn_abi_version() == _lib.ABI_VERSION


def test_debug_entry_points_live_in_the_test_library_only():
    from multimodalemotionrecognition_b200 import _lib

    lib, dbg = _lib.load(), _lib.load_debug()
    declared = _declared_symbols("xattn_fusion_debug.h")
    assert sorted(declared) == sorted(_lib.DEBUG_SYMBOLS), declared
    for name in declared:
        assert hasattr(dbg, name), f"{name} declared in include/xattn_fusion_debug.h but not exported by the test library"
        assert not hasattr(lib, name), f"{name} is test scaffolding and must not ship in the product library"
    for name in _declared_symbols():   # the test library is the product plus the debug entry points
        assert hasattr(dbg, name), name


def test_param_layout_matches_state_dict_shapes():
    from multimodalemotionrecognition_b200 import _lib

    for head in ("concat", "gated"):
        for prior in (True, False):
            m = make_head(head, prior)
            cfg = m._make_cfg(8, 16, 149, _lib.DTYPE_F32, _lib.DTYPE_F32, False, False)
            offs, sizes, total = _lib.xattn_param_layout(cfg)
            sd = m.state_dict()
            used = 0
            for name, off, n in zip(_lib.xattn_param_names(head), offs, sizes):
                assert off % 4 == 0
                if n == 0:
                    assert name not in sd or not prior or head == "concat"
                    continue
                assert sd[name].numel() == n, name
                used += n
            assert used == (398581 if prior and head == "concat" else used)
            assert total >= used
            # every fusion parameter except the unused audio_time_conv is in the flat buffer
            in_flat = {n for n, s in zip(_lib.xattn_param_names(head), sizes) if s}
            assert set(sd) - in_flat == {"audio_time_conv.weight", "audio_time_conv.bias"}


def test_workspace_queries_and_config_validation():
    from multimodalemotionrecognition_b200 import _lib

    m = make_head("concat", True)
    small = _lib.workspace_bytes(m._make_cfg(8, 16, 149, 0, 0, False, False))
    train = _lib.workspace_bytes(m._make_cfg(8, 16, 149, 0, 0, True, True))
    assert train > small > 0
    assert _lib.workspace_bytes(m._make_cfg(16, 16, 149, 0, 0, False, False)) > small
    bad = m._make_cfg(8, 16, 149, 0, 0, False, False)
    bad.H = 3   # 128 % 3 != 0 -> same complaint nn.MultiheadAttention raises at construction
    with pytest.raises(ValueError, match="divisible"):
        _lib.workspace_bytes(bad)
    bad = m._make_cfg(8, 16, 149, 7, 0, False, False)
    with pytest.raises(ValueError):
        _lib.workspace_bytes(bad)
    assert _lib.load().xattn_workspace_bytes(None, None) == _lib.ERR_NULL
    assert "NULL" in _lib.last_error()


def test_no_cpu_fallback():
    m = make_head("concat", True).eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(2, 16, 512), torch.zeros(2, 149, 768))
    with pytest.raises(ValueError):
        m(torch.zeros(2, 512), torch.zeros(2, 149, 768))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from multimodalemotionrecognition_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setenv("XATTN_FUSION_LIB", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="not built"):
        _lib.load()


def test_unsupported_reference_options_raise_cleanly():
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    with pytest.raises(ValueError, match="temporal pooling"):
        XAttnFusionHead(temporal_pooling="nope")                      # the reference's own message (temporal.py:104)
    with pytest.raises(ValueError, match="encoder layers"):
        XAttnFusionHead(temporal_pooling="transformer", temporal_num_layers=9)
    with pytest.raises(ValueError):
        XAttnFusionHead(xattn_head="nope")


def test_reference_shape_test_configuration_builds():
    """The reference's own xattn test builds d_model=8, num_heads=2 (tests/test_attention_integration.py:83-91)."""
    from multimodalemotionrecognition_b200 import FusionModel, _lib

    class V(torch.nn.Module):
        embedding_dim = 16

        def __init__(self):
            super().__init__()
            self.backbone = torch.nn.Identity()

    class A(torch.nn.Module):
        sequence_dim = 8

        def encode_sequence(self, x):
            return x

    fm = FusionModel(A(), V(), num_classes=8, mode="xattn", d_model=8, num_heads=2, audio_n_mels=64).eval()
    cfg = fm._make_cfg(2, 4, 12, _lib.DTYPE_F32, _lib.DTYPE_F32, False, False)
    assert _lib.workspace_bytes(cfg) > 0
    assert fm.mode == "xattn" and fm.pop_alignment_loss() is None
    assert "v2a_attn.in_proj_weight" in fm.state_dict() and "audio_time_conv.weight" in fm.state_dict()


@pytest.mark.parametrize("mode", ["concat", "gated"])
def test_flat_head_seed0_init_matches_reference(mode):
    from multimodalemotionrecognition_b200 import FlatFusionHead, _lib

    blob = load_npz("flat_config5.npz")
    torch.manual_seed(0)
    m = FlatFusionHead(num_classes=8, video_dim=512, audio_dim=768, common_dim=256, mode=mode)
    sd = m.state_dict()
    assert sorted(sd) == list(blob[f"{mode}:keys"])
    for k, d in zip(blob[f"{mode}:keys"], blob[f"{mode}:sha"]):
        assert sha(sd[str(k)]) == str(d), k
    offs, sizes, total = _lib.flat_param_layout(m._flat_cfg(4, 16, 149, 0))
    for name, n in zip(_lib.flat_param_names(mode == "gated"), sizes):
        if n:
            assert sd[name].numel() == n, name
    assert "semantic_alignment.logit_scale" in sd


def test_flat_parameter_views_follow_load_state_dict():
    m = make_head("concat", True)
    flat = m.flat_parameters()
    assert flat.numel() >= 398581
    sd = {k: torch.full_like(v, 0.5) for k, v in m.state_dict().items()}
    m.load_state_dict(sd)
    assert m.flat_parameters() is flat          # views survived the in-place load
    off, n = m._xattn_slices[0]
    assert torch.all(flat[off:off + n] == 0.5)
    g = m.flat_gradients()
    assert m.v_in_proj.weight.grad.data_ptr() == g.data_ptr() + 4 * off


def test_flat_head_training_config_and_modality_dropout_draws():
    """Host logic of the trainable flat heads (no GPU): the config struct carries the training fields behind the inference ones
    (layout as in include/xattn_fusion.h), and the two ModalityDropout decisions are drawn like the reference draws them —
    ``torch.rand(1).item() < p``, audio first (fusion.py:47-53) — so a torch seed decides the same drops on both sides."""
    from multimodalemotionrecognition_b200 import FlatFusionHead, _lib

    assert C.sizeof(_lib.FlatHeadConfig) == 80 and _lib.FlatHeadConfig.training.offset == 44 and _lib.FlatHeadConfig.seed.offset == 64
    m = FlatFusionHead(num_classes=8, video_dim=12, audio_dim=10, common_dim=16, mode="gated", fusion_align_mode="clip", fusion_align_dim=32)
    m.train()
    seen = set()
    for seed in range(40):
        torch.manual_seed(seed)
        want = (int(torch.rand(1).item() < 0.2), int(torch.rand(1).item() < 0.2))
        torch.manual_seed(seed)
        cfg = m._flat_train_cfg(m._flat_cfg(4, 3, 5, _lib.DTYPE_F32))
        assert (cfg.drop_audio, cfg.drop_video) == want and cfg.training == 1 and abs(cfg.p_drop - 0.2) < 1e-7
        assert cfg.precision == _lib.DTYPE_F32 and cfg.offset == seed + 1
        seen.add(want)
    assert len(seen) >= 3
    m.eval()
    cfg = m._flat_train_cfg(m._flat_cfg(4, 3, 5, _lib.DTYPE_F32))
    assert cfg.training == 0 and cfg.drop_audio == 0 and cfg.drop_video == 0
    mc = FlatFusionHead(num_classes=8, video_dim=12, audio_dim=10, common_dim=16, mode="concat", fusion_align_mode="none").train()
    st = torch.get_rng_state()
    cfg = mc._flat_train_cfg(mc._flat_cfg(4, 3, 5, _lib.DTYPE_F32))
    assert torch.equal(st, torch.get_rng_state()) and cfg.drop_audio == 0      # concat mode has no ModalityDropout: no draw


def test_step_args_struct_matches_the_header():
    from multimodalemotionrecognition_b200 import _lib

    a = _lib.XAttnStepArgs
    # 12 pointers, int64, 6 floats, 4 int32, float -> 96 + 8 + 24 + 16 + 4 = 148 -> padded to 152
    assert C.sizeof(a) == 152 and a.workspace_bytes.offset == 96 and a.lr.offset == 104 and a.step.offset == 128
    assert a.phase.offset == 140 and a.grad_scale.offset == 144
