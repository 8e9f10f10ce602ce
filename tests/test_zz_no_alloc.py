"""GPU test (``-m gpu``) of ``xattn_forward_noalloc`` (include/xattn_fusion.h): the forward whose scratch workspace is a
static ``__device__`` array inside the library, for callers that may not allocate.  It has no counterpart in the
reference, so the check is against the existing entry point
``xattn_forward`` on the same inputs: same kernels, same launch order on one stream -> the logits must agree to the
tolerance of the other GPU tests (1e-2 bf16 / 1e-4 fp32; measured: bit-identical).  Two calls in a row on different
inputs show that the shared array's stale contents do not leak into the next call.

Collected last on purpose (file name).  Passes alone in a fresh process and after
the other tests."""
import ctypes as C

import pytest
import torch

from helpers import make_head

pytestmark = pytest.mark.gpu


def _both_entry_points(m, xv, xa, precision):
    """(logits via xattn_forward with a torch workspace, logits via xattn_forward_noalloc) through the C ABI."""
    from multimodalemotionrecognition_b200 import _lib
    from multimodalemotionrecognition_b200.fusion import _io_dtype, _ptr, _stream_ptr

    dev = xv.device
    lib = _lib.load()
    B, T, _ = xv.shape
    Ta = xa.shape[1]
    with torch.cuda.device(dev), torch.no_grad():
        flat = m.flat_parameters()
        cfg = m._make_cfg(B, T, Ta, _io_dtype(xv), precision, False, False)
        packed = m._packed_weights(cfg, flat) if precision == _lib.DTYPE_BF16 else None
        ws = torch.empty(_lib.workspace_bytes(cfg), dtype=torch.uint8, device=dev)
        ref = torch.full((B, m.num_classes), float("nan"), dtype=torch.float32, device=dev)
        got = torch.full((B, m.num_classes), float("nan"), dtype=torch.float32, device=dev)
        _lib.check(lib.xattn_forward(C.byref(cfg), _ptr(flat), _ptr(packed), _ptr(xv), _ptr(xa), _ptr(ref), _ptr(ws), ws.numel(),
                                     _stream_ptr(dev)))
        _lib.check(lib.xattn_forward_noalloc(C.byref(cfg), _ptr(flat), _ptr(packed), _ptr(xv), _ptr(xa), _ptr(got),
                                             _stream_ptr(dev)))
        torch.cuda.synchronize(dev)
    return ref.cpu(), got.cpu()


@pytest.mark.parametrize("precision,B,tol", [("bf16", 8, 1e-2), ("bf16", 64, 1e-2), ("fp32", 8, 1e-4)])
def test_noalloc_forward_matches_xattn_forward_twice_in_a_row(cuda_device, precision, B, tol):
    from multimodalemotionrecognition_b200 import _lib

    prec = _lib.DTYPE_BF16 if precision == "bf16" else _lib.DTYPE_F32
    m = make_head("concat", True, seed=0, precision=precision).to(cuda_device).eval()
    g = torch.Generator().manual_seed(77 + B)
    for call in range(2):   # two calls in a row share the static array; the second sees the first one's leftovers
        xv = torch.randn(B, 16, 512, generator=g)
        xa = torch.randn(B, 149, 768, generator=g)
        if precision == "bf16":
            xv, xa = xv.bfloat16(), xa.bfloat16()
        ref, got = _both_entry_points(m, xv.to(cuda_device).contiguous(), xa.to(cuda_device).contiguous(), prec)
        assert torch.isfinite(ref).all() and torch.isfinite(got).all(), call
        err = (ref - got).abs().max().item()
        assert err <= tol, (precision, B, call, err)


def test_noalloc_rejects_configs_beyond_the_static_scratch(cuda_device):
    from multimodalemotionrecognition_b200 import _lib
    from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr

    lib = _lib.load()
    cap = int(lib.xattn_noalloc_max_workspace_bytes())
    assert cap == 8 * 1024 * 1024
    m = make_head("concat", True, seed=0, precision="bf16").to(cuda_device).eval()
    B = 72
    with torch.cuda.device(cuda_device), torch.no_grad():
        flat = m.flat_parameters()
        cfg = m._make_cfg(B, 16, 149, _lib.DTYPE_BF16, _lib.DTYPE_BF16, False, False)
        assert _lib.workspace_bytes(cfg) > cap
        packed = m._packed_weights(cfg, flat)
        xv = torch.zeros(B, 16, 512, dtype=torch.bfloat16, device=cuda_device)
        xa = torch.zeros(B, 149, 768, dtype=torch.bfloat16, device=cuda_device)
        out = torch.zeros(B, 8, device=cuda_device)
        rc = lib.xattn_forward_noalloc(C.byref(cfg), _ptr(flat), _ptr(packed), _ptr(xv), _ptr(xa), _ptr(out), _stream_ptr(cuda_device))
        assert rc == _lib.ERR_WORKSPACE
        assert "static array" in _lib.last_error()
        train = m._make_cfg(8, 16, 149, _lib.DTYPE_BF16, _lib.DTYPE_F32, True, True)
        rc = lib.xattn_forward_noalloc(C.byref(train), _ptr(flat), None, _ptr(xv), _ptr(xa), _ptr(out), _stream_ptr(cuda_device))
        assert rc == _lib.ERR_SHAPE
