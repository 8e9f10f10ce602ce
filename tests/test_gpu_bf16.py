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
    _lib.check(_lib.load().xattn_debug_gemm_tc(_ptr(x), _ptr(w), _ptr(bias), _ptr(out), int(out_bf16), pitch, M, N, K,
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
