"""Diagnostic: how do the three tcgen05 tf32 GEMM forms read their fp32 operands — truncation or round-to-nearest?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multimodalemotionrecognition_b200 import _lib
from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr
dev = torch.device("cuda:0")
def trunc(x): return (x.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)
def rn(x):
    i = x.contiguous().view(torch.int32)
    r = i + 0x0FFF + ((i >> 13) & 1)
    return (r & ~0x1FFF).view(torch.float32)
def rna(x):   # round half away (cvt.rna)
    i = x.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)
def gemm(form, A, B, C, M, N, K):
    _lib.debug_check(_lib.load_debug().xattn_debug_gemm_tf32(form, _ptr(A), _ptr(B), None, _ptr(C), M, N, K, 0, 0, _stream_ptr(dev)))
    torch.cuda.synchronize()
g = torch.Generator().manual_seed(0)
for M, N, K in ((96, 256, 256), (1536, 128, 512), (4096, 128, 128)):
    # form 0: C[M,N] = A[M,K] W[N,K]^T
    A = torch.randn(M, K, generator=g).to(dev); W = torch.randn(N, K, generator=g).to(dev); C = torch.zeros(M, N, device=dev)
    gemm(0, A, W, C, M, N, K)
    for name, f in (("trunc", trunc), ("rn", rn), ("rna", rna), ("exact", lambda x: x)):
        ref = f(A).double() @ f(W).double().T
        print(f"form0 M={M} N={N} K={K} vs {name:5s}: {(C.double()-ref).abs().max().item()/ref.abs().max().item():.2e}")
    # form 1: C[M,Kout] = dY[M,Nred] W[Nred,Kout]
    dY = torch.randn(M, N, generator=g).to(dev); W2 = torch.randn(N, K, generator=g).to(dev); C = torch.zeros(M, K, device=dev)
    gemm(1, dY, W2, C, M, K, N)
    for name, f in (("trunc", trunc), ("rn", rn), ("rna", rna), ("exact", lambda x: x)):
        ref = f(dY).double() @ f(W2).double()
        print(f"form1 M={M} Nred={N} Kout={K} vs {name:5s}: {(C.double()-ref).abs().max().item()/ref.abs().max().item():.2e}")
    # form 2: C[N,K] += dY[M,N]^T X[M,K]
    X = torch.randn(M, K, generator=g).to(dev); C = torch.zeros(N, K, device=dev)
    gemm(2, dY, X, C, M, N, K)
    for name, f in (("trunc", trunc), ("rn", rn), ("rna", rna), ("exact", lambda x: x)):
        ref = f(dY).double().T @ f(X).double()
        print(f"form2 M={M} N={N} K={K} vs {name:5s}: {(C.double()-ref).abs().max().item()/ref.abs().max().item():.2e}")
