"""xattn_stage1_step / Stage1Trainer(graph=True): the Stage-1 step replayed as a CUDA graph (captured again and patched with
cudaGraphExecUpdate every step, because the Philox offset and the Adam step count are kernel arguments) must be the step the
separate calls run: same losses, same parameters, step after step, with the dropouts ON — a stale offset or step count in the
replayed graph would show at once."""
import numpy as np
import pytest
import torch

from helpers import make_head

pytestmark = pytest.mark.gpu


def _data(dev, B, seed):
    g = torch.Generator().manual_seed(seed)
    xv = torch.randn(B, 16, 512, generator=g).bfloat16().to(dev)
    xa = torch.randn(B, 149, 768, generator=g).bfloat16().to(dev)
    lab = torch.randint(0, 8, (B,), generator=g).to(dev)
    return xv, xa, lab


@pytest.mark.parametrize("precision", ["tf32", "fp32"])
def test_graphed_step_equals_the_separate_calls(cuda_device, precision):
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    torch.manual_seed(1234)      # (the Philox seed of both heads is torch's initial seed)
    heads = [make_head("concat", True, seed=0, precision=precision).to(cuda_device).train() for _ in range(2)]
    # eps = 1e-3: with Adam's default 1e-8 a parameter whose gradient is rounding noise (the attention key bias) moves by ±lr per
    # step on the sign of that noise, and the weight gradients meet in atomics — two eager runs would differ by 1e-3 there too
    eager = Stage1Trainer(heads[0], lr=1e-3, weight_decay=1e-4, eps=1e-3, label_smoothing=0.05, graph=False)
    graphed = Stage1Trainer(heads[1], lr=1e-3, weight_decay=1e-4, eps=1e-3, label_smoothing=0.05, graph=True)
    sets = [_data(cuda_device, 48, 5), _data(cuda_device, 48, 6), _data(cuda_device, 20, 7)]
    order = [0, 1, 0, 1, 2, 2, 0]          # the batch size changes on the way: the cached graph is rebuilt, then patched again
    losses, logits = [], []
    for k, i in enumerate(order):          # no host synchronisation between steps: updates meet launches still in flight
        le = eager.step(*sets[i], training=True)
        logits_e = eager._logits.clone()
        lg = graphed.step(*sets[i], training=True)
        losses.append((le, lg))
        logits.append((logits_e, graphed._logits.clone()))
    ltol = 2e-5 if precision == "fp32" else 2e-4
    for k, (le, lg) in enumerate(losses):
        assert abs(float(le) - float(lg)) <= ltol * max(1.0, abs(float(le))), (k, float(le), float(lg))
        # training-mode logits carry every dropout mask of the step: a stale Philox offset moves them by ~0.1
        assert (logits[k][0] - logits[k][1]).abs().max().item() <= 50 * ltol, k
    assert eager.step_count == graphed.step_count == len(order)
    assert heads[0]._rng_offset == heads[1]._rng_offset == len(order)
    # dW meets in atomics (summation order); under tf32 the first layer's dY is rounded to bf16, where one ulp of noise can flip a
    # rounding (0.4 % of that element): two eager tf32 runs differ by ~1e-5 in the loss after three steps (scripts/exp_ws_poison.py)
    tol = 2e-5 if precision == "fp32" else 1e-3
    sd0, sd1 = heads[0].state_dict(), heads[1].state_dict()
    for name in sd0:
        np.testing.assert_allclose(sd1[name].cpu().numpy(), sd0[name].cpu().numpy(), atol=tol, rtol=0, err_msg=name)
    # the losses moved (the optimizer really stepped) and differ between steps (fresh masks every step)
    vals = [float(l[1]) for l in losses]
    assert len({round(v, 5) for v in vals}) == len(vals)
    # eval-mode step through the same entry point (dropouts identity) against the separate calls
    le = eager.step(*sets[0], training=False)
    lg = graphed.step(*sets[0], training=False)
    assert abs(float(le) - float(lg)) <= ltol * max(1.0, abs(float(le)))


def test_stage1_step_entry_point_rejects_bad_arguments(cuda_device):
    import ctypes as C

    from multimodalemotionrecognition_b200 import _lib

    a = _lib.XAttnStepArgs()
    lib = _lib.load()
    assert lib.xattn_stage1_step(C.byref(a), None) == _lib.ERR_NULL           # cfg is NULL
    head = make_head("concat", True, seed=0, precision="fp32").to(cuda_device)
    cfg = head._make_cfg(4, 16, 149, _lib.DTYPE_F32, _lib.DTYPE_F32, True, True)
    a.cfg = C.pointer(cfg)
    assert lib.xattn_stage1_step(C.byref(a), None) == _lib.ERR_NULL           # buffers are NULL
    assert b"NULL" in lib.xattn_last_error()


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_training_step_on_a_non_default_stream(cuda_device, precision):
    """Every launch of forward / backward must go to the caller's stream (or to the library's side stream, forked from and joined
    to it): on a non-blocking torch stream a kernel that slipped onto the legacy default stream would run unordered.  Gradients
    on a side stream against the same step on the default stream."""
    xv, xa, lab = _data(cuda_device, 32, 11)
    grads = []
    for use_side in (False, True):
        torch.manual_seed(4321)
        head = make_head("concat", True, seed=0, precision=precision).to(cuda_device).train()
        head._rng_offset = 3
        stream = torch.cuda.Stream(cuda_device) if use_side else torch.cuda.current_stream(cuda_device)
        stream.wait_stream(torch.cuda.current_stream(cuda_device))
        with torch.cuda.stream(stream):
            for _ in range(3):                       # (repeat: a racing kernel has to lose only once)
                head.zero_grad(set_to_none=True)
                loss = torch.nn.functional.cross_entropy(head(xv, xa), lab)
                head._rng_offset = 3
                loss.backward()
        torch.cuda.current_stream(cuda_device).wait_stream(stream)
        torch.cuda.synchronize(cuda_device)
        grads.append({n: p.grad.detach().clone() for n, p in head.named_parameters() if p.grad is not None})
    tol = 1e-6 if precision == "fp32" else 2e-5
    for n, g0 in grads[0].items():
        assert (g0 - grads[1][n]).abs().max().item() <= tol * max(1.0, g0.abs().max().item()), n


def test_host_batch_prefetcher_delivers_each_batch_once_and_intact(cuda_device):
    """HostBatchPrefetcher: the copy of batch i+1 runs beside the work on batch i; every ``get`` returns exactly the batch of the
    matching ``put``, also when the consumer is slow (a long kernel between get and release) or fast."""
    from multimodalemotionrecognition_b200.training import HostBatchPrefetcher

    g = torch.Generator().manual_seed(3)
    host = [(torch.randn(64, 149, 768, generator=g).bfloat16().pin_memory(), torch.randint(0, 8, (64,), generator=g).pin_memory())
            for _ in range(5)]
    pre = HostBatchPrefetcher(cuda_device)
    pre.put(*host[0])
    sums = []
    big = torch.randn(4096, 4096, device=cuda_device)
    for i in range(5):
        x, lab = pre.get()
        if i % 2 == 0:
            big = big @ big * 1e-4                # a slow consumer: the next copy must not overtake the read below
        sums.append((x.float().sum(), lab.sum(), x[7, 3, :5].float().clone()))
        pre.release()
        if i + 1 < 5:
            pre.put(*host[i + 1])
    torch.cuda.synchronize(cuda_device)
    for i, (s, ls, piece) in enumerate(sums):
        assert abs(float(s) - float(host[i][0].float().sum())) <= 1e-2 * max(1.0, abs(float(s))), i
        assert int(ls) == int(host[i][1].sum()), i
        assert torch.equal(piece.cpu(), host[i][0][7, 3, :5].float()), i
