"""GPU tests (``-m gpu``) of the data-parallel Stage-1 step: the parameters after one
batch-sharded step — per-rank forward/backward on a contiguous shard, loss gradient pre-scaled by 1/world, ONE sum
all-reduce of the flat gradient bucket, fused Adam — equal the parameters after one single-rank step on the whole batch.

* ``test_sharded_gradients_sum_to_full_batch_gradients``: the arithmetic of the all-reduce emulated on ONE GPU (always runs).
* ``test_two_rank_nccl_step_equals_single_rank_step``: two processes, two GPUs, NCCL (skipped on a 1-GPU box; run with
  ``gpurun --gpus 2``).  Also covers the bucketed / overlapped all-reduce against the single-bucket one."""
import os
import socket

import pytest
import torch

from helpers import make_head

pytestmark = pytest.mark.gpu
# Gradients are compared at fp32 summation-order level.  Parameters get a looser bound: Adam's first step is
# lr * g / (|g| + eps), so a coordinate whose gradient is of the order of eps = 1e-8 turns a 1e-10 difference in g into a
# 1e-5 difference in the parameter (measured worst case over the 400k parameters: 1.5e-6).
GRAD_TOL = 2e-6
PARAM_TOL = 2e-5


def _batch(B, seed=21):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(B, 16, 512, generator=g), torch.randn(B, 149, 768, generator=g), torch.randint(0, 8, (B,), generator=g))


def _single_rank_step(dev, xv, xa, lab, precision="fp32"):
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    m = make_head("concat", True, seed=0, precision=precision).to(dev).train()
    tr = Stage1Trainer(m, lr=1e-3, weight_decay=1e-4)
    loss = tr.step(xv.to(dev), xa.to(dev), lab.to(dev), training=False)
    torch.cuda.synchronize(dev)
    return float(loss), m.flat_parameters().detach().cpu().clone(), m.flat_gradients(attach=False).detach().cpu().clone()


def test_sharded_gradients_sum_to_full_batch_gradients(cuda_device):
    from multimodalemotionrecognition_b200.training import Stage1Trainer, shard_range

    B, world = 48, 2
    xv, xa, lab = _batch(B)
    loss_full, p_full, g_full = _single_rank_step(cuda_device, xv, xa, lab)
    total = torch.zeros_like(g_full)
    losses = []
    for r in range(world):
        s, e = shard_range(B, r, world)
        m = make_head("concat", True, seed=0, precision="fp32").to(cuda_device).train()
        tr = Stage1Trainer(m)
        losses.append(float(tr.forward_backward(xv[s:e].to(cuda_device), xa[s:e].to(cuda_device), lab[s:e].to(cuda_device),
                                                training=False)))
        total += m.flat_gradients(attach=False).detach().cpu() / world     # what grad_scale = 1/world + the sum all-reduce produce
    assert abs(sum(losses) / world - loss_full) <= 1e-6
    scale = g_full.abs().max().item()
    assert (total - g_full).abs().max().item() <= 2e-6 * max(scale, 1.0), ((total - g_full).abs().max().item(), scale)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, B, bucketed, out_path):
    import torch.distributed as dist

    from multimodalemotionrecognition_b200.training import Stage1Trainer, shard_range

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        xv, xa, lab = _batch(B)
        s, e = shard_range(B, rank, world)
        m = make_head("concat", True, seed=0, precision="fp32").to(dev).train()
        tr = Stage1Trainer(m, lr=1e-3, weight_decay=1e-4, overlap_allreduce=bucketed)
        loss = tr.step(xv[s:e].to(dev), xa[s:e].to(dev), lab[s:e].to(dev), training=False)
        torch.cuda.synchronize(dev)
        flat = m.flat_parameters().detach()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        gflat = m.flat_gradients(attach=False).detach()
        lt = torch.tensor([float(loss)], device=dev)
        dist.all_reduce(lt)
        if rank == 0:
            torch.save({"params": [t.cpu() for t in gathered], "grads": gflat.cpu(), "mean_loss": float(lt) / world}, out_path)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("bucketed", [False, True])
def test_two_rank_nccl_step_equals_single_rank_step(cuda_device, tmp_path, bucketed):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    B, world = 48, 2
    out_path = str(tmp_path / "ranks.pt")
    mp.spawn(_rank_main, args=(world, _free_port(), B, bucketed, out_path), nprocs=world, join=True)
    got = torch.load(out_path)
    xv, xa, lab = _batch(B)
    loss_full, p_full, g_full = _single_rank_step(cuda_device, xv, xa, lab)
    assert abs(got["mean_loss"] - loss_full) <= 1e-6
    gerr = (got["grads"] - g_full).abs().max().item()
    assert gerr <= GRAD_TOL * max(g_full.abs().max().item(), 1.0), gerr       # the all-reduced bucket == the full-batch gradient
    for r, p in enumerate(got["params"]):
        assert (p - p_full).abs().max().item() <= PARAM_TOL, (r, (p - p_full).abs().max().item())
    assert torch.equal(got["params"][0], got["params"][1])      # replicas stay bit-identical


def _rank_main_steps(rank, world, port, B, steps, out_path):
    import torch.distributed as dist

    from multimodalemotionrecognition_b200.training import Stage1Trainer, shard_range

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        s, e = shard_range(B, rank, world)
        m = make_head("concat", True, seed=0, precision="tf32").to(dev).train()
        tr = Stage1Trainer(m, lr=1e-3, weight_decay=1e-4, eps=1e-3)            # graph on: two graphs around the all-reduce
        losses = []
        for k in range(steps):
            xv, xa, lab = _batch(B, seed=21 + (k % 2))
            losses.append(tr.step(xv[s:e].bfloat16().to(dev), xa[s:e].bfloat16().to(dev), lab[s:e].to(dev), training=False))
        torch.cuda.synchronize(dev)
        flat = m.flat_parameters().detach()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        lt = torch.stack(losses)
        dist.all_reduce(lt)
        if rank == 0:
            torch.save({"params": [t.cpu() for t in gathered], "mean_losses": (lt / world).cpu()}, out_path)
    finally:
        dist.destroy_process_group()


def test_two_rank_graphed_steps_track_the_single_rank_separate_calls(cuda_device, tmp_path):
    """Four data-parallel tf32 steps on two GPUs — the first through the separate calls, the rest through the two CUDA graphs of
    ``xattn_stage1_step`` (phase 1 | tail-bucket all-reduce beside phase 2 | head-bucket all-reduce | Adam) — against four
    single-rank steps on the whole batch through the separate calls: losses and parameters."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    from multimodalemotionrecognition_b200.training import Stage1Trainer

    B, world, steps = 48, 2, 4
    out_path = str(tmp_path / "ranks_steps.pt")
    mp.spawn(_rank_main_steps, args=(world, _free_port(), B, steps, out_path), nprocs=world, join=True)
    got = torch.load(out_path)
    m = make_head("concat", True, seed=0, precision="tf32").to(cuda_device).train()
    tr = Stage1Trainer(m, lr=1e-3, weight_decay=1e-4, eps=1e-3, graph=False)
    ref = []
    for k in range(steps):
        xv, xa, lab = _batch(B, seed=21 + (k % 2))
        ref.append(float(tr.step(xv.bfloat16().to(cuda_device), xa.bfloat16().to(cuda_device), lab.to(cuda_device), training=False)))
    p_full = m.flat_parameters().detach().cpu()
    for k in range(steps):
        assert abs(float(got["mean_losses"][k]) - ref[k]) <= 2e-4 * max(1.0, abs(ref[k])), (k, float(got["mean_losses"][k]), ref[k])
    assert ref[-1] < ref[0]
    for r, p in enumerate(got["params"]):
        assert (p - p_full).abs().max().item() <= 1e-3, (r, (p - p_full).abs().max().item())
    assert torch.equal(got["params"][0], got["params"][1])      # replicas stay bit-identical
