"""Host-side multi-rank logic on CPU (gloo, world_size 2): batch sharding, the ragged logits gather, and the gradient
bucket all-reduce convention of Stage1Trainer (loss gradients pre-scaled by 1/world, bucket summed).  The compute path
itself needs a B200 and is covered by the ``-m gpu`` tests; nothing here launches a kernel."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from multimodalemotionrecognition_b200.training import Stage1Trainer, gather_batch, gather_logits, shard_range


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _FakeHead:
    """Just the flat-buffer surface Stage1Trainer.allreduce_gradients touches."""

    def __init__(self, grads):
        self._g = grads

    def flat_gradients(self, attach=True):
        return self._g


def _worker(rank: int, world: int, port: int, total: int):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. contiguous shards cover the batch exactly once, remainder to the low ranks
        s, e = shard_range(total, rank, world)
        full = torch.arange(total * 8, dtype=torch.float32).view(total, 8)
        local = full[s:e].clone()
        # 2. ragged gather reproduces the single-process logits on every rank
        got = gather_logits(local, total)
        assert torch.equal(got, full), (rank, got.shape)
        # 3. gradient bucket: each rank's gradient of its local mean loss scaled by 1/world, summed == global mean
        g_local = torch.full((10,), float(rank + 1)) / world
        tr = Stage1Trainer(_FakeHead(g_local))
        assert tr._world() == world
        tr.allreduce_gradients()
        expect = sum(float(r + 1) for r in range(world)) / world
        assert torch.allclose(g_local, torch.full((10,), expect))
        # 4. the two-bucket form (tail bucket in flight from forward_backward, head bucket + join in allreduce_gradients)
        g2 = torch.full((10,), float(rank + 1)) / world
        tr2 = Stage1Trainer(_FakeHead(g2), overlap_allreduce=True)
        tr2._tail_offset = 4
        tr2._tail_work = dist.all_reduce(g2[4:], op=dist.ReduceOp.SUM, async_op=True)
        tr2.allreduce_gradients()
        assert tr2._tail_work is None and torch.allclose(g2, torch.full((10,), expect))
        # 5. differentiable batch gather for the cross-sample alignment loss: a B x B term on the gathered rows, each rank adding
        #    1/world of it, gives every rank's rows the gradient of the single-process loss
        gen = torch.Generator().manual_seed(5)
        xa_full = torch.randn(total, 6, generator=gen, dtype=torch.float64)
        xv_full = torch.randn(total, 6, generator=gen, dtype=torch.float64)

        def clip_like(a, v):
            logits = a @ v.t()
            return (logits.logsumexp(dim=1) - logits.diagonal()).mean() + (logits.logsumexp(dim=0) - logits.diagonal()).mean()

        a_ref = xa_full.clone().requires_grad_(True)
        v_ref = xv_full.clone().requires_grad_(True)
        clip_like(a_ref, v_ref).backward()
        a_loc = xa_full[s:e].clone().requires_grad_(True)
        v_loc = xv_full[s:e].clone().requires_grad_(True)
        loss = clip_like(gather_batch(a_loc, total), gather_batch(v_loc, total)) / world
        loss.backward()
        if e > s:
            assert torch.allclose(a_loc.grad, a_ref.grad[s:e], atol=1e-12) and torch.allclose(v_loc.grad, v_ref.grad[s:e], atol=1e-12)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total", [8, 7, 1])
def test_world2_gloo_shard_gather_allreduce(total):
    mp.spawn(_worker, args=(2, _free_port(), total), nprocs=2, join=True)


def test_shard_range_partitions():
    for total in (0, 1, 5, 4096, 16385):
        for world in (1, 2, 3, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and b - a >= d - c >= 0
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1


def test_frozen_embedding_cache_runs_backbones_once_per_sample():
    """Stage-1 embeddings are computed the first time a dataset index shows up and served from the cache
    (bf16) afterwards; host-side logic, CPU tensors."""
    import torch
    from torch import nn

    from multimodalemotionrecognition_b200.training import FrozenEmbeddingCache

    calls = {"video": 0, "audio": 0}

    class Video(nn.Module):
        embedding_dim = 12

        def __init__(self):
            super().__init__()

            class Backbone(nn.Module):
                def forward(self, x):
                    calls["video"] += x.shape[0]
                    return x.flatten(1)[:, :12]
            self.backbone = Backbone()

    class Audio(nn.Module):
        def encode_sequence(self, a):
            calls["audio"] += a.shape[0]
            return a.view(a.shape[0], 5, 10)

    class Model:
        video_model, audio_model = Video(), Audio()

    cache = FrozenEmbeddingCache(num_samples=9, T=3, Ta=5, Dv=12, Da=10, device="cpu")
    assert cache.nbytes() == 9 * (3 * 12 + 5 * 10) * 2
    g = torch.Generator().manual_seed(0)
    data_v = torch.randn(9, 3, 3, 2, 2, generator=g)
    data_a = torch.randn(9, 50, generator=g)
    idx = torch.tensor([4, 1, 7])
    v, a = cache.fetch_or_compute(Model, idx, data_v[idx], data_a[idx])
    assert v.dtype == torch.bfloat16 and v.shape == (3, 3, 12) and a.shape == (3, 5, 10)
    assert calls == {"video": 9, "audio": 3} and (cache.hits, cache.misses) == (0, 3)
    assert torch.equal(v, data_v[idx].reshape(9, -1)[:, :12].view(3, 3, 12).bfloat16())
    idx2 = torch.tensor([1, 2, 4])
    v2, a2 = cache.fetch_or_compute(Model, idx2, data_v[idx2], data_a[idx2])
    assert calls == {"video": 12, "audio": 4} and (cache.hits, cache.misses) == (2, 4)   # only sample 2 was new
    assert torch.equal(a2[0], a[1]) and torch.equal(v2[2], v[0])
    assert cache.missing(torch.arange(9)).tolist() == [True, False, False, True, False, True, True, False, True]


def test_hostmem_numa_binding_against_a_fake_sysfs(tmp_path, monkeypatch):
    """hostmem: cpulist parsing, GPU -> NUMA node lookup and the affinity call, against a fabricated sysfs tree."""
    import os

    from multimodalemotionrecognition_b200 import hostmem

    assert hostmem._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert hostmem._parse_cpulist("") == set()
    monkeypatch.setattr(hostmem, "pci_address", lambda device: "0000:1b:00.0")
    dev = tmp_path / "bus" / "pci" / "devices" / "0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("-1\n")
    assert hostmem.numa_node_of(0, str(tmp_path)) is None            # virtualised box: nothing to bind to
    assert hostmem.bind_to_device_numa_node(0, str(tmp_path)) is None
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices" / "system" / "node" / "node1"
    node.mkdir(parents=True)
    mine = sorted(os.sched_getaffinity(0))
    (node / "cpulist").write_text(f"{mine[0]}\n")
    calls = []
    monkeypatch.setattr(os, "sched_setaffinity", lambda pid, cpus: calls.append((pid, set(cpus))))
    assert hostmem.bind_to_device_numa_node(0, str(tmp_path)) == 1
    assert calls == [(0, {mine[0]})]
    (node / "cpulist").write_text("100000\n")                         # no CPU of that node is ours: leave the process alone
    assert hostmem.bind_to_device_numa_node(0, str(tmp_path)) is None
