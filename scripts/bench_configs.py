#!/usr/bin/env python
"""Measurements of BASELINE.json configs[2..4] (the bench.py line covers configs[1]); one JSON object per line on rank 0.

    python scripts/bench_configs.py [sweep] [train] [flat]                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_configs.py sweep train                                     # N GPUs, one rank per GPU

* sweep (configs[2]): xattn fusion forward bf16, total batch B in {1,2,4,...,16384} cut into contiguous shards over the
  ranks (B//G each, remainder to the low ranks); the timed region holds the forward of the local shard AND the logits
  all-gather.  Ranks whose shard is empty still join the gather.
* train (configs[3]): Stage-1 step (zero-grad + forward + CE + backward + gradient all-reduce + Adam) at B=1024 per GPU,
  reference dropout defaults (0.1 / 0.1 / 0.2 / 0.1), lr 1e-3, wd 1e-4.
* flat  (configs[4]): concat and gated flat heads with the CLIP-style alignment projections, forward at B=4096 from the same
  sequence embeddings as configs[1] (temporal mean fused in) and from pre-pooled embeddings.
All times are CUDA events on the launching stream after warm-up, max over ranks.  Inputs larger than L2 are rotated
between two sets; small-batch points are latency measurements and sit in L2 by nature (said in each line).
"""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist

from multimodalemotionrecognition_b200 import FlatFusionHead, XAttnFusionHead, _lib
from multimodalemotionrecognition_b200.training import Stage1Trainer, gather_logits, shard_range

T, TA, DV, DA, NCLS = 16, 149, 512, 768, 8
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)


def timed(fn, steps, warmup):
    """ms per call of fn(i): CUDA events around `steps` calls, max over ranks."""
    for i in range(warmup):
        fn(i)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(d):
    if rank == 0:
        print(json.dumps(d), flush=True)


def make_inputs(b, nsets, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    sets = []
    for _ in range(nsets):
        sets.append((torch.randn(b, T, DV, device=dev, generator=g).bfloat16(), torch.randn(b, TA, DA, device=dev, generator=g).bfloat16(),
                     torch.randint(0, NCLS, (b,), device=dev, generator=g)))
    return sets


def run_sweep():
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True, precision="bf16").to(dev).eval()
    B = 1
    while B <= 16384:
        s, e = shard_range(B, rank, world)
        b = e - s
        in_bytes = b * (T * DV + TA * DA) * 2
        nsets = 2 if in_bytes > 64e6 else 1
        sets = make_inputs(max(b, 0), nsets, 1 + rank) if b > 0 else None

        def step(i):
            with torch.no_grad():
                if b > 0:
                    xv, xa, _ = sets[i % nsets]
                    out = head(xv, xa)
                else:
                    out = torch.empty((0, NCLS), dtype=torch.float32, device=dev)
                return gather_logits(out, B)

        steps = 100 if B <= 1024 else 30
        ms = timed(step, steps, 10)
        ms_graph = None
        if 0 < b <= 1024:   # launch-bound sizes: the same forward replayed as one CUDA graph over static buffers (GraphedForward)
            gf = head.graphed_forward(b)
            gf.video.copy_(sets[0][0])
            gf.audio.copy_(sets[0][1])
            ms_graph = timed(lambda i: gather_logits(gf(), B), steps, 10)
        elif b == 0:
            ms_graph = timed(step, steps, 10)
        emit({"config": "configs[2] batched inference sweep", "B_total": B, "n_gpus": world, "shard": f"{b} clips on rank 0",
              "ms_per_step": ms, "clips_per_s": B / (ms * 1e-3), "ms_per_step_cuda_graph": ms_graph,
              "clips_per_s_cuda_graph": None if ms_graph is None else B / (ms_graph * 1e-3),
              "includes": "forward + logits all-gather" if world > 1 else "forward",
              "l2": "two input sets rotated" if nsets == 2 else "inputs fit L2 (latency point)", "dtype": "bf16"})
        del sets
        B *= 2


def run_train():
    B = 1024
    sets = make_inputs(B, 2, 1 + rank)
    for prec in ("fp32", "tf32"):
        torch.manual_seed(0)
        head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True, precision=prec).to(dev).train()
        tr = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4)
        n0 = _lib.launch_count()
        ms = timed(lambda i: tr.step(*sets[i % 2], training=True), 20, 5)
        launches = (_lib.launch_count() - n0) // 25
        # the same step without the collective, to show what the all-reduce costs
        ms_local = timed(lambda i: (tr.forward_backward(*sets[i % 2], training=True), tr.optimizer_step()), 20, 3)
        emit({"config": "configs[3] Stage-1 training step", "precision": prec, "batch_per_gpu": B, "n_gpus": world, "ms_per_step": ms,
              "clips_per_s": B * world / (ms * 1e-3), "ms_per_step_without_allreduce": ms_local, "kernel_launches_per_step": launches,
              "step": "forward(train, dropouts 0.1/0.1/0.2/0.1) + CE + backward + NCCL all-reduce of the 398,581-float bucket + Adam",
              "dtype": "fp32 SIMT kernels" if prec == "fp32" else "fp32 tensors, Linears on tcgen05 (tf32 products; bf16 first layer), fp32 accumulate",
              "tflops": 150_416_000 * B / (ms * 1e-3) / 1e12})


def run_flat():
    B = 4096
    sets = make_inputs(B, 2, 1 + rank)
    pooled = [(xa.float().mean(1).bfloat16(), xv.float().mean(1).bfloat16()) for xv, xa, _ in sets]
    for mode in ("concat", "gated"):
        torch.manual_seed(0)
        head = FlatFusionHead(num_classes=NCLS, mode=mode, fusion_align_mode="clip").to(dev).eval()
        with torch.no_grad():
            ms_seq = timed(lambda i: head(sets[i % 2][0], sets[i % 2][1]), 30, 5)
            ms_pool = timed(lambda i: head.flat_forward_pooled(*pooled[i % 2]), 100, 10)
            ms_align = timed(lambda i: head.alignment_loss(sets[i % 2][0], sets[i % 2][1]), 20, 3)
        gbs = B * (T * DV + TA * DA) * 2 / (ms_seq * 1e-3) / 1e9
        emit({"config": "configs[4] flat heads with CLIP-style alignment projection", "mode": mode, "B": B, "n_gpus": world,
              "sequence_inputs": {"ms_per_step": ms_seq, "clips_per_s": B * world / (ms_seq * 1e-3), "input_gbs_per_gpu": gbs,
                                  "note": "temporal mean fused in; HBM-bound (245,248 B/clip)"},
              "pooled_inputs": {"ms_per_step": ms_pool, "clips_per_s": B * world / (ms_pool * 1e-3), "note": "latency-bound"},
              "alignment_loss": {"ms_per_call": ms_align, "note": "B x B CLIP-style loss (training-only output, reported separately): "
                                 "temporal means + 2 alignment projections + 2 B x B cosine products on tcgen05 (tf32) + row passes; "
                                 f"{2 * 2 * B * B * 256 / (ms_align * 1e-3) / 1e12:.1f} TFLOP/s on the B x B products alone"},
              "dtype": "bf16 embeddings and operands (tcgen05 kind::f16), fp32 accumulate"})


if __name__ == "__main__":
    what = sys.argv[1:] or ["sweep", "train", "flat"]
    if "sweep" in what:
        run_sweep()
    if "train" in what:
        run_train()
    if "flat" in what:
        run_flat()
    if world > 1:
        dist.destroy_process_group()
