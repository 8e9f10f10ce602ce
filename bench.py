#!/usr/bin/env python
"""bench.py — xattn fusion clips/sec on 1..8 B200 (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload forward|train]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic clips: the xattn fusion head forward (emotion prior
on, concat head, T=16, Ta=149, bf16 embeddings) at B=4096 clips per GPU — BASELINE.json configs[1]; with N>1 each rank
owns its own B=4096 shard (weak scaling, no data-path collective).

* ``value``  : whole-job clips/s with the inputs already resident in HBM (CUDA events, max over ranks).
* ``e2e``    : the same metric through the public API with HOST (pinned) buffers, H2D of the embeddings and D2H of the
               logits inside the timed region (``XAttnFusionHead.predict_host``).
* ``roofline``: algorithmic FLOPs of the timed launches / their CUDA-event duration vs MEASURED_PEAKS.json.
* ``cpu_baseline``: oracle/torch_port.py (the reference's op sequence on stock ATen CPU kernels) on this box's host
               cores over a bounded sample.  ``--impl reference`` times only that arm.

Nothing here reads /root/reference.  oracle/ is executed only for the cpu_baseline / reference arm.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

T, TA, DV, DA, NCLS = 16, 149, 512, 768, 8
FLOP_FWD_UNFOLDED = 60_605_760      # reference op count, prior on, concat head
FLOP_FWD_FOLDED = 55_723_328        # with audio_seq_proj∘a_in_proj folded (what the bf16 path executes)
FLOP_TRAIN = 150_416_000            # fwd + bwd, first-layer dX excluded
BYTES_FWD_BF16 = T * DV * 2 + TA * DA * 2 + NCLS * 4   # 245,280 B/clip
E_, H_, HW_ = 128, 4, 256


def stage_model():
    """Algorithmic bytes and executed FLOPs per clip of each kernel of the bf16 forward.
    Bytes count every tensor the kernel must read or write once (weights, a few hundred KB per launch, excluded)."""
    tee = 2 * T * E_ * E_                      # one [T,E]x[E,E] projection
    qt = 4 * T * E_ * 2                        # one folded per-clip operand [H*T, E] bf16
    return {
        "gemm_v0": {"bytes": T * DV * 2 + T * E_ * 4, "flops": 2 * T * DV * E_},
        "gemm_a0": {"bytes": TA * DA * 2 + TA * E_ * 2 + TA * 8 + 6 * E_ * 4, "flops": 2 * TA * DA * E_ + 4 * TA * E_},
        "video_pre": {"bytes": T * E_ * 4 + qt, "flops": 2 * tee},
        "prior": {"bytes": 4 * E_ * 4 + E_ * 4 + 16, "flops": 2 * (2 * E_ * 64 + 64 * 8)},
        "v2a_attn": {"bytes": TA * E_ * 2 + 2 * qt + TA * 4, "flops": 2 * (2 * H_ * T * TA * E_)},
        "video_update": {"bytes": 3 * qt + T * E_ * 4 + E_ * 4, "flops": 7 * tee},
        "a2v_attn": {"bytes": TA * E_ * 2 + 2 * qt + TA * 4 + 6 * E_ * 4, "flops": 2 * (2 * TA * H_ * T * E_)},
        "head": {"bytes": E_ * 4 + 6 * E_ * 4 + NCLS * 4, "flops": 2 * (2 * E_ * HW_ + HW_ * NCLS)},
    }


# dram__bytes_read.sum + dram__bytes_write.sum per launch at B=4096 from the committed `ncu --set full` capture
# (profiles/r01_ncu_full_bf16_v4.csv); reported as roofline.traffic only when the bench runs that batch.
NCU_TRAFFIC_B4096 = {"gemm_v0": 71.3e6, "gemm_a0": 1094.4e6, "video_pre": 52.1e6, "prior": 14.7e6, "v2a_attn": 270.3e6,
                     "a2v_attn": 306.1e6, "head": 19.0e6}

METRIC = "xattn fusion clips/sec at 1/2/4/8 B200; % of tensor/HBM roofline vs CPU ref"


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU during the timed region (NVML; nvidia-smi as a backstop)."""

    def __init__(self, index: int) -> None:
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    _REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting",
                0x10: "sync_boost"}

    def _loop(self) -> None:
        n = self._nvml
        while not self._stop.is_set():
            try:
                self.samples.append(int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
                r = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h)) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                for bit, name in self._REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.004)

    def start(self) -> None:
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
        if not self.samples:   # backstop: one nvidia-smi query
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                a, b = [int(x) for x in out.strip().split(",")]
                self.samples, self.max_mhz = [a], b
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's op sequence on the host cores (oracle/torch_port.py)
# ---------------------------------------------------------------------------------------------------------------------
def cpu_reference_clips_per_s(sample_b: int, steps: int, warmup: int, budget_s: float, train: bool = False):
    import torch

    from oracle import torch_port
    from multimodalemotionrecognition_b200 import XAttnFusionHead

    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True)   # parameters only (CPU)
    sd = {k: v.detach().clone() for k, v in head.state_dict().items()}
    g = torch.Generator().manual_seed(1)
    xv = torch.randn(sample_b, T, DV, generator=g).bfloat16().float()
    xa = torch.randn(sample_b, TA, DA, generator=g).bfloat16().float()
    labels = torch.randint(0, NCLS, (sample_b,), generator=g)
    if train:
        for k in sd:
            sd[k].requires_grad_(True)
        opt = torch.optim.Adam([p for k, p in sd.items() if not k.startswith("audio_time_conv.")], lr=1e-3, weight_decay=1e-4)
        run = lambda: torch_port.stage1_step(sd, opt, xv, xa, labels, num_heads=4, training=True)
    else:
        def run():
            with torch.inference_mode():
                return torch_port.xattn_forward(sd, xv, xa, num_heads=4, head="concat", use_prior=True)
    for _ in range(max(1, warmup)):
        run()
    times = []
    t_begin = time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s:
            break
    total = sum(times)
    return sample_b * len(times) / total, cores, len(times), total / len(times) * 1e3


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    train = args.workload == "train"
    sample_b = 256 if train else 512
    v, cores, n, ms = cpu_reference_clips_per_s(sample_b, args.steps, min(args.warmup, 3), budget_s=150.0, train=train)
    sample = (f"{n} steps x B={sample_b} clips (bounded sample of the B={args.batch}/GPU workload), fp32, "
              f"torch CPU {cores} threads, oracle/torch_port.py")
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "clips/s", "n_gpus": args.gpus, "steps": n,
        "warmup": min(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args),
        "cpu_baseline": {"value": v, "unit": "clips/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, batch=None) -> dict:
    b = args.batch if batch is None else batch
    if args.workload == "train":
        name = (f"Stage-1 training step (fwd+CE+bwd+allreduce+Adam) of the xattn fusion head, B={b}/GPU, T={T}, Ta={TA}, "
                "prior on, concat head (BASELINE configs[3])")
    else:
        name = (f"xattn fusion forward, emotion-prior bias on, concat head, B={b}/GPU, video [B,{T},{DV}] + audio "
                f"[B,{TA},{DA}] bf16, {NCLS} classes (BASELINE configs[1])")
    return {"workload": name, "batch_per_gpu": b, "T": T, "Ta": TA, "precision": args.precision,
            "l2": "inputs (1.0 GB/step at B=4096) exceed the 126 MB L2; 2 input sets rotated between steps",
            "parallelism": f"batch-sharded x{args.gpus}, weights replicated"}


# ---------------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------------
def run_b200_arm(args) -> None:
    import torch
    import torch.distributed as dist

    from multimodalemotionrecognition_b200 import XAttnFusionHead, _lib
    from multimodalemotionrecognition_b200.hostmem import bind_to_device_numa_node
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (one process per GPU)")
        args.gpus = world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # one rank per GPU: keep the rank (and the pinned e2e buffers it allocates) on the GPU's own socket
    numa_node = bind_to_device_numa_node(local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()
    B = args.batch
    train = args.workload == "train"

    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True,
                           precision=args.precision).to(dev)
    head.train(train)
    nsets = 2
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    sets = []
    for _ in range(nsets):
        xv = torch.randn(B, T, DV, device=dev, generator=g).bfloat16()
        xa = torch.randn(B, TA, DA, device=dev, generator=g).bfloat16()
        sets.append((xv, xa, torch.randint(0, NCLS, (B,), device=dev, generator=g)))
    trainer = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4) if train else None

    def step(i: int):
        xv, xa, lab = sets[i % nsets]
        if train:
            return trainer.step(xv, xa, lab, training=True)
        with torch.no_grad():
            return head(xv, xa)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(max(args.warmup, 3)):
        step(i)
    barrier()
    sampler = ClockSampler(local)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count()
    sampler.start()
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    barrier()
    clocks = sampler.stop()
    launches = _lib.launch_count() - n0
    ms_total = ev0.elapsed_time(ev1)
    tt = torch.tensor([ms_total], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_step = float(tt.item()) / args.steps
    value = B * world / (ms_step * 1e-3)

    # ---- roofline ---------------------------------------------------------------------------------------------------
    # The dominant kernel of the step, timed live: a second pass of the same K steps with the library's per-kernel CUDA-event
    # brackets on (xattn_stage_timing; events on the launching stream).  `value` above comes from the unbracketed pass.
    ms_local = ms_total / args.steps
    folded = args.precision == "bf16"
    flop_clip = FLOP_TRAIN if train else (FLOP_FWD_FOLDED if folded else FLOP_FWD_UNFOLDED)
    tflops = flop_clip * B / (ms_local * 1e-3) / 1e12
    gbs = BYTES_FWD_BF16 * B / (ms_local * 1e-3) / 1e9
    whole = {"kernel": "whole forward (all launches of one step)" if not train else "whole Stage-1 step",
             "tflops": tflops, "tensor_frac": tflops / peaks["bf16_tflops_sustained"], "flop_per_clip": flop_clip,
             "gbs": gbs, "hbm_frac": gbs / peaks["hbm_gbs"], "bytes_per_clip": BYTES_FWD_BF16,
             "peaks": f"sustained bf16 + hbm, {peaks['source']}"}
    stages = None
    if folded and not train:
        _lib.stage_timing(True)
        sev0, sev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sev0.record()
        for i in range(args.steps):
            step(i)
        sev1.record()
        torch.cuda.synchronize(dev)
        rec = _lib.stage_timing_read()
        _lib.stage_timing(False)
        model = stage_model()
        tot = sum(ms for ms, _ in rec.values()) or 1.0
        stages = []
        for name, (ms, n) in rec.items():
            if n == 0:
                continue
            us = ms / n * 1e3
            m = model[name]
            g_ = m["bytes"] * B / (us * 1e-6) / 1e9
            t_ = m["flops"] * B / (us * 1e-6) / 1e12
            stages.append({"kernel": name, "us": us, "share": ms / tot, "launches": n, "bytes_per_clip": m["bytes"],
                           "flops_per_clip": m["flops"], "gbs": g_, "hbm_frac": g_ / peaks["hbm_gbs"], "tflops": t_,
                           "tensor_frac": t_ / peaks["bf16_tflops"]})
        whole["bracketed_pass_ms_per_step"] = sev0.elapsed_time(sev1) / args.steps
    if stages:
        top = max(stages, key=lambda d: d["us"])
        hbm_bound = top["hbm_frac"] >= top["tensor_frac"]
        roofline = {
            "bound": "hbm" if hbm_bound else "tensor",
            "achieved": top["gbs"] if hbm_bound else top["tflops"],
            "peak": peaks["hbm_gbs"] if hbm_bound else peaks["bf16_tflops"],
            "unit": "GB/s" if hbm_bound else "TFLOP/s",
            "frac": top["hbm_frac"] if hbm_bound else top["tensor_frac"],
            "traffic": NCU_TRAFFIC_B4096.get(top["kernel"]) if B == 4096 else None,
            "kernel": top["kernel"], "us_per_launch": top["us"], "share_of_step": top["share"],
            "algorithmic_bytes_per_launch": top["bytes_per_clip"] * B, "algorithmic_flops_per_launch": top["flops_per_clip"] * B,
            "peak_kind": f"burst (kernel timed alone between events), {peaks['source']}",
            "stages": stages, "whole_step": whole,
        }
    else:
        roofline = {
            "bound": "tensor", "achieved": tflops, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
            "frac": tflops / peaks["bf16_tflops_sustained"], "traffic": None, "kernel": whole["kernel"], "whole_step": whole,
        }

    # ---- e2e: host (pinned) buffers through the public API -------------------------------------------------------------
    e2e = None
    if not train:
        hv = torch.empty((B, T, DV), dtype=torch.bfloat16).pin_memory()
        ha = torch.empty((B, TA, DA), dtype=torch.bfloat16).pin_memory()
        hv.copy_(sets[0][0])
        ha.copy_(sets[0][1])
        e2e_steps = max(3, min(args.steps, 10))
        for _ in range(2):
            head.predict_host(hv, ha)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            out = head.predict_host(hv, ha)    # returns host logits; synchronises on its D2H
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": B * world * e2e_steps / float(dt.item()), "unit": "clips/s",
               "h2d_bytes_per_step": (hv.numel() + ha.numel()) * 2 * world, "d2h_bytes_per_step": out.numel() * 4 * world,
               "steps": e2e_steps, "api": "XAttnFusionHead.predict_host (pinned host in, host logits out)",
               "numa_node_rank0": numa_node}

    # ---- CPU baseline (rank 0, N=1 only) ----------------------------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sb = 256 if train else 512
        v, cores, n, _ = cpu_reference_clips_per_s(sb, steps=40, warmup=1, budget_s=15.0, train=train)
        cpu = {"value": v, "unit": "clips/s", "cores": cores, "kind": "port",
               "sample": f"{n} x B={sb} clips of the same workload, fp32, torch CPU {cores} threads (oracle/torch_port.py)"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"bf16": "bf16", "fp32": "f32", "tf32": "tf32"}[args.precision], "data": "synthetic",
            "config": workload_config(args), "clocks": clocks, "gpu_launches": int(launches),
            "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=["forward", "train"], default="forward")
    ap.add_argument("--batch", type=int, default=None, help="clips per GPU (default 4096 forward, 1024 train)")
    ap.add_argument("--precision", choices=["bf16", "fp32", "tf32"], default=None,
                    help="forward: bf16 (default) | fp32; train: tf32 (default: fp32 tensors, Linears on tcgen05) | fp32 (SIMT parity path)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.batch is None:
        args.batch = 1024 if args.workload == "train" else 4096
    if args.precision is None:
        args.precision = "tf32" if args.workload == "train" else "bf16"
    if args.workload == "train" and args.precision == "bf16":
        raise SystemExit("the bf16 tcgen05 path is inference-only; train with --precision tf32 or fp32")
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
