#!/usr/bin/env python
"""bench.py — xattn fusion clips/sec on 1..8 B200 (BASELINE.json metric), ONE JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload forward|train|sweep|flat]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path over one batch of synthetic clips.  Workloads (BASELINE.json configs):
  forward (default, configs[1]) : xattn fusion head forward, emotion prior on, concat head, T=16, Ta=149, bf16 embeddings,
                                  B=4096 clips per GPU; N>1: each rank owns its own shard (weak scaling, no data-path collective)
  train   (configs[3])          : Stage-1 step (forward + CE + backward + NCCL gradient all-reduce + Adam), B=1024 per GPU
  sweep   (configs[2])          : batched inference B=1..16384 sharded over the ranks, logits all-gather inside the timed region
  flat    (configs[4])          : concat / gated flat heads with the CLIP-style alignment projection, B=4096, forward

Keys of the line:
* ``value``       whole-job clips/s with the inputs already resident in HBM (CUDA events on the launching stream, max over ranks).
* ``e2e``         the same metric through the public API with HOST (pinned) buffers: H2D of the step's inputs and D2H of its result
                  inside the timed region (``XAttnFusionHead.predict_host``; for ``train``: embeddings + labels in, loss out).
* ``roofline``    ``frac`` is the WHOLE-STEP fraction: SURVEY 8d algorithmic bytes (or FLOPs) of one step / its device time vs
                  MEASURED_PEAKS.json.  ``dominant_kernel`` describes the slowest kernel (timed live with the library's CUDA-event
                  brackets in a second pass), ``stages`` every kernel; ``traffic`` = DRAM bytes of one step from the committed
                  ncu capture named in ``traffic_source`` (null when the batch differs from the captured one).
* ``cpu_baseline`` the reference's own module (``oracle/_ref``, staged unmodified by ``oracle/stage_ref.py``; ``kind:
                  "reference"``) or, when absent, the port ``oracle/torch_port.py`` (``kind: "port"``) on this box's host cores.
                  ``--impl reference`` times only that arm.  ``parity`` (forward) compares the timed GPU path with it.

Nothing here reads /root/reference.  oracle/ is executed only for the cpu_baseline / reference arm (timing + checker).
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
FLOP_FWD_REFERENCE = 60_605_760     # the reference's op count, prior on, concat head
FLOP_TRAIN = 150_416_000            # fwd + bwd, first-layer dX excluded
BYTES_FWD_BF16 = T * DV * 2 + TA * DA * 2 + NCLS * 4   # 245,280 B/clip: inputs once + fp32 logits
BYTES_TRAIN_BF16 = 2 * (T * DV * 2 + TA * DA * 2) + NCLS * 4 + 8   # inputs read twice (forward, first-layer dW) + labels
E_, H_, HW_ = 128, 4, 256
TRAFFIC_FILE = ROOT / "profiles" / "r02_ncu_traffic_bf16.json"   # written from an `ncu --set full` capture (profiles/README.md)
TRAFFIC_FILE_TRAIN = ROOT / "profiles" / "r02_ncu_traffic_train_tf32.json"   # scripts/ncu_traffic_train.py (one metrics pass)
METRIC = "xattn fusion clips/sec at 1/2/4/8 B200; % of tensor/HBM roofline vs CPU ref"


def stage_model():
    """Algorithmic bytes and EXECUTED FLOPs per clip of each kernel of the bf16 forward.  Bytes: every tensor
    the kernel must read or write once (weights, a few hundred KB per launch, excluded).  The FLOPs are the products the
    kernels run after the algebra of DESIGN 4.1 (sum 46.1 M/clip, against the reference's 60.6 M)."""
    tee = 2 * T * E_ * E_                      # one [T,E]x[E,E] projection
    qt = 4 * T * E_ * 2                        # one folded per-clip operand [H*T, E] bf16
    return {
        "gemm_v0": {"bytes": T * DV * 2 + T * E_ * 4, "flops": 2 * T * DV * E_},
        "gemm_a0": {"bytes": TA * DA * 2 + TA * E_ * 2 + TA * 8 + 6 * E_ * 4, "flops": 2 * TA * DA * E_ + 4 * TA * E_,
                    "algorithmic_bytes": TA * DA * 2},   # SURVEY 8d counts only the audio embeddings it reads
        "video_pre": {"bytes": T * E_ * 4 + qt, "flops": 2 * tee},
        "prior": {"bytes": 4 * E_ * 4 + E_ * 4 + 16, "flops": 2 * (2 * E_ * 64 + 64 * 8)},
        "v2a_attn": {"bytes": TA * E_ * 2 + 2 * qt + TA * 4, "flops": 2 * (2 * H_ * T * TA * E_)},
        "video_update": {"bytes": 3 * qt + T * E_ * 4 + E_ * 4, "flops": 7 * tee},
        "a2v_attn": {"bytes": TA * E_ * 2 + 2 * qt + TA * 4 + 6 * E_ * 4, "flops": 2 * (2 * TA * H_ * T * E_)},
        "head": {"bytes": E_ * 4 + 6 * E_ * 4 + NCLS * 4, "flops": 2 * (2 * E_ * HW_ + HW_ * NCLS)},
    }


FLOP_FWD_EXECUTED = sum(v["flops"] for v in stage_model().values())


def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def load_traffic(batch: int):
    """Per-kernel and per-step DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) from the committed ncu capture."""
    if not TRAFFIC_FILE.exists():
        return None
    d = json.loads(TRAFFIC_FILE.read_text())
    return d if int(d.get("batch", -1)) == batch else None


def load_traffic_train(batch: int, precision: str):
    """DRAM bytes of one Stage-1 step (every launch between two Adam kernels) from the committed ncu metrics pass."""
    if not TRAFFIC_FILE_TRAIN.exists():
        return None
    d = json.loads(TRAFFIC_FILE_TRAIN.read_text())
    return d if int(d.get("batch", -1)) == batch and d.get("precision") == precision else None


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
# reference arm / cpu baseline: the reference's own module on the host cores (oracle/_ref), else the port
# ---------------------------------------------------------------------------------------------------------------------
class CpuReference:
    """The reference's xattn path on this box's host cores with all the threads it can use.  ``kind`` "reference": the
    unmodified ``FusionModel`` staged in oracle/_ref (identity-stub backbones); "port": oracle/torch_port.py (same ATen ops)."""

    def __init__(self) -> None:
        import torch

        from oracle import ref_runner, torch_port
        from multimodalemotionrecognition_b200 import XAttnFusionHead

        cores = os.cpu_count() or 1
        try:
            cores = len(os.sched_getaffinity(0))
        except Exception:
            pass
        torch.set_num_threads(cores)
        self.cores = cores
        self.torch = torch
        torch.manual_seed(0)
        head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True)   # parameters only (CPU)
        self.sd = {k: v.detach().clone() for k, v in head.state_dict().items()}
        self.port = torch_port
        self.module = None
        if ref_runner.available():
            self.module = ref_runner.build_reference_module(self.sd)
        self.kind = "reference" if self.module is not None else "port"
        self.what = ("the reference's own FusionModel (oracle/_ref/src/models/fusion.py, identity-stub backbones)"
                     if self.module is not None else "oracle/torch_port.py (the reference's op sequence on stock ATen kernels)")
        self._opt = None

    def inputs(self, b: int, seed: int = 1):
        torch = self.torch
        g = torch.Generator().manual_seed(seed)
        xv = torch.randn(b, T, DV, generator=g).bfloat16().float()
        xa = torch.randn(b, TA, DA, generator=g).bfloat16().float()
        return xv, xa, torch.randint(0, NCLS, (b,), generator=g)

    def forward(self, xv, xa, sd=None):
        torch = self.torch
        with torch.inference_mode():
            if self.module is not None and sd is None:
                self.module.eval()
                return self.module(xv[..., None, None], xa)
            return self.port.xattn_forward(sd or self.sd, xv, xa, num_heads=4, head="concat", use_prior=True)

    def train_step(self, xv, xa, labels):
        torch = self.torch
        if self.module is not None:   # train.py:204-228 on the reference module, reference dropout defaults
            if self._opt is None:
                self.module.train()
                params = [p for n, p in self.module.named_parameters() if not n.startswith("audio_time_conv.")]
                self._opt = torch.optim.Adam(params, lr=1e-3, weight_decay=1e-4)
            self._opt.zero_grad()
            loss = torch.nn.functional.cross_entropy(self.module(xv[..., None, None], xa), labels)
            loss.backward()
            self._opt.step()
            return loss.detach()
        if self._opt is None:
            for k in self.sd:
                self.sd[k].requires_grad_(True)
            self._opt = torch.optim.Adam([p for k, p in self.sd.items() if not k.startswith("audio_time_conv.")], lr=1e-3,
                                         weight_decay=1e-4)
        return self.port.stage1_step(self.sd, self._opt, xv, xa, labels, num_heads=4, training=True)

    def time(self, run, steps: int, warmup: int, budget_s: float):
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
        return len(times), sum(times) / len(times)


def cpu_baseline(args, budget_s: float, steps: int, warmup: int):
    """(cpu_baseline dict, CpuReference) for the arm's workload; forward also reports the reference's own B=8 case."""
    ref = CpuReference()
    train = args.workload == "train"
    b = args.batch
    xv, xa, lab = ref.inputs(b)
    run = (lambda: ref.train_step(xv, xa, lab)) if train else (lambda: ref.forward(xv, xa))
    n, sec = ref.time(run, steps, warmup, budget_s)
    out = {"value": b / sec, "unit": "clips/s", "cores": ref.cores, "kind": ref.kind, "ms_per_step": sec * 1e3,
           "sample": f"{n} steps x B={b} clips (the arm's full batch), fp32, torch CPU {ref.cores} threads, {ref.what}"}
    if not train:
        xv8, xa8, _ = ref.inputs(8)
        n8, sec8 = ref.time(lambda: ref.forward(xv8, xa8), 50, 3, 5.0)
        out["b8"] = {"value": 8 / sec8, "unit": "clips/s", "ms_per_step": sec8 * 1e3,
                     "sample": f"{n8} steps x B=8 (BASELINE configs[0], the reference's own CPU-runnable case)"}
    return out, ref


def run_reference_arm(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload in ("sweep", "flat"):
        print(json.dumps({"impl": "reference", "unavailable": f"the reference arm covers forward and train; {args.workload} is GPU-side only"}))
        return
    cb, _ = cpu_baseline(args, budget_s=150.0, steps=args.steps, warmup=min(args.warmup, 3))
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "clips/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 3), "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args),
        "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "clips/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, batch=None) -> dict:
    b = args.batch if batch is None else batch
    names = {
        "train": (f"Stage-1 training step (fwd+CE+bwd+allreduce+Adam) of the xattn fusion head, B={b}/GPU, T={T}, Ta={TA}, "
                  "prior on, concat head, reference dropouts (BASELINE configs[3])"),
        "forward": (f"xattn fusion forward, emotion-prior bias on, concat head, B={b}/GPU, video [B,{T},{DV}] + audio "
                    f"[B,{TA},{DA}] bf16, {NCLS} classes (BASELINE configs[1])"),
        "sweep": "batched inference sweep B=1..16384 of the xattn fusion forward, batch-sharded, logits all-gather timed (BASELINE configs[2])",
        "flat": f"gated and concat flat heads with the CLIP-style alignment projection, forward, B={b}/GPU (BASELINE configs[4])",
    }
    return {"workload": names[args.workload], "batch_per_gpu": b, "T": T, "Ta": TA, "precision": args.precision,
            "l2": "inputs (1.0 GB/step at B=4096; 0.25 GB at B=1024) exceed the 126 MB L2; 2 input sets rotated between steps",
            "parallelism": f"batch-sharded x{args.gpus}, weights replicated"}


# ---------------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------------
class Ctx:
    """Per-process state of the B200 arm: device, ranks, timing helpers."""

    def __init__(self, args) -> None:
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            if self.world == 1 and args.gpus > 1:
                raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (one process per GPU)")
            args.gpus = self.world
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        from multimodalemotionrecognition_b200.hostmem import bind_to_device_numa_node
        # one rank per GPU: keep the rank (and the pinned e2e buffers it allocates) on the GPU's own socket
        self.numa_node = bind_to_device_numa_node(self.local) if self.world > 1 else None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.peaks = load_peaks()

    def barrier(self) -> None:
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x: float) -> float:
        t = self.torch.tensor([x], device=self.dev, dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps: int, warmup: int) -> float:
        """ms per call of fn(i): CUDA events around `steps` calls bracketed by barrier + synchronize, max over ranks."""
        torch = self.torch
        for i in range(warmup):
            fn(i)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1) / steps)

    def finish(self) -> None:
        if self.world > 1:
            self.dist.destroy_process_group()


def make_sets(cx: Ctx, b: int, nsets: int = 2):
    torch = cx.torch
    g = torch.Generator(device=cx.dev).manual_seed(1 + cx.rank)
    sets = []
    for _ in range(nsets):
        xv = torch.randn(b, T, DV, device=cx.dev, generator=g).bfloat16()
        xa = torch.randn(b, TA, DA, device=cx.dev, generator=g).bfloat16()
        sets.append((xv, xa, torch.randint(0, NCLS, (b,), device=cx.dev, generator=g)))
    return sets


def base_line(cx: Ctx, args, value, ms_step, clocks, launches, dtype=None):
    return {"metric": METRIC, "value": value, "unit": "clips/s", "n_gpus": cx.world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": dtype or {"bf16": "bf16", "fp32": "f32", "tf32": "tf32"}[args.precision], "data": "synthetic",
            "config": workload_config(args), "clocks": clocks, "gpu_launches": int(launches)}


def run_forward(cx: Ctx, args) -> None:
    torch = cx.torch
    from multimodalemotionrecognition_b200 import XAttnFusionHead, _lib

    peaks, B, dev = cx.peaks, args.batch, cx.dev
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True, precision=args.precision).to(dev).eval()
    sets = make_sets(cx, B)

    def step(i):
        with torch.no_grad():
            return head(sets[i % 2][0], sets[i % 2][1])

    for i in range(max(args.warmup, 3)):
        step(i)
    cx.barrier()
    sampler = ClockSampler(cx.local)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count()
    sampler.start()
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    cx.barrier()
    clocks = sampler.stop()
    launches = _lib.launch_count() - n0
    ms_local = ev0.elapsed_time(ev1) / args.steps
    ms_step = cx.max_over_ranks(ms_local)
    value = B * cx.world / (ms_step * 1e-3)

    # ---- roofline: the whole step against SURVEY 8d's algorithmic bytes / FLOPs; per-kernel brackets in a second pass -------
    folded = args.precision == "bf16"
    flops_exec = FLOP_FWD_EXECUTED if folded else FLOP_FWD_REFERENCE
    gbs = BYTES_FWD_BF16 * B / (ms_local * 1e-3) / 1e9
    tf_exec = flops_exec * B / (ms_local * 1e-3) / 1e12
    traffic = load_traffic(B) if folded else None
    roofline = {
        "bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"],
        "traffic": None if traffic is None else traffic["step_total_bytes"],
        "traffic_source": None if traffic is None else traffic["source"],
        "kernel": "whole forward (every launch of one step)", "algorithmic_bytes_per_clip": BYTES_FWD_BF16,
        "algorithmic_bytes_per_step": BYTES_FWD_BF16 * B, "peak_kind": f"measured copy bandwidth ({peaks['source']})",
        "why_hbm": "roofline = slower of FLOPs / bf16 peak and bytes / HBM: 245,280 B/clip is 37.5 ns of HBM, the executed "
                   f"{flops_exec / 1e6:.1f} MFLOP/clip {flops_exec / peaks['bf16_tflops'] / 1e3:.1f} ns of tensor time",
        "tensor": {"executed_flops_per_clip": flops_exec, "reference_flops_per_clip": FLOP_FWD_REFERENCE, "tflops_executed": tf_exec,
                   "frac_of_sustained_peak": tf_exec / peaks["bf16_tflops_sustained"], "frac_of_burst_peak": tf_exec / peaks["bf16_tflops"]},
    }
    if folded:
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
            st = {"kernel": name, "us": us, "share": ms / tot, "launches": n, "own_io_bytes_per_clip": m["bytes"],
                  "flops_per_clip": m["flops"], "gbs_own_io": g_, "hbm_frac_own_io": g_ / peaks["hbm_gbs"], "tflops": t_,
                  "tensor_frac_burst": t_ / peaks["bf16_tflops"]}
            if traffic is not None and name in traffic.get("kernels", {}):
                st["dram_bytes_ncu"] = traffic["kernels"][name]
            stages.append(st)
        top = max(stages, key=lambda d: d["us"])
        alg = model[top["kernel"]].get("algorithmic_bytes", model[top["kernel"]]["bytes"])
        roofline["dominant_kernel"] = {
            "kernel": top["kernel"], "us_per_launch": top["us"], "share_of_step": top["share"], "bound": "hbm",
            "achieved_gbs_algorithmic": alg * B / (top["us"] * 1e-6) / 1e9,
            "frac_algorithmic": alg * B / (top["us"] * 1e-6) / 1e9 / peaks["hbm_gbs"],
            "algorithmic_bytes_per_launch": alg * B, "achieved_gbs_own_io": top["gbs_own_io"], "frac_own_io": top["hbm_frac_own_io"],
            "traffic": top.get("dram_bytes_ncu"),
            "note": "algorithmic = SURVEY 8d bytes this kernel is the reader of (the audio embeddings); own_io adds the a0 rows it writes",
        }
        roofline["stages"] = stages
        roofline["bracketed_pass_ms_per_step"] = sev0.elapsed_time(sev1) / args.steps

    # ---- e2e: host (pinned) buffers through the public API ----------------------------------------------------------
    e2e_all = {}
    for eb in ([B] if cx.world > 1 else [B, 256, 8]):
        hv = torch.empty((eb, T, DV), dtype=torch.bfloat16).pin_memory()
        ha = torch.empty((eb, TA, DA), dtype=torch.bfloat16).pin_memory()
        hv.copy_(sets[0][0][:eb])
        ha.copy_(sets[0][1][:eb])
        e2e_steps = max(3, min(args.steps, 10)) if eb == B else 50
        for _ in range(3):
            head.predict_host(hv, ha)
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            out = head.predict_host(hv, ha)    # returns host logits; synchronises on its D2H
        cx.barrier()
        dt = cx.max_over_ranks(time.perf_counter() - t0)
        e2e_all[eb] = {"value": eb * cx.world * e2e_steps / dt, "unit": "clips/s", "ms_per_call": dt / e2e_steps * 1e3,
                       "h2d_bytes_per_step": (hv.numel() + ha.numel()) * 2 * cx.world, "d2h_bytes_per_step": out.numel() * 4 * cx.world,
                       "steps": e2e_steps}
        if eb != B:   # the same with two batches in flight (predict_host_async): batch i+1 is issued before batch i is collected
            cx.barrier()
            t0 = time.perf_counter()
            prev = None
            for _ in range(e2e_steps):
                cur = head.predict_host_async(hv, ha)
                if prev is not None:
                    prev.result()
                prev = cur
            prev.result()
            cx.barrier()
            dt2 = cx.max_over_ranks(time.perf_counter() - t0)
            e2e_all[eb]["two_in_flight"] = {"value": eb * cx.world * e2e_steps / dt2, "ms_per_call": dt2 / e2e_steps * 1e3}
    e2e = dict(e2e_all[B], api="XAttnFusionHead.predict_host (pinned host in, host logits out)", numa_node_rank0=cx.numa_node)
    if len(e2e_all) > 1:   # the worker batch sizes of the reference's serving path (inference_worker.py:33-34), latency view
        e2e["worker_batches"] = {f"B={k}": v for k, v in e2e_all.items() if k != B}

    # ---- CPU baseline + parity of the timed path against it (rank 0, N=1 only) -------------------------------------
    cpu = parity = None
    if cx.rank == 0 and cx.world == 1 and not args.no_cpu_baseline:
        cpu, ref = cpu_baseline(args, budget_s=20.0, steps=40, warmup=1)
        if folded:
            # the quantisation-matched oracle: the fp32 reference arithmetic on the same bf16-rounded inputs and weights
            xv, xa = sets[0][0], sets[0][1]
            with torch.no_grad():
                got = head(xv, xa).cpu()
            sdq = {k: (v.bfloat16().float() if v.ndim >= 2 else v) for k, v in ref.sd.items()}
            want = ref.forward(xv.float().cpu(), xa.float().cpu(), sd=sdq)
            err = (got - want).abs().max().item()
            top2 = want.topk(2, dim=1).values
            margin = top2[:, 0] - top2[:, 1]
            agree = (got.argmax(1) == want.argmax(1)).float()
            parity = {"oracle": "fp32 port of the reference on the same bf16-rounded inputs and weights (oracle/torch_port.py)",
                      "clips": B, "max_abs_logit_err": err, "raw_top1_agreement": agree.mean().item(),
                      "near_tie_fraction_margin_lt_2err": (margin < 2 * err).float().mean().item(),
                      "largest_margin_among_disagreements": margin[agree == 0].max().item() if (agree == 0).any() else 0.0,
                      "note": "random-init weights, iid N(0,1) inputs: ~1-2 % of the clips have a top1-top2 margin below 1e-3 by "
                              "construction; the trained-fixture gate (>= 99.9 % raw) is tests/test_gpu_bf16.py"}
    if cx.rank == 0:
        line = base_line(cx, args, value, ms_step, clocks, launches)
        line.update({"roofline": roofline, "e2e": e2e, "cpu_baseline": cpu, "parity": parity})
        print(json.dumps(line), flush=True)


def run_train(cx: Ctx, args) -> None:
    torch, dist = cx.torch, cx.dist
    from multimodalemotionrecognition_b200 import XAttnFusionHead, _lib
    from multimodalemotionrecognition_b200.fusion import _ptr, _stream_ptr
    from multimodalemotionrecognition_b200.training import Stage1Trainer

    peaks, B, dev = cx.peaks, args.batch, cx.dev
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True, precision=args.precision).to(dev).train()
    tr = Stage1Trainer(head, lr=1e-3, weight_decay=1e-4, overlap_allreduce=not args.no_overlap)
    sets = make_sets(cx, B)

    def step(i):
        return tr.step(*sets[i % 2], training=True)

    for i in range(max(args.warmup, 3)):
        step(i)
    cx.barrier()
    sampler = ClockSampler(cx.local)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = _lib.launch_count()
    sampler.start()
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    cx.barrier()
    clocks = sampler.stop()
    launches = _lib.launch_count() - n0
    ms_local = ev0.elapsed_time(ev1) / args.steps
    ms_step = cx.max_over_ranks(ms_local)
    value = B * cx.world / (ms_step * 1e-3)

    # phases of the step, bracketed with CUDA events on the launching stream in a second pass (the collective is synchronous here)
    names = ["forward", "ce_loss", "backward", "allreduce", "adam"]
    acc = {n: 0.0 for n in names}
    evs = []
    for i in range(args.steps):
        xv, xa, lab = sets[i % 2]
        e = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        e[0].record()
        logits, state = head._launch_forward(xv, xa, save=True, training=True)
        e[1].record()
        grads = head.flat_gradients(attach=True)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dlogits = torch.empty_like(logits)
        _lib.check(_lib.load().xattn_ce_loss_fwd_bwd(_ptr(logits), _ptr(lab), B, NCLS, 0.0, 1.0 / cx.world, _ptr(loss), _ptr(dlogits),
                                                     _stream_ptr(dev)))
        e[2].record()
        head._launch_backward(state, dlogits, grads)
        e[3].record()
        if cx.world > 1:
            dist.all_reduce(grads)
        e[4].record()
        tr.optimizer_step()
        e[5].record()
        evs.append(e)
    torch.cuda.synchronize(dev)
    for e in evs:
        for k, n in enumerate(names):
            acc[n] += e[k].elapsed_time(e[k + 1])
    tot = sum(acc.values()) or 1.0
    stages = [{"phase": n, "us": acc[n] / args.steps * 1e3, "share": acc[n] / tot} for n in names]
    tflops = FLOP_TRAIN * B / (ms_local * 1e-3) / 1e12
    traffic = load_traffic_train(B, args.precision)
    gbs = BYTES_TRAIN_BF16 * B / (ms_local * 1e-3) / 1e9
    roofline = {
        "bound": "tensor", "achieved": tflops, "peak": peaks["bf16_tflops_sustained"], "unit": "TFLOP/s",
        "frac": tflops / peaks["bf16_tflops_sustained"], "traffic": None if traffic is None else traffic["step_total_bytes"],
        "traffic_source": None if traffic is None else traffic["source"], "kernel": "whole Stage-1 step (every launch)",
        "algorithmic_flops_per_clip": FLOP_TRAIN, "peak_kind": f"sustained bf16 ({peaks['source']}); the Linears run as tf32 products "
        "(half the bf16 rate), the first layer as bf16",
        "hbm": {"algorithmic_bytes_per_clip": BYTES_TRAIN_BF16, "gbs": gbs, "frac": gbs / peaks["hbm_gbs"],
                # the same step on the DRAM bytes it actually moves (fp32 activations are stored and re-read for the backward)
                "gbs_on_traffic": None if traffic is None else traffic["step_total_bytes"] / (ms_local * 1e-3) / 1e9,
                "frac_on_traffic": None if traffic is None else traffic["step_total_bytes"] / (ms_local * 1e-3) / 1e9 / peaks["hbm_gbs"]},
        "stages": stages, "unbracketed_ms_per_step": ms_local,
        "note": "per-kernel times of the step: profiles/ ncu launch list (README); phases here are CUDA-event brackets of a second pass "
                "through the SEPARATE calls (eager launches, synchronous all-reduce): they sum to more than the graphed step",
    }
    # e2e: pinned host embeddings + labels in, loss out
    hv = torch.empty((B, T, DV), dtype=torch.bfloat16).pin_memory()
    ha = torch.empty((B, TA, DA), dtype=torch.bfloat16).pin_memory()
    hl = torch.empty((B,), dtype=torch.int64).pin_memory()
    hv.copy_(sets[0][0]); ha.copy_(sets[0][1]); hl.copy_(sets[0][2])

    from multimodalemotionrecognition_b200.training import HostBatchPrefetcher

    def e2e_step():
        dv = hv.to(dev, non_blocking=True)
        da = ha.to(dev, non_blocking=True)
        dl = hl.to(dev, non_blocking=True)
        return float(tr.step(dv, da, dl, training=True))     # .item(): the loss comes back to the host every step (train.py:230)

    pre = HostBatchPrefetcher(dev)

    def e2e_step_prefetch():   # the copy of the NEXT batch is in flight while this one trains; still one .item() per step
        dv, da, dl = pre.get()
        loss = tr.step(dv, da, dl, training=True)
        pre.release()
        pre.put(hv, ha, hl)
        return float(loss)

    def timed(fn, n):
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        cx.barrier()
        return cx.max_over_ranks(time.perf_counter() - t0)

    e2e_steps = max(3, min(args.steps, 20))
    for _ in range(3):
        e2e_step()
    dt_serial = timed(e2e_step, e2e_steps)
    pre.put(hv, ha, hl)
    for _ in range(3):
        e2e_step_prefetch()
    dt = timed(e2e_step_prefetch, e2e_steps)
    e2e = {"value": B * cx.world * e2e_steps / dt, "unit": "clips/s", "ms_per_step": dt / e2e_steps * 1e3,
           "h2d_bytes_per_step": ((hv.numel() + ha.numel()) * 2 + hl.numel() * 8) * cx.world, "d2h_bytes_per_step": 4 * cx.world,
           "steps": e2e_steps,
           "api": "HostBatchPrefetcher (pinned host -> device copy of the next batch beside the step) + Stage1Trainer.step; loss read "
                  "back with .item() every step",
           "serial_copy_then_step": {"value": B * cx.world * e2e_steps / dt_serial, "ms_per_step": dt_serial / e2e_steps * 1e3}}
    cpu = None
    if cx.rank == 0 and cx.world == 1 and not args.no_cpu_baseline:
        cpu, _ = cpu_baseline(args, budget_s=20.0, steps=20, warmup=1)
    if cx.rank == 0:
        line = base_line(cx, args, value, ms_step, clocks, launches)
        line["collective"] = (None if cx.world == 1 else
                              ("NCCL all-reduce of the flat fp32 gradient bucket (398,581 floats) in two buckets: attention / norm / "
                               "head / prior gradients on a side stream beside the input-projection backward, then the input projections"
                               if not args.no_overlap else "one NCCL all-reduce of the flat fp32 gradient bucket after the backward"))
        line["step_api"] = ("Stage1Trainer.step -> xattn_stage1_step: the step's launches captured every step and replayed as CUDA graph(s) "
                            "(cudaGraphExecUpdate; one graph on a single rank, two around the overlapped all-reduce)"
                            if (cx.world == 1 or not args.no_overlap) else "Stage1Trainer.step, separate calls (--no-overlap)")
        line.update({"roofline": roofline, "e2e": e2e, "cpu_baseline": cpu})
        print(json.dumps(line), flush=True)


def run_sweep(cx: Ctx, args) -> None:
    torch = cx.torch
    from multimodalemotionrecognition_b200 import XAttnFusionHead, _lib
    from multimodalemotionrecognition_b200.training import gather_logits, shard_range

    dev = cx.dev
    torch.manual_seed(0)
    head = XAttnFusionHead(num_classes=NCLS, audio_n_mels=DA, xattn_use_emotion_prior=True, precision="bf16").to(dev).eval()
    points = []
    n0 = _lib.launch_count()
    sampler = ClockSampler(cx.local)
    sampler.start()
    Bt = 1
    while Bt <= 16384:
        s, e = shard_range(Bt, cx.rank, cx.world)
        b = e - s
        nsets = 2 if b * (T * DV + TA * DA) * 2 > 64e6 else 1
        sets = make_sets(cx, b, nsets) if b > 0 else None

        def step(i):
            with torch.no_grad():
                out = head(sets[i % nsets][0], sets[i % nsets][1]) if b > 0 else torch.empty((0, NCLS), dtype=torch.float32, device=dev)
                return gather_logits(out, Bt)

        steps = 100 if Bt <= 1024 else 30
        ms = cx.timed(step, steps, 10)
        ms_graph = None
        # launch-bound sizes: the same forward as one CUDA graph over static buffers (GraphedForward).  The decision must be the
        # same on every rank (each timed call holds a collective): rank 0 owns the largest shard; ranks with an empty shard join
        # the gathers with their eager (empty) step.
        s0, e0 = shard_range(Bt, 0, cx.world)
        if e0 - s0 <= 1024:
            if b > 0:
                gf = head.graphed_forward(b)
                gf.video.copy_(sets[0][0])
                gf.audio.copy_(sets[0][1])
                ms_graph = cx.timed(lambda i: gather_logits(gf(), Bt), steps, 10)
            else:
                ms_graph = cx.timed(step, steps, 10)
        points.append({"B_total": Bt, "clips_rank0": b, "ms": ms, "clips_per_s": Bt / (ms * 1e-3), "ms_cuda_graph": ms_graph,
                       "l2": "two input sets rotated" if nsets == 2 else "inputs fit L2 (latency point)"})
        del sets
        Bt *= 2
    clocks = sampler.stop()
    if cx.rank == 0:
        best = max(points, key=lambda p: p["clips_per_s"])
        args.batch = best["B_total"] // cx.world
        line = base_line(cx, args, best["clips_per_s"], best["ms"], clocks, _lib.launch_count() - n0, dtype="bf16")
        line["steps"], line["scaling"] = 30, "strong"
        line["includes"] = "forward + logits all-gather" if cx.world > 1 else "forward"
        line["value_is"] = f"the best point of the sweep (B_total={best['B_total']})"
        line["sweep"] = points
        line.update({"roofline": None, "e2e": None, "cpu_baseline": None})
        print(json.dumps(line), flush=True)


def run_flat(cx: Ctx, args) -> None:
    torch = cx.torch
    from multimodalemotionrecognition_b200 import FlatFusionHead, _lib

    dev, B, peaks = cx.dev, args.batch, cx.peaks
    sets = make_sets(cx, B)
    pooled = [(xa.float().mean(1).bfloat16(), xv.float().mean(1).bfloat16()) for xv, xa, _ in sets]
    n0 = _lib.launch_count()
    sampler = ClockSampler(cx.local)
    sampler.start()
    modes = {}
    for mode in ("concat", "gated"):
        torch.manual_seed(0)
        head = FlatFusionHead(num_classes=NCLS, mode=mode, fusion_align_mode="clip").to(dev).eval()
        with torch.no_grad():
            ms_seq = cx.timed(lambda i: head(sets[i % 2][0], sets[i % 2][1]), args.steps, max(args.warmup, 3))
            ms_pool = cx.timed(lambda i: head.flat_forward_pooled(*pooled[i % 2]), 100, 10)
            ms_align = cx.timed(lambda i: head.alignment_loss(sets[i % 2][0], sets[i % 2][1]), 20, 3)
        gbs = B * (T * DV + TA * DA) * 2 / (ms_seq * 1e-3) / 1e9
        modes[mode] = {"sequence_inputs": {"ms_per_step": ms_seq, "clips_per_s": B * cx.world / (ms_seq * 1e-3), "input_gbs_per_gpu": gbs,
                                           "hbm_frac": gbs / peaks["hbm_gbs"]},
                       "pooled_inputs": {"ms_per_step": ms_pool, "clips_per_s": B * cx.world / (ms_pool * 1e-3), "note": "latency-bound"},
                       "alignment_loss_ms": ms_align}
    clocks = sampler.stop()
    if cx.rank == 0:
        c = modes["concat"]["sequence_inputs"]
        line = base_line(cx, args, c["clips_per_s"], c["ms_per_step"], clocks, _lib.launch_count() - n0, dtype="bf16")
        line["value_is"] = "concat flat head from sequence embeddings (temporal mean fused in)"
        line["modes"] = modes
        line["roofline"] = {"bound": "hbm", "achieved": c["input_gbs_per_gpu"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                            "frac": c["hbm_frac"], "traffic": None, "kernel": "whole flat-head forward",
                            "algorithmic_bytes_per_clip": (T * DV + TA * DA) * 2}
        line.update({"e2e": None, "cpu_baseline": None})
        print(json.dumps(line), flush=True)


def run_b200_arm(args) -> None:
    cx = Ctx(args)
    try:
        {"forward": run_forward, "train": run_train, "sweep": run_sweep, "flat": run_flat}[args.workload](cx, args)
    finally:
        cx.finish()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", choices=["forward", "train", "sweep", "flat"], default="forward")
    ap.add_argument("--batch", type=int, default=None, help="clips per GPU (default 4096 forward / flat, 1024 train)")
    ap.add_argument("--precision", choices=["bf16", "fp32", "tf32"], default=None,
                    help="forward: bf16 (default) | fp32; train: tf32 (default: fp32 tensors, Linears on tcgen05) | fp32 (SIMT parity path)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="train: one all-reduce after the backward instead of two overlapped buckets")
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
