#!/usr/bin/env python
"""ncu launch list (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum; --csv) of the bf16 forward -> the JSON
bench.py reads for roofline.traffic: DRAM bytes per launch of each kernel of ONE forward and their sum.

    python scripts/ncu_traffic.py profiles/r02_launches_bf16.csv 4096 > profiles/r02_ncu_traffic_bf16.json

Kernels are mapped to the stage names of xattn_stage_name() by their function names; the LAST complete forward in the list is used
(earlier ones include first-touch effects)."""
import collections
import csv
import json
import sys

src, batch = sys.argv[1], int(sys.argv[2])
rows = [r for r in csv.reader(open(src)) if len(r) > 12]
hdr = rows[0]
ci = {n: hdr.index(n) for n in ("ID", "Kernel Name", "Metric Name", "Metric Value")}
launch = collections.OrderedDict()
for r in rows[1:]:
    d = launch.setdefault(int(r[ci["ID"]]), {"name": r[ci["Kernel Name"]]})
    d[r[ci["Metric Name"]]] = float(r[ci["Metric Value"]].replace(",", ""))


def stage_of(name: str):
    if "gemm_tc_kernel<0" in name:
        return "gemm_v0"
    if "gemm_tc_kernel<3" in name:
        return "gemm_a0"
    if "gemm_tc_kernel<4" in name or "gemm_tc_kernel<5" in name:
        return "video_update"
    for key, st in (("video_pre", "video_pre"), ("prior_kernel", "prior"), ("v2a_tc", "v2a_attn"), ("a2v_tc", "a2v_attn"),
                    ("head_pooled", "head")):
        if key in name:
            return st
    return None


seq = [(i, stage_of(d["name"]), d) for i, d in launch.items() if stage_of(d["name"]) is not None]
heads = [k for k, (_, st, _) in enumerate(seq) if st == "head"]
assert heads, "no complete forward in the launch list"
end = heads[-1]
start = heads[-2] + 1 if len(heads) > 1 else 0
kernels, times = collections.OrderedDict(), collections.OrderedDict()
for _, st, d in seq[start:end + 1]:
    kernels[st] = kernels.get(st, 0.0) + d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    times[st] = times.get(st, 0.0) + d.get("gpu__time_duration.sum", 0.0) / 1e3
out = {"batch": batch, "source": src, "metric": "dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu, --clock-control none)",
       "kernels": kernels, "kernel_us_under_ncu": times, "step_total_bytes": sum(kernels.values()),
       "algorithmic_bytes": 245280 * batch, "ratio_to_algorithmic": sum(kernels.values()) / (245280 * batch)}
print(json.dumps(out, indent=1))
