#!/usr/bin/env python
"""ncu launch list (gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum; --csv) of a few Stage-1 steps -> the JSON
bench.py reads for the train line's roofline.traffic: DRAM bytes of ONE step (every launch between two Adam kernels; the LAST
complete step of the list, which runs as the CUDA graph) and the same per kernel function.

    python scripts/ncu_traffic_train.py profiles/r02_traffic_train_tf32.csv 1024 tf32 > profiles/r02_ncu_traffic_train_tf32.json
"""
import collections
import csv
import json
import re
import sys

src, batch, precision = sys.argv[1], int(sys.argv[2]), sys.argv[3]
rows = [r for r in csv.reader(open(src)) if len(r) > 12]
hdr = rows[0]
ci = {n: hdr.index(n) for n in ("ID", "Kernel Name", "Metric Name", "Metric Value")}
launch = collections.OrderedDict()
for r in rows[1:]:
    d = launch.setdefault(int(r[ci["ID"]]), {"name": r[ci["Kernel Name"]]})
    d[r[ci["Metric Name"]]] = float(r[ci["Metric Value"]].replace(",", ""))
seq = list(launch.values())
adam = [i for i, d in enumerate(seq) if "adam_kernel" in d["name"]]
assert len(adam) >= 2, "need two complete steps in the launch list"
step = seq[adam[-2] + 1:adam[-1] + 1]


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name.replace("xf::<unnamed>::", "").replace("xf::", "")


kern = collections.OrderedDict()
for d in step:
    k = kern.setdefault(short(d["name"]), {"launches": 0, "dram_bytes": 0.0, "us_under_ncu": 0.0})
    k["launches"] += 1
    k["dram_bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
    k["us_under_ncu"] += d.get("gpu__time_duration.sum", 0.0) / 1e3
for k in kern.values():
    k["gbs_under_ncu"] = k["dram_bytes"] / (k["us_under_ncu"] * 1e-6) / 1e9 if k["us_under_ncu"] > 0 else None
total = sum(k["dram_bytes"] for k in kern.values())
alg = 490536 * batch
out = {"batch": batch, "precision": precision, "source": src,
       "metric": "dram__bytes_read.sum + dram__bytes_write.sum per launch (ncu, --clock-control none), summed over one step",
       "launches_per_step": len(step), "step_total_bytes": total, "step_us_under_ncu": sum(k["us_under_ncu"] for k in kern.values()),
       "algorithmic_bytes": alg, "ratio_to_algorithmic": total / alg,
       "kernels": collections.OrderedDict(sorted(kern.items(), key=lambda kv: -kv[1]["dram_bytes"]))}
print(json.dumps(out, indent=1))
