#!/usr/bin/env python
"""Per-source-line warp-stall samples of one kernel:  ncu_lines.py REPORT.ncu-rep KERNEL_REGEX OBJECT.o [main_file_substr]
Joins `ncu --page source --csv` (SASS order) with `nvdisasm -g` line info of the same function (same instruction order)."""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, kre, obj = sys.argv[1:4]
main = sys.argv[4] if len(sys.argv) > 4 else os.path.basename(obj).replace(".o", ".cu")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
# several launches / template instances may match: keep the first block whose full name matches NCU_LINES_NAME (default: first)
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
want = os.environ.get("NCU_LINES_NAME", "")
pick = next((k for k in range(len(starts) - 1) if want in rows[starts[k]][1]), 0)
rows = rows[starts[pick]:starts[pick + 1]]
kname = rows[0][1]
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
si, ii, so = idx["# Samples"], idx["Instructions Executed"], idx["Source"]
ins = [(int(r[si] or 0), int(r[ii] or 0), r[so]) for r in rows[2:] if len(r) > si]
base = re.sub(r"[<(].*", "", kname).split("::")[-1]
mangled = os.environ.get("NCU_LINES_MANGLED", "")   # e.g. ILi4ELi6ELi1E to pick one template instance
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
    cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
    dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
lines, cur_file, cur_line, main_line, active = [], "?", 0, 0, False
for l in dis.splitlines():
    if l.startswith(".text."):
        active = base in l and mangled in l
        continue
    if not active:
        continue
    m = re.match(r'\s+//## File "([^"]+)", line (\d+)', l)
    if m:
        cur_file, cur_line = m.group(1), int(m.group(2))
        if main in cur_file:
            main_line = cur_line
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append((main_line, os.path.basename(cur_file), cur_line))
n = min(len(lines), len(ins))
print(f"kernel {kname[:80]}: {len(ins)} ncu instrs, {len(lines)} disasm instrs")
tot_s = sum(x[0] for x in ins) or 1; tot_i = sum(x[1] for x in ins) or 1
agg = collections.OrderedDict()
for k in range(n):
    a = agg.setdefault(lines[k][0], [0, 0]); a[0] += ins[k][0]; a[1] += ins[k][1]
srcl = open([p for p in (os.path.join(os.path.dirname(os.path.abspath(obj)), main), main) if os.path.exists(p)][0]).read().splitlines()
for ln, (s, i) in sorted(agg.items()):
    if s / tot_s >= 0.004 or i / tot_i >= 0.01:
        print(f"{ln:4d} smp {100*s/tot_s:5.1f}% ins {100*i/tot_i:5.1f}% | {srcl[ln-1].strip()[:100] if 0 < ln <= len(srcl) else ''}")
