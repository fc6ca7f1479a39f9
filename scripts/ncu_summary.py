#!/usr/bin/env python
"""Summarise a gpurun_out/<tag>/ directory into profiles/: launch list shares (launches.csv) and the
per-kernel metrics of the --set full capture (prof_step.ncu-rep).
usage: ncu_summary.py gpurun_out/<tag> profiles/<prefix> "<command that was profiled>" """
import csv, io, json, subprocess, sys, collections
src, dst, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
# ---- launch list -----------------------------------------------------------------------------------------
rows = []
with open(f"{src}/launches.csv") as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.reader(lines)
hdr = next(rd)
ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
ui = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rd:
    if r[mi] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    if r[ui] in ("ns", "nsecond"):
        v /= 1000.0
    elif r[ui] in ("ms", "msecond"):
        v *= 1000.0
    name = r[ki].split("(")[0][:90]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
with open(dst + "_launches.md", "w") as f:
    f.write(f"# ncu launch list\n\n`ncu --metrics gpu__time_duration.sum --clock-control none -c 400 {cmd}`\n"
            "(first 400 launches: setup + relabel + steps; cold-cache, serialised - compare shares, not absolutes)\n\n"
            "| kernel | launches | total us | share |\n|---|---:|---:|---:|\n")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{k}` | {n} | {us:.1f} | {100 * us / tot:.1f}% |\n")
# ---- full capture ------------------------------------------------------------------------------------------
out = subprocess.run(["ncu", "-i", f"{src}/prof_step.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
want = {"gpu__time_duration.sum": "duration", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct", "launch__registers_per_thread": "regs",
        "launch__grid_size": "grid", "launch__block_size": "block", "lts__t_sector_hit_rate.pct": "l2_hit_pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct"}
idx = {n: hdr.index(n) for n in want if n in hdr}
res = []
for r in rows[2:]:
    e = {"kernel": r[hdr.index("Kernel Name")].split("(")[0]}
    for n, i in idx.items():
        e[want[n]] = f"{r[i]} {units[i]}".strip()
    res.append(e)
json.dump({"command": cmd, "kernels": res}, open(dst + "_ncu_full.json", "w"), indent=1)
with open(dst + "_ncu_full.md", "w") as f:
    f.write(f"# ncu --set full (one step, cache flushed between replays)\n\n`{cmd}`\n\n")
    cols = ["kernel"] + list(want.values())
    f.write("| " + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
    for e in res:
        f.write("| " + " | ".join(str(e.get(c, "")) for c in cols) + " |\n")
print(open(dst + "_ncu_full.md").read())
