#!/usr/bin/env python
"""Top stall locations of one kernel from an ncu report (source page, SASS rows).
usage: ncu_hot.py report.ncu-rep kernel_name [top]"""
import csv, subprocess, sys, io
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"]
si = hdr.index("# Samples") if "# Samples" in hdr else hdr.index("Warp Stall Sampling (All Samples)")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si] or 0) for r in body)
print("total samples", tot, "instructions", len(body))
agg = {}
for r in body:
    for i in stall_cols:
        agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
print("stall totals:", sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:10])
order = sorted(range(len(body)), key=lambda i: -int(body[i][si] or 0))[:top]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[j] or 0), hdr[j]) for j in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {int(r[si]):6d} {100*int(r[si])/max(tot,1):5.1f}%  {r[1].strip()[:90]:90s} {st}")
