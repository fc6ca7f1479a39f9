"""Per-kernel time shares of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python scripts/launch_shares.py FILE"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 2:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    n = r[ki].split("(")[0][-60:]
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
print("| kernel | launches | us | share |\n|---|---:|---:|---:|")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"| `{k.strip()}` | {v[0]} | {v[1]:.1f} | {100 * v[1] / tot:.1f} % |")
print(f"| total | | {tot:.1f} | |")
