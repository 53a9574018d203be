"""Summarise an ncu --csv launch list (gpu__time_duration.sum) per kernel: tools/launch_summary.py file.csv nsteps"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
nsteps = float(sys.argv[2]) if len(sys.argv) > 2 else 1
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1, "ms": 1e3}.get(r[ui], 1)
    a = agg.setdefault(r[ki][:70], [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
for k, v in agg.items():
    print(f"{k:72s} n/step={v[0] / nsteps:5.1f} {v[1] / nsteps:10.1f} us/step {100 * v[1] / tot:5.1f}%")
print(f"total {tot / nsteps:.1f} us/step")
