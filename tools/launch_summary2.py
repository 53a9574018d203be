"""Per-kernel time and warp instructions from an ncu --csv launch list taken with
--metrics gpu__time_duration.sum,smsp__inst_executed.sum.   usage: tools/launch_summary2.py file.csv nsteps"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
nsteps = float(sys.argv[2]) if len(sys.argv) > 2 else 1
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, mi, vi, ui = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    a = agg.setdefault(r[ki][:62], [0.0, 0.0, 0])
    v = float(r[vi].replace(",", ""))
    if r[mi].startswith("gpu__time"):
        a[0] += v * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(r[ui], 1)
        a[2] += 1
    else:
        a[1] += v
tot = sum(v[0] for v in agg.values())
for k, v in agg.items():
    print(f"{k:64s} n/step={v[2] / nsteps:4.1f} {v[0] / nsteps:8.1f} us/step {100 * v[0] / tot:5.1f}%  {v[1] / nsteps / 1e6:8.2f} M warp-inst")
print(f"total {tot / nsteps:.1f} us/step")
