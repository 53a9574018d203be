"""Per-source-line instruction / stall-sample table of one kernel from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`.  usage: tools/ncu_regions.py src.csv substr [file]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
only = sys.argv[3] if len(sys.argv) > 3 else None
func = None
agg, samp, src = collections.Counter(), collections.Counter(), {}
for r in rows:
    if not r:
        continue
    if r[0] == "Function Name":
        func = r[1]
        continue
    if r[0] == "File Path":
        fname = r[1].split('/')[-1]
        continue
    if r[0] == "Line No":
        H = r
        ie, sm = H.index("Instructions Executed"), H.index("# Samples")
        continue
    if want not in (func or ""):
        continue
    if r[0].isdigit():
        k = (fname, int(r[0]))
        agg[k] += int(r[ie]) if r[ie].isdigit() else 0
        samp[k] += int(r[sm]) if r[sm].isdigit() else 0
        src[k] = r[1]
tot, ts = sum(agg.values()) or 1, sum(samp.values()) or 1
print(f"total {tot / 1e6:.2f} M warp instructions, {ts} samples")
for (f, l), v in sorted(agg.items(), key=lambda kv: (kv[0][0] != (only or kv[0][0]), kv[0][0], kv[0][1])):
    if v > 0.004 * tot or samp[(f, l)] > 0.01 * ts:
        print(f"{f[:14]:14s}{l:5d} {v / 1e6:7.2f}M {100 * v / tot:5.1f}% s={100 * samp[(f, l)] / ts:5.1f}% | {src[(f, l)].strip()[:90]}")
