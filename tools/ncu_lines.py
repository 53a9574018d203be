"""Per-source-line instruction / stall-sample totals from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`.   usage: tools/ncu_lines.py src.csv [substr-of-function]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
agg = {}
H = None
fname = func = ""
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        func = r[1]
        continue
    if r[0] == "Line No":
        H = r
        ie, sm = H.index("Instructions Executed"), H.index("# Samples")
        continue
    if H is None or want not in func or not r[0].isdigit():
        continue
    a = agg.setdefault((fname, int(r[0])), [0, 0, r[1]])
    a[0] += int(r[ie]) if r[ie].isdigit() else 0
    a[1] += int(r[sm]) if r[sm].isdigit() else 0
tot = sum(v[0] for v in agg.values()) or 1
tots = sum(v[1] for v in agg.values()) or 1
print("total warp-instr", tot, "samples", tots)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    print(f"{100 * v[0] / tot:5.1f}% inst {100 * v[1] / tots:5.1f}% samp {k[0]}:{k[1]:4d} | {v[2].strip()[:100]}")
