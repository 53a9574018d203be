"""DRAM bytes per launch of every kernel in an `ncu --set full` report -> profiles/rN_traffic.json.
usage: ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_traffic.py raw.csv "source note" > profiles/r2_traffic.json"""
import csv
import json
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
H, U = rows[0], rows[1]
ki, ri, wi, ti = H.index("Kernel Name"), H.index("dram__bytes_read.sum"), H.index("dram__bytes_write.sum"), H.index("gpu__time_duration.sum")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out, dur = {}, {}
for r in rows[2:]:
    name = re.sub(r"^void ", "", r[ki]).split("(")[0].split("<")[0]
    b = float(r[ri].replace(",", "")) * scale[U[ri]] + float(r[wi].replace(",", "")) * scale[U[wi]]
    out.setdefault(name, []).append(b)
    dur.setdefault(name, []).append(float(r[ti].replace(",", "")))
print(json.dumps({"source": sys.argv[2] if len(sys.argv) > 2 else "", "unit": "bytes (dram__bytes_read.sum + dram__bytes_write.sum), mean over the captured launches",
                  "dram_bytes_per_launch": {k: int(sum(v) / len(v)) for k, v in out.items()},
                  "launches_captured": {k: len(v) for k, v in out.items()},
                  "gpu_time_us_under_ncu": {k: sum(v) / len(v) for k, v in dur.items()}}, indent=1))
