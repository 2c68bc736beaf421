#!/usr/bin/env python3
"""profiles/r2_traffic.json from `ncu --set full` reports: DRAM bytes (read + write) of ONE launch of each captured kernel.
   python tools/ncu_traffic.py <images_per_launch> report1.ncu-rep [report2.ncu-rep ...]
bench.py quotes the entry of the dominant kernel as roofline.traffic (and null when there is none)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_path = os.path.join(ROOT, "profiles", "r2_traffic.json")
res = json.load(open(out_path)) if os.path.exists(out_path) else {}
imgs = int(sys.argv[1])
for rep in sys.argv[2:]:
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[0]
    ik, ir, iw, it = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    units = rows[1]

    def to_bytes(v, u):
        v = float(v.replace(",", ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    for r in rows[2:]:
        name = r[ik].split("(")[0].split("::")[-1]
        res[name] = {"dram_bytes_per_launch": int(to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])), "images_per_launch": imgs,
                     "duration": r[it] + " " + units[it], "report": os.path.basename(rep)}
json.dump(res, open(out_path, "w"), indent=1)
print(json.dumps(res, indent=1))
