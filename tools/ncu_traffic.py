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
    # utilisation figures bench.py quotes as roofline.issue (first header that matches each substring)
    want = {"issue_slots_busy_pct": ["sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct"],
            "sm_throughput_pct": ["sm__throughput.avg.pct_of_peak_sustained_elapsed"],
            "alu_pipe_pct": ["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct"],
            "l1tex_throughput_pct": ["l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct"],
            "lsu_shared_wavefronts_pct": ["l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"],
            "dram_throughput_pct": ["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct"],
            "warp_instructions": ["smsp__inst_executed.sum", "sm__inst_executed.sum"],
            "ipc_active": ["sm__inst_executed.avg.per_cycle_active"],
            "registers_per_thread": ["launch__registers_per_thread"],
            "achieved_occupancy_pct": ["sm__warps_active.avg.pct_of_peak_sustained_active"]}
    cols = {}
    for key, subs in want.items():
        for sub in subs:
            hit = [i for i, h in enumerate(hdr) if h.startswith(sub)]
            if hit:
                cols[key] = hit[0]
                break
    for r in rows[2:]:
        name = r[ik].split("(")[0].split("::")[-1].replace("void ", "").split("<")[0].strip()
        ent = {"dram_bytes_per_launch": int(to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])), "images_per_launch": imgs,
               "duration": r[it] + " " + units[it], "report": os.path.basename(rep)}
        for key, i in cols.items():
            try:
                ent[key] = float(r[i].replace(",", ""))
            except ValueError:
                pass
        res[name] = ent
json.dump(res, open(out_path, "w"), indent=1)
print(json.dumps(res, indent=1))
