#!/usr/bin/env python3
"""Per-source-line instruction / stall-sample shares of one kernel from an ncu report:
   ncu -i X.ncu-rep --page source --print-source cuda,sass --csv > src.csv ; python tools/ncu_lines.py src.csv [file-substring] [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cur, hdr, per = None, None, {}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr) - 2: continue
    if r[0] == "" : continue            # SASS rows under a line
    try:
        ln = int(r[0]); n = int(r[hdr.index("Instructions Executed")]); s = int(r[hdr.index("# Samples")])
    except ValueError:
        continue
    per[(cur, ln)] = (r[1], n, s)
tot = sum(v[1] for v in per.values()); ts = sum(v[2] for v in per.values())
print("total warp instructions %d, samples %d" % (tot, ts))
files = {}
for (f, ln), (src, n, s) in per.items():
    files.setdefault(f, [0, 0]); files[f][0] += n; files[f][1] += s
for f, v in files.items(): print("  %-70s instr %5.1f%% samples %5.1f%%" % (f[-70:], 100 * v[0] / tot, 100 * v[1] / max(ts, 1)))
sel = [(k, v) for k, v in per.items() if want in k[0]]
for (f, ln), (src, n, s) in sorted(sel, key=lambda t: -t[1][1])[:top]:
    print("%5d %6.2f%% %6.2f%%  %s" % (ln, 100 * n / tot, 100 * s / max(ts, 1), src.strip()[:120]))
