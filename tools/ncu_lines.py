#!/usr/bin/env python
"""Hottest CUDA source lines of an `ncu --page source --print-source cuda,sass --csv` dump: executed warp instructions and
stall samples per (file, line).  usage: ncu_lines.py dump.csv [top]"""
import csv, sys, os
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; H = None; out = []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = os.path.basename(r[1]); continue
    if r and r[0] == "Line No": H = r; ie = H.index("Instructions Executed"); isamp = H.index("# Samples"); continue
    if H and len(r) > ie and r[2] == "-" and r[0].isdigit():
        try: out.append((int(r[ie]), int(r[isamp]), cur, int(r[0]), r[1].strip()))
        except ValueError: pass
ti = sum(o[0] for o in out); ts = sum(o[1] for o in out)
print("total inst", ti, "samples", ts)
for n, s, f, ln, src in sorted(out, key=lambda o: -o[1])[:top]:
    print(f"{s/ts:6.3f} samp {n/ti:6.3f} inst  {f}:{ln}  {src[:110]}")
