#!/usr/bin/env python
"""Summarise an `ncu --page source --print-source cuda,sass --csv` dump: executed warp instructions and stall samples
per CUDA source line (top N).  usage: ncu_lines.py dump.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = []; cur_file = ""
H = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": H = {h: i for i, h in enumerate(r)}; ie = r.index("Instructions Executed"); ss = r.index("# Samples"); continue
    if H is None or len(r) < 10 or not r[0].isdigit(): continue
    try: out.append((cur_file, int(r[0]), r[1].strip()[:90], int(r[ie]), int(r[ss])))
    except ValueError: pass
ti = sum(o[3] for o in out); ts = sum(o[4] for o in out)
print(f"total warp-instr {ti}  samples {ts}")
for f, ln, src, n, s in sorted(out, key=lambda o: -o[3])[:N]:
    print(f"{f}:{ln:4d} inst {n/ti:6.3f} samp {s/ts:6.3f} | {src}")
