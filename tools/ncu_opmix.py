#!/usr/bin/env python
"""Dynamic opcode mix from an `ncu --page source --print-source cuda,sass --csv` dump: executed warp instructions per SASS
opcode (and per hot SASS range).  usage: ncu_opmix.py dump.csv [rows_per_unit]  (rows_per_unit: divide counts to get per-warp-row)"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
unit = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
H = rows[2]; ie = H.index("Instructions Executed")
ops = collections.Counter(); seen = set(); tot = 0
for r in rows:
    if len(r) > ie and r[2].startswith("0x") and r[2] not in seen:
        seen.add(r[2])
        t = r[3].split()
        if not t: continue
        op = t[1] if t[0].startswith("@") else t[0]
        op = op.split(".")[0]
        n = int(r[ie]); ops[op] += n; tot += n
print("total", tot, "per unit", tot / unit)
for op, n in ops.most_common(45):
    print(f"{op:12s} {n/tot:6.3f} {n/unit:8.1f}")
