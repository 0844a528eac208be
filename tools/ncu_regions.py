#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` (SASS view) dump of one kernel: share of executed warp instructions and of
stall samples between consecutive BAR.SYNC instructions, with the top opcodes and stall reasons of each region.
usage: ncu_regions.py dump.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
H = rows[1]; idx = {h: i for i, h in enumerate(H)}
ie, ss = idx["Instructions Executed"], idx["# Samples"]
stall_cols = [(h, i) for h, i in idx.items() if h.startswith("stall_") and "Not Issued" not in h]
seg = 0; acc = collections.defaultdict(lambda: [0, 0, collections.Counter(), collections.Counter()])
for r in rows[2:]:
    if len(r) < len(H): continue
    n, s, t = int(r[ie]), int(r[ss]), r[1].strip()
    op = t.split(); o = (op[1] if op[0].startswith("@") else op[0]).split(".")[0]
    a = acc[seg]; a[0] += n; a[1] += s; a[2][o] += n
    for h, i in stall_cols:
        try: a[3][h] += int(r[i])
        except ValueError: pass
    if "BAR.SYNC" in t: seg += 1
tot = sum(v[0] for v in acc.values()); st = sum(v[1] for v in acc.values())
print(f"total warp-instr {tot}  samples {st}")
for k, v in acc.items():
    print(f"region {k}: inst {v[0]/tot:.3f} samp {v[1]/st:.3f} |", " ".join(f"{o}:{n/tot:.3f}" for o, n in v[2].most_common(7)))
    print("          stalls:", " ".join(f"{h[6:]}:{n/st:.3f}" for h, n in v[3].most_common(5)))
