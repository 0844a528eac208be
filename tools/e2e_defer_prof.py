#!/usr/bin/env python
"""deferred stream-of-batches loop (bench.py's e2e): wall time per step and the sum of kernel times inside it (vslam_profile),
with and without the look-ahead upload / the list copies, to see what the copies cost the kernels."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import visualslam_b200 as vs
from visualslam_b200 import synth
nf, h, w = 256, 1080, 1920
ctx = vs.Context(0)
uniq = [synth.frame(i, h, w) for i in range(8)]
host = torch.empty((nf, h, w), dtype=torch.uint8, pin_memory=True); hnp = host.numpy()
for i in range(nf): hnp[i] = uniq[i % 8]
dev = host.to("cuda:0")
cap = 40000 * nf * h * w // 1000000
kp = [torch.empty((cap, 6), dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]
co = [torch.empty((cap * 4, 2), dtype=torch.int32, pin_memory=True).numpy() for _ in range(2)]
kd = torch.zeros((cap, 6), dtype=torch.int32, device="cuda:0"); cd = torch.zeros((cap * 4, 2), dtype=torch.int32, device="cuda:0")
def run(label, upload, host_lists, n=6):
    it = [0]; pend = [ctx.upload(hnp) if upload else dev]
    def step():
        i = it[0]; it[0] += 1
        up = pend[0]
        if upload: pend[0] = ctx.upload(hnp)
        if host_lists:
            ctx.dog_harris(up, kp_out=kp[i & 1], co_out=co[i & 1], compact=True, defer=True)
            if i: ctx.finish()
        else:
            ctx.dog_harris(up, kp_out=kd, co_out=cd, compact=True)
    step(); step(); torch.cuda.synchronize()
    ctx.profile(True)
    t0 = time.perf_counter()
    for _ in range(n): step()
    if host_lists: ctx.finish()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / n * 1e3
    p = ctx.profile_read(); ctx.profile(False)
    ks = sum(v["ms"] for v in p.values()) / n
    print("%-34s step %.2f ms, kernels %.2f ms" % (label, ms, ks), {k[:12]: round(v["ms"] / n, 2) for k, v in p.items() if v["launches"]})
run("resident, device lists", False, False)
run("resident, host lists (deferred)", False, True)
run("look-ahead upload, device lists", True, False)
run("look-ahead upload, host (deferred)", True, True)
