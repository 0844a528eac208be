#!/usr/bin/env python
"""combined call, resident frames, host lists: step time for the VSLAM_HOST_CHUNKS in the environment; also per-kernel profile"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import visualslam_b200 as vs
from visualslam_b200 import synth
nf, h, w = 256, 1080, 1920
ctx = vs.Context(0)
uniq = [synth.frame(i, h, w) for i in range(8)]
dev = torch.from_numpy(np.stack([uniq[i % 8] for i in range(nf)], 0)).cuda()
cap = 40000 * nf * h * w // 1000000
kp_host = torch.empty((cap, 6), dtype=torch.int32, pin_memory=True).numpy()
co_host = torch.empty((cap * 4, 2), dtype=torch.int32, pin_memory=True).numpy()
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
f = lambda: ctx.dog_harris(dev, kp_out=kp_host, co_out=co_host, compact=True)
ms = t(f)
ctx.profile(True); f(); p = ctx.profile_read(); ctx.profile(False)
print("chunks", os.environ.get("VSLAM_HOST_CHUNKS", "6"), "step %.2f ms" % ms, {k[:14]: round(v["ms"], 2) for k, v in p.items() if v["launches"]})
