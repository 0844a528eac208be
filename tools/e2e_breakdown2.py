#!/usr/bin/env python
"""Where the end-to-end step goes, round 2 (run under gpurun): vslam_dog_harris with device / host lists, resident / uploaded
frames, look-ahead uploads.  CUDA-event + wall timing of 5 steps each."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import visualslam_b200 as vs
from visualslam_b200 import synth
nf, h, w = 256, 1080, 1920
stream = torch.cuda.Stream()
ctx = vs.Context(0, stream=stream.cuda_stream)
uniq = [synth.frame(i, h, w) for i in range(8)]
host = torch.empty((nf, h, w), dtype=torch.uint8, pin_memory=True); hnp = host.numpy()
for i in range(nf): hnp[i] = uniq[i % 8]
dev = host.to("cuda:0")
cap = 40000 * nf * h * w // 1000000
kp_host = torch.empty((cap, 6), dtype=torch.int32, pin_memory=True).numpy()
co_host = torch.empty((cap * 4, 2), dtype=torch.int32, pin_memory=True).numpy()
kp_dev = torch.zeros((cap, 6), dtype=torch.int32, device="cuda:0")
co_dev = torch.zeros((cap * 4, 2), dtype=torch.int32, device="cuda:0")
def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("separate, device lists      %.2f ms" % t(lambda: (ctx.dog_keypoints(dev, out=kp_dev), ctx.harris(dev, want_maps=False, device_out={"corners": co_dev}, compact=True))))
print("combined, device lists      %.2f ms" % t(lambda: ctx.dog_harris(dev, kp_out=kp_dev, co_out=co_dev, compact=True)))
print("combined, host lists        %.2f ms" % t(lambda: ctx.dog_harris(dev, kp_out=kp_host, co_out=co_host, compact=True)))
print("upload only                 %.2f ms" % t(lambda: (ctx.upload(hnp), ctx.sync())))
print("upload + combined, dev lists %.2f ms" % t(lambda: ctx.dog_harris(ctx.upload(hnp), kp_out=kp_dev, co_out=co_dev, compact=True)))
print("upload + combined, host     %.2f ms" % t(lambda: ctx.dog_harris(ctx.upload(hnp), kp_out=kp_host, co_out=co_host, compact=True)))
pend = [ctx.upload(hnp)]
def look():
    up = pend[0]; pend[0] = ctx.upload(hnp)
    return ctx.dog_harris(up, kp_out=kp_host, co_out=co_host, compact=True)
print("look-ahead upload, host     %.2f ms" % t(look))
def look_dev():
    up = pend[0]; pend[0] = ctx.upload(hnp)
    return ctx.dog_harris(up, kp_out=kp_dev, co_out=co_dev, compact=True)
print("look-ahead upload, dev      %.2f ms" % t(look_dev))
