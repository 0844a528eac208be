#!/usr/bin/env python
"""Where the end-to-end step goes (run under gpurun): each detector alone, device-resident vs host frames / host lists."""
import sys, time, os
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
co_host = torch.empty((cap * 4, 3), dtype=torch.int32, pin_memory=True).numpy()
kp_dev = torch.zeros((cap, 6), dtype=torch.int32, device="cuda:0")
co_dev = torch.zeros((cap * 4, 3), dtype=torch.int32, device="cuda:0")
def t(fn, n=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
print("dog device           %.2f ms" % t(lambda: ctx.dog_keypoints(dev, out=kp_dev)))
print("harris device corner %.2f ms" % t(lambda: ctx.harris(dev, want_maps=False, device_out={"corners": co_dev})))
print("upload only          %.2f ms" % t(lambda: (ctx.upload(hnp), ctx.sync() if hasattr(ctx, "sync") else torch.cuda.synchronize())))
print("upload + dog host    %.2f ms" % t(lambda: ctx.dog_keypoints(ctx.upload(hnp), out=kp_host)))
print("upload + harris host %.2f ms" % t(lambda: ctx.harris(ctx.upload(hnp), want_maps=False, corners_out=co_host)))
print("dog device -> host   %.2f ms" % t(lambda: ctx.dog_keypoints(dev, out=kp_host)))
print("harris dev -> host   %.2f ms" % t(lambda: ctx.harris(dev, want_maps=False, corners_out=co_host)))
def both():
    up = ctx.upload(hnp); ctx.dog_keypoints(up, out=kp_host); ctx.harris(up, want_maps=False, corners_out=co_host)
print("e2e both             %.2f ms" % t(both))
