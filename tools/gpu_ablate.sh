#!/bin/bash
for v in "" "VSLAM_UMMA_NODETECT=1"; do
echo "== $v"
env $v timeout 300 python - <<'PY' 2>&1 | tail -3
import numpy as np, sys
sys.path.insert(0, '.')
import visualslam_b200 as vs
from visualslam_b200 import synth
import torch
ctx = vs.Context(0)
f = np.stack([synth.frame(i, 1080, 1920) for i in range(8)], 0)
d = torch.from_numpy(np.concatenate([f] * 8, 0)).cuda()
ctx.dog_keypoints(d)
ctx.profile(True)
for _ in range(3):
    r = ctx.dog_keypoints(d)
p = ctx.profile_read()
print("octave kernels ms per 64 frames:", sum(p[k]["ms"] for k in ("octave_fused_k","octave_mma_k","octave_umma_k")) / 3, "keypoints", sum(r["counts"]))
PY
done
