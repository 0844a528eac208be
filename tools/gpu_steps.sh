#!/bin/bash
timeout 300 python - <<'PY' 2>&1 | tail -20
import numpy as np, sys
sys.path.insert(0, '.')
import visualslam_b200 as vs
from visualslam_b200 import synth
ctx = vs.Context(0)
import torch
for H in (64, 128, 192, 256, 512, 1080):
    f = np.stack([synth.frame(i, H, 1920) for i in range(8)], 0)
    f = np.concatenate([f] * 8, 0)
    d = torch.from_numpy(f).cuda()
    ctx.dog_keypoints(d)
    ctx.profile(True)
    for _ in range(3):
        ctx.dog_keypoints(d)
    p = ctx.profile_read()
    ctx.profile(False)
    k = p["octave_fused_k"]
    print(H, "octave0 rows", 2 * H, "steps/CTA", (2 * H + 127) // 128, "ms per call (4 octave launches)", k["ms"] / 3)
PY
