#!/bin/bash
VSLAM_UMMA_DBG=1 timeout 300 python - <<'PY' 2>&1 | tail -12
import numpy as np, sys
sys.path.insert(0, '.')
import visualslam_b200 as vs
from visualslam_b200 import synth
ctx = vs.Context(0)
f = np.stack([synth.frame(i, 1080, 1920) for i in range(8)], 0)
f = np.concatenate([f] * 4, 0)
for _ in range(2):
    r = ctx.dog_keypoints(f)
print(sum(r["counts"]))
PY
