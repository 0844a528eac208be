"""Small run of every kernel on the timed path (+ the descriptor stage) for compute-sanitizer (tools/sanitize.sh)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visualslam_b200 as vs
from visualslam_b200 import synth

ctx = vs.Context(0)
frames = np.stack([synth.frame(i, 270, 480) for i in range(2)], 0)      # all four octave_mma_k variants + upsample + pads + compaction
r = ctx.dog_keypoints(frames)
h = ctx.harris(frames)
pyr = ctx.build_pyramid(frames, 4, 1.6, materialise=0x78)                # octave_fused_k, grad_all_k
d = ctx.dog_detect(pyr)                                                  # dog_detect_k
desc = ctx.describe_keypoints(pyr, d["keypoints"], d["counts"])          # orient_hist_k, sift_k
print("keypoints", r["counts"], "corners", h["counts"], "descriptors", len(desc["descriptors"]), "launches", ctx.launches)
pyr.close()
ctx.close()
