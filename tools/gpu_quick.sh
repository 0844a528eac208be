#!/bin/bash
# quick GPU check of the DoG path against the oracle (small image), under a timeout
timeout 300 python - <<'PY' 2>&1 | tail -30
import numpy as np, sys, time
sys.path.insert(0, '.')
import visualslam_b200 as vs
from oracle import c_oracle as O
from visualslam_b200 import synth
ctx = vs.Context(0)
for (h, w) in ((192, 256), (300, 217), (540, 960)):
    img = synth.frame(0, h, w)
    res = ctx.dog_keypoints(img)
    kc, cs, _ = O.dog_keypoints(img)
    k = res["keypoints"]
    got = np.stack([k[n] for n in ("row", "col", "value", "padding", "octave", "level")], 1)
    ok = np.array_equal(got, kc)
    print((h, w), "keypoints", len(got), "oracle", len(kc), "EQUAL" if ok else "DIFFER")
    if not ok:
        for o in range(4):
            a = got[got[:, 4] == o]; b = kc[kc[:, 4] == o]
            print("  octave", o, len(a), len(b), np.array_equal(a, b))
            if not np.array_equal(a, b):
                sa = set(map(tuple, a)); sb = set(map(tuple, b))
                print("   only gpu", sorted(sa - sb)[:6]); print("   only oracle", sorted(sb - sa)[:6])
ctx.close()
PY
