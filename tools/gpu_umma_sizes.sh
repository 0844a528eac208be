#!/bin/bash
VSLAM_UMMA=1 timeout 600 python - <<'PY' 2>&1 | tail -30
import numpy as np, sys
sys.path.insert(0, '.')
import visualslam_b200 as vs
from visualslam_b200 import synth
from oracle import c_oracle as O
c = vs.Context(0)
for (h, w) in ((192, 256), (300, 217), (70, 9), (9, 70), (64, 49), (129, 97), (5, 7), (257, 96)):
    img = synth.frame(h + w, h, w)
    try:
        r = c.dog_keypoints(img); k = r['keypoints']
    except Exception as e:
        print((h, w), "EXC", e); continue
    got = np.stack([k[n] for n in ('row', 'col', 'value', 'padding', 'octave', 'level')], 1)
    want = O.dog_keypoints(img)[0]
    ok = np.array_equal(got, want)
    print((h, w), len(got), len(want), "EQUAL" if ok else "DIFFER")
    if not ok:
        for o in range(4):
            a = got[got[:, 4] == o]; b = want[want[:, 4] == o]
            if not np.array_equal(a, b):
                sa = set(map(tuple, a)); sb = set(map(tuple, b))
                print("  octave", o, len(a), len(b), "only gpu", sorted(sa - sb)[:5], "only oracle", sorted(sb - sa)[:5])
PY
