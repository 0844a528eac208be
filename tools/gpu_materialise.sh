#!/bin/bash
# materialise-mode line of the sweep only (vslam_pyramid_build with every gradient map resident + vslam_dog_detect); $1 = frames (8)
python - "${1:-8}" <<'PY'
import sys, json
sys.path.insert(0, ".")
import bench, argparse
a = argparse.Namespace(e2e_mode="auto")
R = bench.Runner(a, 0, 1, 0)
r = R.materialise(1080, 1920, int(sys.argv[1]), 3, 1)
print(json.dumps({k: r[k] for k in r if k in ("value", "ms_per_step", "build_hbm_frac", "kernel_ms_per_step", "build_achieved_GBps")}))
PY
