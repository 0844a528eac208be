#!/bin/bash
# A/B of the Harris kernel variants: per-kernel ms of a short device-resident bench
for v in "VSLAM_HARRIS_FAST=0" "VSLAM_HARRIS_CTAS=3" "VSLAM_HARRIS_CTAS=4" "VSLAM_HARRIS_CTAS=5"; do
  env $v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/ab.json 2> gpurun_out/ab.err
  python - "$v" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open("gpurun_out/ab.json") if l.startswith("{")][-1])
    k = d["roofline"]["kernel_ms_per_step"]
    print(sys.argv[1], "step", round(d["ms_per_step"], 2), "harris", [round(v, 3) for n, v in k.items() if "harris" in n], "corners", d["corners_per_step"])
except Exception as e:
    print(sys.argv[1], "failed", e); print(open("gpurun_out/ab.err").read()[-800:])
PY
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k harris 2>&1 | tail -3
