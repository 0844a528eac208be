#!/bin/bash
# quick parity check + a short device-resident bench; prints step time and per-kernel times
bash tools/gpu_quick.sh
python bench.py --steps ${1:-5} --warmup 3 --no-cpu-baseline --no-e2e ${2:-} > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err
python - <<'PY'
import json
try:
    d = json.load(open("gpurun_out/bench_short.json"))
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "kp", d["keypoints_per_step"], "corners", d["corners_per_step"])
    print(d["roofline"]["kernel_ms_per_step"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_short.err").read()[-1500:])
PY
