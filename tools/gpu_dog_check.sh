#!/bin/bash
# DoG path check after an octave-kernel change: the DoG parity tests, then a short device-resident bench with per-kernel times
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "dog or keypoints or 1080p or tcgen05 or pyramid" 2>&1 | tail -4
python bench.py --steps ${1:-5} --warmup 3 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err
python - <<'PY'
import json
try:
    d = json.loads([l for l in open("gpurun_out/bench_short.json") if l.startswith("{")][-1])
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "kp", d["keypoints_per_step"], "corners", d["corners_per_step"])
    print(d["roofline"]["kernel_ms_per_step"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_short.err").read()[-1500:])
PY
