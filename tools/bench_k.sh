#!/bin/bash
# kernel times of one bench configuration (run under gpurun): prints ms per step of every kernel class
python bench.py --steps ${1:-5} --warmup 3 --no-cpu-baseline --no-e2e ${@:2} | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],2), {k: round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()})"
