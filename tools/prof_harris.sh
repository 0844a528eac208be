#!/bin/bash
# ncu --set full capture of one Harris launch (32 frames of 1080p) after warm-up; $1 = report name, env passes through
python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:harris_ -s 1 -c 1 -f -o gpurun_out/${1:-prof_harris} \
    python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
