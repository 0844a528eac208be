#!/bin/bash
# ncu --set full capture of octave_mma_k launches after warm-up (run under gpurun); the report lands in gpurun_out/
# $1 = launches to skip (default 4 = the warm-up step's four octaves), $2 = report name, $3 = launches to capture (default 1)
python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:octave_mma_k -s ${1:-4} -c ${3:-1} -f -o gpurun_out/${2:-prof_mma} \
    python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e --no-sweep > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
