#!/bin/bash
# ncu --set full capture of the first octave_mma_k launch after warm-up (run under gpurun); the report lands in gpurun_out/
python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:octave_mma_k -s ${1:-4} -c 1 -f -o gpurun_out/${2:-prof_mma} \
    python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
