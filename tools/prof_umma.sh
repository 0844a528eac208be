#!/bin/bash
# ncu --set full capture of the octave_umma_k launches (octaves 0 and 1) of one warm-up step (run under gpurun)
python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:octave_umma_k -c 2 -f -o gpurun_out/${1:-prof_umma} \
    python bench.py --steps 1 --warmup 1 --frames 32 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
