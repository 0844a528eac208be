#!/bin/bash
# ncu --set full of dog_detect_k in materialise mode (run under gpurun after the plain run exited 0)
bash tools/gpu_materialise.sh 8 > gpurun_out/mat_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:dog_detect_k -c 2 -f -o gpurun_out/prof_detect bash tools/gpu_materialise.sh 8 > gpurun_out/ncu_detect.log 2>&1
tail -2 gpurun_out/ncu_detect.log
