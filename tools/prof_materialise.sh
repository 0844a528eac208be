#!/bin/bash
# ncu --set full of the materialise-mode kernels (grad_all_k, octave_fused_k: first launches after warm-up)
bash tools/gpu_materialise.sh > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'grad_all_k|octave_fused_k' -s 10 -c 6 -f -o gpurun_out/prof_mat bash tools/gpu_materialise.sh > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
