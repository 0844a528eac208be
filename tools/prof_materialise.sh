#!/bin/bash
# ncu --set full of the materialise-mode kernels (octave_mma_k MAT instantiations + grad_all_k: launches of one step after warm-up)
bash tools/gpu_materialise.sh > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'grad_all_k|octave_mma_k|octave_fused_k' -s 8 -c 8 -f -o gpurun_out/prof_mat bash tools/gpu_materialise.sh > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
