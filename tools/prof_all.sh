#!/bin/bash
# One profiling pass of the bench command (run under gpurun): plain run, launch list, then ncu --set full of the hot kernels
# of the timed step (upsample, 4 x (reflect_pad + octave), harris = 10 launches; the 10 of the parity step and the 10 of the
# warm-up step are skipped).
set -x
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-sweep"
$CMD > gpurun_out/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'octave_mma_k|octave_fused_k|harris_fast_k|harris_march_k|upsample2x_k|reflect_pad_k' -s 20 -c 10 -f -o gpurun_out/prof_all $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
