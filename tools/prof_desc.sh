#!/bin/bash
# Profiling pass of the descriptor stage (run under gpurun): plain run, launch list, ncu --set full of both kernels.
CMD="python tools/bench_descriptors.py --frames 8 --cpu-frames 0"
$CMD > gpurun_out/desc_plain.json 2> gpurun_out/desc_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'orient_hist_k|sift_k' --csv --log-file gpurun_out/launches_desc.csv $CMD > gpurun_out/ncu_desc_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'orient_hist_k|sift_k' -s 5 -c 5 -f -o gpurun_out/prof_desc $CMD > gpurun_out/ncu_desc_full.log 2>&1
tail -2 gpurun_out/ncu_desc_full.log
