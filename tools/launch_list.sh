#!/bin/bash
# ncu launch list (gpu__time_duration) of one small bench step (run under gpurun): per-launch kernel times into gpurun_out/$1.csv
CMD="python bench.py --steps 1 --warmup 1 --frames ${2:-64} --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/${1:-launches}.csv $CMD > gpurun_out/ncu_list.log 2>&1
python - <<PY
import csv
rows = [r for r in csv.reader(open("gpurun_out/${1:-launches}.csv")) if len(r) > 10 and r[0].isdigit()]
for r in rows[len(rows)//2:]:
    print(r[4][:60], r[-1])
PY
