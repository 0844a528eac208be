#!/usr/bin/env python
"""Summarise an `ncu --set full` report of the bench step (tools/prof_all.sh) into profiles/: per-launch headline metrics
as CSV and the measured DRAM traffic per launch as JSON (bench.py reads it for roofline.traffic).
usage: ncu_summary.py report.ncu-rep out_summary.csv traffic.json key"""
import csv, json, subprocess, sys
rep, out_csv, out_json, key = sys.argv[1:5]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
want = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__shared_mem_per_block_dynamic',
        'launch__shared_mem_per_block_static', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed'] + \
       [f'smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio' for s in
        ('barrier', 'short_scoreboard', 'long_scoreboard', 'wait', 'mio_throttle', 'math_pipe_throttle', 'not_selected', 'no_instruction')]
idx = {n: i for i, n in enumerate(h)}
with open(out_csv, 'w', newline='') as f:
    w = csv.writer(f)
    w.writerow(['metric', 'unit'] + [f'launch{i}' for i in range(len(rows) - 2)])
    for n in want:
        if n in idx: w.writerow([n, u[idx[n]]] + [r[idx[n]] for r in rows[2:]])
def val(r, n):
    return float(r[idx[n]]) * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}.get(u[idx[n]], 1)
tr = {}
for r in rows[2:]:
    k = r[idx['Kernel Name']]
    name = ('octave_mma_k' if 'octave_mma' in k else 'octave_umma_k' if 'octave_umma' in k else 'octave_fused_k' if 'octave_' in k else
            'harris_fast_k' if 'harris_fast' in k else 'harris_march_k' if 'harris' in k else 'upsample2x_k' if 'upsample' in k else 'reflect_pad_k')
    tr.setdefault(name, []).append(val(r, 'dram__bytes_read.sum') + val(r, 'dram__bytes_write.sum'))
t = {k: sum(v) / len(v) for k, v in tr.items()}
t['source'] = (f"{out_csv}: dram__bytes_read.sum + dram__bytes_write.sum per launch (octave and pad kernels: mean of the four launches of a "
               "step), ncu --set full --clock-control none, bench.py --steps 1 --warmup 1")
json.dump({key: t}, open(out_json, 'w'), indent=1)
print(t)
