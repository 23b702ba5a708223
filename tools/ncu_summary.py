#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` export: per launch duration, DRAM traffic, DRAM %, tensor-pipe %,
occupancy, registers.  usage: ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv"""
import csv
import sys

want = ['Kernel Name', 'launch__grid_size', 'launch__registers_per_thread', 'gpu__time_duration.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct']
with open(sys.argv[1]) as f:
    r = csv.reader(f)
    hdr = next(r)
    units = next(r)
    idx = {w: hdr.index(w) for w in want if w in hdr}
    print(' | '.join(f'{w.split(".")[0][-28:]}[{units[i]}]' for w, i in idx.items()))
    for row in r:
        print(' | '.join(row[i].split('(')[0][:30] for i in idx.values()))
