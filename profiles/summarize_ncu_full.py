#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one row per launch) into the handful of metrics DESIGN.md cites.
  ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv; python profiles/summarize_ncu_full.py /tmp/raw.csv"""
import csv
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__cluster_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active"]
rows = list(csv.reader(open(sys.argv[1])))
h, u = rows[0], rows[1]
for n, r in enumerate(rows[2:]):
    print(f"\n== launch {n + 1}: {r[h.index('Kernel Name')].split('(')[0]}")
    for w in WANT:
        if w in h:
            i = h.index(w)
            print(f"   {w:72s} {r[i]:>16s} {u[i]}")
