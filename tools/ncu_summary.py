#!/usr/bin/env python
"""Selected columns of an ncu report's raw page, one row per launch:
    ncu -i X.ncu-rep --page raw --csv | python tools/ncu_summary.py > profiles/<name>.csv"""
import csv
import sys

WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "sm__cycles_elapsed.max.per_second", "sm__cycles_active.avg",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg", "sm__pipe_tensor_subpipe_imma_cycles_active_realtime.avg",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__inst_executed.sum"]
rows = list(csv.reader(sys.stdin))
h = rows[0]


def find(name):
    for i, c in enumerate(h):
        if c == name or c.endswith("." + name):
            return i
    return None


idx = [(w, find(w)) for w in WANT]
idx = [(w, i) for w, i in idx if i is not None]
out = csv.writer(sys.stdout)
out.writerow([w for w, _ in idx])
for r in rows[1:]:
    out.writerow([r[i] for _, i in idx])
