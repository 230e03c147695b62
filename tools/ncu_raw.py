#!/usr/bin/env python
"""Key metrics per launch from `ncu -i X.ncu-rep --page raw --csv`.  usage: ncu_raw.py file.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "l1tex__throughput.avg.pct_of_peak_sustained_active"]
idx = [(k, hdr.index(k)) for k in keys if k in hdr]
for r in rows[2:]:
    print(" | ".join("%s=%s%s" % (k.split(".")[0][-28:], r[i][:34], units[i] if k != "Kernel Name" else "") for k, i in idx))
