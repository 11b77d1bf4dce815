#!/usr/bin/env python
"""Print the key metrics of an `ncu --page raw --csv` dump: python scripts/ncu_keys.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, vals = rows[0], rows[1], rows[2]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit", "smsp__inst_executed.sum", "sm__pipe_fp64_cycles_active.avg.pct",
        "dmma_cycles_active.avg.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__issue_active.avg.pct", "lts__t_sector_hit_rate.pct", "launch__grid_size",
        "sm__inst_executed_pipe_lsu.avg.pct", "sm__inst_executed_pipe_alu.avg.pct",
        "smsp__average_warps_issue_stalled", "sm__pipe_shared_cycles_active.avg.pct", "sm__inst_executed_pipe_fma.avg.pct",
        "l1tex__throughput.avg.pct", "lts__throughput.avg.pct"]
for h, u, v in zip(hdr, units, vals):
    if any(k in h for k in keys) and "not_issued" not in h:
        print(h, u, v)
