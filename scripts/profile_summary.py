#!/usr/bin/env python
"""Text summary of one ncu report for profiles/: headline metrics of the raw page + hottest source lines.
usage: profile_summary.py report.ncu-rep > profiles/<name>.txt"""
import csv, io, subprocess, sys, os
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2:]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__inst_executed.sum", "sm__inst_executed.avg.per_cycle_active", "sm__instruction_throughput.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__maximum_warps_per_active_cycle_pct", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed_pipe_fp64.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
print("# %s" % os.path.basename(rep))
for v in vals:
    d = dict(zip(hdr, v)); u = dict(zip(hdr, units))
    print("kernel: %s   grid %s block %s" % (d.get("Kernel Name"), d.get("Grid Size"), d.get("Block Size")))
    for k in want:
        if k in d: print("  %-70s %s %s" % (k, d[k], u.get(k, "")))
here = os.path.dirname(os.path.abspath(__file__))
print()
print(subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), rep, "30"], capture_output=True, text=True).stdout)
