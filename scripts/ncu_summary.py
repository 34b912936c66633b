#!/usr/bin/env python
"""Condense an .ncu-rep (ncu --set full) into the CSV kept under profiles/: one row per captured launch."""
import csv, io, subprocess, sys
METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
           "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed", "l1tex__m_xbar2l1tex_read_bytes.sum"]
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, body = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
keep = [m for m in METRICS if m in idx]
with open(out, "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["Kernel Name"] + keep)
    w.writerow([""] + [units[idx[m]] for m in keep])
    for r in body:
        w.writerow([r[idx["Kernel Name"]][:60]] + [r[idx[m]] for m in keep])
