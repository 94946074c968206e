#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) into the handful of metrics the roofline needs."""
import csv, subprocess, sys, io
rep = sys.argv[1]
pat = sys.argv[2:] or []
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct", "sm__inst_executed_pipe_fma", "lts__t_bytes.sum",
        "l1tex__t_bytes.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_srcunit_tex.sum", "sm__cycles_elapsed.avg ",
        "smsp__average_warp", "smsp__warp_issue_stalled", "l1tex__data_pipe_lsu_wavefronts.sum", "sm__pipe_tensor",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
        "smsp__cycles_active.avg", "sm__inst_executed.sum", "lts__t_sectors.sum", "dram__sectors"] + pat
for v in rows[2:]:
    print("==", v[4][:70], "grid", v[8], "block", v[7])
    for i, h in enumerate(hdr):
        name = h.split(".", 2)[-1] if h.count(".") >= 2 and h.split(".")[1].startswith("Triage") else h
        if any(k in h for k in KEYS):
            print(f"  {h} [{units[i]}] = {v[i]}")
