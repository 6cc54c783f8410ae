#!/usr/bin/env python3
"""Summarise an .ncu-rep (ncu --set full) into a markdown table of the metrics this project tracks.
usage: ncu_summary.py <report.ncu-rep> [title] > profiles/xxx.md"""
import csv, subprocess, sys, io
rep = sys.argv[1]
title = sys.argv[2] if len(sys.argv) > 2 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'launch__waves_per_multiprocessor',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores', 'l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct',
        'l1tex__t_sector_pipe_lsu_mem_local_op_st_hit_rate.pct', 'l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum']
want += [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
ki = hdr.index('Kernel Name')
print("# %s\n" % title)
print("| metric | unit | " + " | ".join(d[ki].split('(')[0] for d in data) + " |")
print("|---|---|" + "---|" * len(data))
for w in want:
    if w in hdr:
        i = hdr.index(w)
        vals = [d[i] for d in data]
        if all(v in ('0', '0.000000') for v in vals) and 'stalled' in w:
            continue
        print("| %s | %s | %s |" % (w.replace('smsp__average_warps_issue_stalled_', 'stall_').replace('_per_issue_active.ratio', ''), units[i], " | ".join(v[:14] for v in vals)))
