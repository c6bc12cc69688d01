#!/usr/bin/env python
"""Print the key ncu metrics (one column per captured launch) of a report.  usage: ncu_keys.py report.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__inst_executed.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'launch__grid_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'lts__t_sectors.sum', 'lts__t_requests.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']
want += [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for k in want:
    if k in hdr:
        i = hdr.index(k)
        vals = [r[i] for r in rows[2:]]
        try:
            if all(float(v) < 0.05 for v in vals) and 'stalled' in k:
                continue
        except ValueError:
            pass
        print(f"{k.replace('smsp__average_warps_issue_stalled_', 'stall:').replace('_per_issue_active.ratio', '')} [{units[i]}] {vals}")
