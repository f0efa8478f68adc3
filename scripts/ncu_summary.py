#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU needed): scripts/ncu_summary.py <report> [kernel-substring]"""
import csv, subprocess, sys
KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'lts__t_sectors_srcunit_tex_op_write.sum', 'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_op_write.sum', 'lts__t_sectors_op_read.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_active.avg',
        'smsp__issue_active.avg.pct_of_peak_sustained_active']
def main():
    rep = sys.argv[1]
    sub = sys.argv[2] if len(sys.argv) > 2 else ''
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[h.index('Kernel Name')]
        if sub not in name:
            continue
        print('## ' + name[:100])
        for k in KEYS:
            if k in h:
                print(f"{k:80s} {r[h.index(k)]:>18s} {units[h.index(k)]}")
        for i, k in enumerate(h):
            if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio') and float(r[i] or 0) >= 0.2:
                print(f"{k:80s} {r[i]:>18s}")
        break
main()
