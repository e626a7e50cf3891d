#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page raw --csv` output: one column per captured launch."""
import csv, sys
WANT = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'gpu__time_duration.sum',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum',
        'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_atom.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps', 'sm__maximum_warps_per_active_cycle_pct']
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
for w in WANT:
    if w in h:
        i = h.index(w)
        cell = (lambda v: v.replace("void ", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", "")[:64]) if w == 'Kernel Name' else (lambda v: v[:44])
        print("%-66s %-8s %s" % (w, rows[1][i], " | ".join(cell(r[i]) for r in rows[2:])))
