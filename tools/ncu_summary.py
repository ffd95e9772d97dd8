#!/usr/bin/env python
"""Print the metrics we track from an `ncu --page raw --csv` dump (one column per captured launch)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
keep = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg',
        'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg', 'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_shared.avg',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum', 'smsp__inst_executed_op_shared_atom.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum',
        'smsp__warps_eligible.avg.per_cycle_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'sm__cycles_active.avg']
keep += [h for h in hdr if 'issue_stalled' in h and 'per_issue_active' in h]
for k in keep:
    if k in hdr:
        i = hdr.index(k)
        print("%-82s %-10s %s" % (k[-82:], units[i][:10], " | ".join(r[i][:22] for r in rows[2:])))
