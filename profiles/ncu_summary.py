#!/usr/bin/env python3
"""Print the metrics we track from an .ncu-rep (run where ncu is installed):  ncu_summary.py rep [rep...]"""
import csv, subprocess, sys, io
KEEP = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
        'smsp__inst_executed.sum', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__warps_eligible.avg.per_cycle_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_lsu.sum', 'smsp__inst_executed_op_ldgsts.sum', 'local_load', 'smsp__inst_executed_op_local']
for rep in sys.argv[1:]:
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print('==', rep)
    for r in rows[2:]:
        print('kernel:', r[hdr.index('Kernel Name')][:100])
        for i, h in enumerate(hdr):
            if h in KEEP or ('issue_stalled' in h and h.endswith('per_issue_active.ratio')):
                try:
                    v = float(r[i].replace(',', ''))
                except ValueError:
                    continue
                if 'issue_stalled' in h and v < 0.15:
                    continue
                print('  %-95s %-12s %s' % (h, units[i], r[i]))
