#!/usr/bin/env python3
"""Summarise an .ncu-rep: tools/ncu_summary.py report.ncu-rep  (key metrics + top stalls per captured kernel)."""
import csv
import io
import subprocess
import sys

out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__inst_executed.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct',
        'smsp__thread_inst_executed_per_inst_executed.ratio']
for r in rows[2:]:
    print('=====', r[hdr.index('Kernel Name')])
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f'  {w:88s} {r[i]} {units[i]}')
    items = []
    for i, h in enumerate(hdr):
        if 'issue_stalled' in h and 'ratio' in h and 'not_issued' not in h:
            try:
                items.append((float(r[i]), h))
            except ValueError:
                pass
    print('  top stalls (warps per issue-active cycle):')
    for v, h in sorted(items, reverse=True)[:8]:
        print(f'    {v:6.3f} ' + h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''))
