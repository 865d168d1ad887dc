#!/usr/bin/env python
"""Markdown summary of one kernel from an .ncu-rep (ncu --set full): headline counters, traffic, per-issue stalls.

    python tools/ncu_summary.py prof.ncu-rep "title" > profiles/rNN_ncu_full_summary.md
"""
import csv
import io
import subprocess
import sys

path, title = sys.argv[1], sys.argv[2]
raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = {h: (u, v) for h, u, v in zip(rows[0], rows[1], rows[2])}
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_warps',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__icc_request_hit_rate.pct',
        'gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'sass__inst_executed_register_spilling']
print('# %s\n' % title)
print('| metric | value | unit |\n|---|---|---|')
for k in keys:
    if k in d:
        print('| %s | %s | %s |' % (k, d[k][1], d[k][0]))
print('\nWarps stalled per issued instruction (smsp__average_warps_issue_stalled_*_per_issue_active):\n\n| reason | ratio |\n|---|---|')
st = [(h, float(v)) for h, (u, v) in d.items() if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for h, v in sorted(st, key=lambda x: -x[1])[:10]:
    print('| %s | %.2f |' % (h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v))
