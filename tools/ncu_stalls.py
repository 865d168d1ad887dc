#!/usr/bin/env python
"""Print the headline counters and the per-issue stall breakdown of one kernel from an .ncu-rep (ncu --set full).

    python tools/ncu_stalls.py prof.ncu-rep [more.ncu-rep ...]
"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed.avg.per_cycle_elapsed', 'smsp__inst_executed.sum', 'sm__icc_request_hit_rate.pct',
        'gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed', 'gcc__cache_requests_type_instruction.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum']
cols = []
for path in sys.argv[1:]:
    raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    d = dict(zip(rows[0], rows[2]))
    cols.append((path, d))
names = KEYS + sorted(k for k in cols[0][1] if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio'))
print('%-70s' % 'metric' + ''.join('%16s' % p.split('/')[-1][:15] for p, _ in cols))
for k in names:
    short = k.replace('smsp__average_warps_issue_stalled_', 'stall/issue: ').replace('_per_issue_active.ratio', '')
    print('%-70s' % short[:70] + ''.join('%16s' % d.get(k, '-')[:14] for _, d in cols))
