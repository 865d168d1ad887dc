#!/usr/bin/env python
"""Cold code per tick: SASS instructions that a warp executes at most a few times per physics tick (straight-line code that
misses the instruction caches) against the loop bodies it executes hundreds of times, per device function, from an ncu
--page source --csv dump joined with nvdisasm -g by instruction index (as tools/ncu_by_function.py).

    python tools/ncu_cold_code.py dis.txt src.csv '_Z10env_kernelILi3ELi32EEv7KParams' <ticks in the captured launch> [--launch i]
"""
import collections
import csv
import re
import sys

dis, src, kern, ticks = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4])
which = int(sys.argv[sys.argv.index('--launch') + 1]) if '--launch' in sys.argv else 0
ins, infn, cur, loc = [], False, 'kernel body', ('?', 0)
for ln in open(dis):
    if ln.startswith('.text.'):
        infn = ln.strip() == '.text.%s:' % kern; cur = 'kernel body'; continue
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        loc = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'^(\$[^:]+):', ln)
    if m:
        cur = m.group(1).split('$')[-1]; continue
    if re.match(r'^\s+/\*[0-9a-f]{4,5}\*/', ln):
        ins.append((cur, loc))
rows = list(csv.reader(open(src)))
secs = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
st = secs[which]; hdr = rows[st + 1]; body = rows[st + 2: secs[which + 1] if which + 1 < len(secs) else None]
ie, isamp = hdr.index('Instructions Executed'), hdr.index('# Samples')
data = [(int(r[ie]), int(r[isamp])) for r in body if len(r) > ie and r[ie].isdigit()]
assert len(ins) == len(data), (len(ins), len(data))
byf = collections.defaultdict(lambda: [0.0, 0, 0, 0.0])
for (fn, _), (c, s) in zip(ins, data):
    k = fn[:44]
    if c > 3 * ticks:
        byf[k][2] += s; byf[k][3] += c
    else:
        byf[k][0] += min(c, ticks) / ticks; byf[k][1] += s
tc, ts, th = sum(v[0] for v in byf.values()), sum(v[1] for v in byf.values()), sum(v[2] for v in byf.values())
texec = sum(c for c, _ in data)
print('%s, %.3g ticks in the launch: %.3g warp-instructions executed, %.0f per tick' % (kern, ticks, texec, texec / ticks))
print('cold code (executed <= 3 times per tick): %.0f distinct instructions per tick, %.1f %% of the stall samples; loops: %.1f %% of the samples'
      % (tc, 100.0 * ts / (ts + th), 100.0 * th / (ts + th)))
print('%-46s %10s %12s %12s' % ('function', 'cold/tick', 'cold samples', 'loop samples'))
for k, v in sorted(byf.items(), key=lambda x: -(x[1][1] + x[1][2])):
    if v[1] + v[2] > 0.001 * (ts + th):
        print('%-46s %10.0f %11.1f%% %11.1f%%' % (k, v[0], 100.0 * v[1] / (ts + th), 100.0 * v[2] / (ts + th)))
