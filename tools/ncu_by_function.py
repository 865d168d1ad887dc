#!/usr/bin/env python
"""Attribute an ncu per-SASS-instruction dump (--page source --csv) to device functions and source lines.

    cuobjdump -xelf all urgym_b200/liburgym_b200.so && nvdisasm -g -c capi.sm_100a.cubin > dis.txt
    ncu -i prof.ncu-rep --page source --csv > src.csv
    python tools/ncu_by_function.py dis.txt src.csv '_Z10env_kernelILi0EEv7KParams' [--lines 25]

Joins by instruction index (same binary, same order) and prints executed warp-instructions, stall samples and the
top stall reasons per SASS sub-function, per source file, and for the hottest source lines.
"""
import collections
import csv
import re
import sys

dis, src, kern = sys.argv[1], sys.argv[2], sys.argv[3]
nlines = int(sys.argv[sys.argv.index('--lines') + 1]) if '--lines' in sys.argv else 25
which = int(sys.argv[sys.argv.index('--launch') + 1]) if '--launch' in sys.argv else 0

# ---- nvdisasm: per instruction (function label, file, line)
ins = []
infn, cur_fn, cur_loc = False, None, ('?', 0)
for ln in open(dis):
    if ln.startswith('.text.'):
        infn = ln.strip() == '.text.%s:' % kern
        cur_fn = 'kernel body'
        continue
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur_loc = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'^(\$[^:]+):', ln)
    if m:
        lab = m.group(1)
        cur_fn = lab.split('$')[-1]
        continue
    if re.match(r'^_Z\w+:$', ln.strip()) and kern not in ln:
        infn = False
        continue
    if re.match(r'\s*/\*[0-9a-f]{4,}\*/', ln):
        ins.append((cur_fn, cur_loc[0], cur_loc[1], ln.split('*/', 1)[1].strip()))

rows = list(csv.reader(open(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
seg = rows[starts[which] + 1:(starts[which + 1] if which + 1 < len(starts) else len(rows))]
hdr, body = seg[0], seg[1:]
ci = hdr.index('Instructions Executed'); cs = hdr.index('# Samples')
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
assert len(body) == len(ins), (len(body), len(ins))
tot_i = sum(int(r[ci]) for r in body); tot_s = sum(int(r[cs]) for r in body)
print('kernel %s launch %d: %d SASS instructions, %.3g warp-instructions executed, %d samples' % (kern, which, len(ins), tot_i, tot_s))


def agg(keyf):
    d = collections.defaultdict(lambda: [0, 0, collections.Counter(), 0])
    for (fn, f, l, txt), r in zip(ins, body):
        k = keyf(fn, f, l)
        d[k][0] += int(r[ci]); d[k][1] += int(r[cs]); d[k][3] += 1
        for i, h in stall_cols:
            v = int(r[i] or 0)
            if v:
                d[k][2][h] += v
    return d


def show(title, d, n=100):
    print('\n== ' + title)
    for k, v in sorted(d.items(), key=lambda x: -x[1][1])[:n]:
        top = ', '.join('%s %.0f%%' % (h[6:], 100.0 * c / max(v[1], 1)) for h, c in v[2].most_common(3))
        print('%-46s sass %5d  inst %5.1f%%  samples %5.1f%%  [%s]' % (str(k)[:46], v[3], 100.0 * v[0] / tot_i, 100.0 * v[1] / tot_s, top))


show('by SASS function', agg(lambda fn, f, l: fn))
show('by source file (innermost inlined frame)', agg(lambda fn, f, l: f))
show('hottest source lines', agg(lambda fn, f, l: '%s:%d' % (f, l)), nlines)
