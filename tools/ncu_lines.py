#!/usr/bin/env python3
"""Top source lines of a kernel by warp-stall samples: tools/ncu_lines.py report.ncu-rep kernel_regex [n]"""
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kern, '--print-source', 'cuda,sass'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = [i for i, r in enumerate(rows) if '# Samples' in r][0]
h = rows[hi]
iS, iE = h.index('# Samples'), h.index('Instructions Executed')
stall_cols = [i for i, c in enumerate(h) if c.startswith('stall_') and 'Not Issued' not in c]
agg, tot = [], 0
for r in rows[hi + 1:]:
    if len(r) <= iS or r[0] == '':
        continue
    try:
        n = int(r[iS] or 0)
    except ValueError:
        continue
    st = {}
    for i in stall_cols:
        try:
            v = int(r[i] or 0)
        except ValueError:
            v = 0
        if v:
            st[h[i][6:]] = v
    agg.append((n, int(r[0]), r[1][:90], int(r[iE] or 0), st))
    tot += n
print('total samples', tot)
for n, ln, src, e, st in sorted(agg, reverse=True)[:top]:
    t4 = sorted(st.items(), key=lambda x: -x[1])[:4]
    print(f'{100 * n / tot:5.1f}% L{ln:4d} inst={e:10d} {src[:72]:72s} {t4}')
