#!/usr/bin/env python3
"""Dynamic SASS opcode histogram from `ncu -i X.ncu-rep --page source --csv --kernel-name regex:K` output."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[hi]
iS, iE = hdr.index('Source'), hdr.index('Instructions Executed')
iW, iX = hdr.index('L1 Wavefronts Shared'), hdr.index('L1 Wavefronts Shared Excessive')
cnt, wf, tot = collections.Counter(), collections.Counter(), 0
for r in rows[hi + 1:]:
    if len(r) <= iE:
        continue
    s = r[iS].strip()
    try:
        n = int(r[iE] or 0)
    except ValueError:
        continue
    m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)', s)
    if not m:
        continue
    op = m.group(2)
    if op in ('LDS', 'STS', 'LDG', 'STG'):
        op += m.group(3)
    cnt[op] += n
    tot += n
    try:
        if r[iW] and int(r[iW]):
            wf[op] += int(r[iW])
            wf[op + ' (excess)'] += int(r[iX] or 0)
    except ValueError:
        pass
print('total warp instructions', tot)
for k, v in cnt.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f'{k:16s} {v:12d} {100 * v / tot:5.1f}%')
print('shared wavefronts:', dict(wf))
