#!/usr/bin/env python3
"""Per-stage (BAR.SYNC-delimited) stall samples and instruction counts from an ncu source-page CSV."""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r)
hdr = rows[hi]
iS, iE, iN, iA = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Address')
names = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
idx = {n: hdr.index(n) for n in names}
new = lambda: dict(n=0, e=0, s=collections.Counter(), ops=collections.Counter(), first=None)
seen, seg, cur, tot = set(), [], new(), 0
for r in rows[hi + 1:]:
    if len(r) <= iE or r[iA] in seen or not r[iE].isdigit():
        continue
    seen.add(r[iA])
    e, n, src = int(r[iE] or 0), int(r[iN] or 0), r[iS].strip()
    op = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_]+)', src).group(2)
    if cur['first'] is None:
        cur['first'] = r[iA][-5:]
    cur['n'] += n
    cur['e'] += e
    cur['ops'][op] += e
    tot += n
    for k in names:
        cur['s'][k[6:]] += int(r[idx[k]] or 0)
    if op == 'BAR':
        seg.append(cur)
        cur = new()
seg.append(cur)
print('total samples', tot)
for g in seg:
    print(g['first'], 'samples %6d (%4.1f%%)' % (g['n'], 100 * g['n'] / max(1, tot)), 'instr %9d' % g['e'],
          dict(g['s'].most_common(4)), dict(g['ops'].most_common(7)))
