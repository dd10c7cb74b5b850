#!/usr/bin/env python
"""Group the SASS of one kernel (ncu --page source --csv) into regions of similar execution count."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r)
hdr = rows[hi]
ia, iex, ith = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Avg. Threads Executed')
d = []
for r in rows[hi + 1:]:
    try:
        d.append((r[ia], int(r[iex] or 0), float(r[ith] or 0)))
    except (ValueError, IndexError):
        pass
tot = sum(x[1] for x in d)
print('total warp instructions', tot, 'SASS lines', len(d))
i = 0
while i < len(d):
    j = i
    while j + 1 < len(d) and abs(d[j + 1][1] - d[i][1]) <= 0.15 * max(d[i][1], 1):
        j += 1
    ex = sum(x[1] for x in d[i:j + 1])
    th = sum(x[2] * x[1] for x in d[i:j + 1]) / max(ex, 1)
    if ex / max(tot, 1) > 0.012:
        print('lines %4d-%4d n=%4d exec~%8d thr=%4.1f share=%5.1f%%  %s' % (i, j, j - i + 1, d[i][1], th, 100 * ex / tot, d[i][0][:70]))
    i = j + 1
