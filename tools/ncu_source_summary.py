#!/usr/bin/env python
"""Summarise `ncu --page source --csv` output: top stall sites and the instruction mix."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hi = next(i for i, r in enumerate(rows) if 'Source' in r and 'Address' in r)
hdr = rows[hi]
ia, ist, iex = hdr.index('Source'), hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Instructions Executed')
ith = hdr.index('Avg. Threads Executed')
data = []
for r in rows[hi + 1:]:
    if len(r) <= iex:
        continue
    try:
        data.append((int(r[ist] or 0), int(r[iex] or 0), r[ia], r[ith]))
    except ValueError:
        pass
tot = sum(d[0] for d in data) or 1
totex = sum(d[1] for d in data) or 1
print('stall samples', tot, 'warp instructions', totex, 'SASS lines', len(data))
for s, e, src, th in sorted(data, reverse=True)[:top]:
    print('%6d %5.1f%% ex=%9d thr=%5s  %s' % (s, 100 * s / tot, e, th, src[:100]))
mix = collections.Counter()
for s, e, src, th in data:
    parts = src.split()
    op = parts[1] if parts and parts[0].startswith('@') and len(parts) > 1 else (parts[0] if parts else '?')
    mix[op.split('.')[0]] += e
print(' '.join('%s:%.1f%%' % (k, 100 * v / totex) for k, v in mix.most_common(18)))
