#!/usr/bin/env python
"""Static SASS instruction count per source line for one kernel (from `nvdisasm -g -c` output)."""
import collections
import re
import sys

path, kernel = sys.argv[1], sys.argv[2]
src_file = sys.argv[3] if len(sys.argv) > 3 else None
counts = collections.Counter()
cur = None
active = False
for line in open(path):
    if line.startswith('.text.'):
        active = kernel in line
        continue
    if not active:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).rsplit('/', 1)[-1], int(m.group(2)))
        continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', line) and cur:
        counts[cur] += 1
tot = sum(counts.values())
print('static SASS instructions:', tot)
lines = {}
if src_file:
    for i, l in enumerate(open(src_file), 1):
        lines[i] = l.rstrip()
for (f, ln), c in sorted(counts.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    if c >= 4:
        txt = lines.get(ln, '')[:90] if src_file and f == src_file.rsplit('/', 1)[-1] else ''
        print('%-14s %4d  %4d  %s' % (f, ln, c, txt))
