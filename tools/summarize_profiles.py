#!/usr/bin/env python
"""Turn the ncu artefacts of a round into the tracked summaries under profiles/.

  python tools/summarize_profiles.py <round tag> <launches.csv> <full.ncu-rep>

Writes profiles/<tag>_launches.csv (copy), profiles/<tag>_launches.md (per-kernel share of the
bench command's device time), profiles/<tag>_kernels.md (key metrics of the full capture) and
profiles/accum_traffic.json (dram bytes of k_accum per launch, read by bench.py).
"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

tag, launches_csv, rep = sys.argv[1:4]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(root, 'profiles')
os.makedirs(out, exist_ok=True)
shutil.copy(launches_csv, os.path.join(out, tag + '_launches.csv'))

rows = [r for r in csv.reader(open(launches_csv)) if len(r) > 5]
hdr = next(r for r in rows if 'Kernel Name' in r)
ik, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.OrderedDict()
for r in rows:
    if r is hdr or r[0] == 'ID':
        continue
    try:
        v = float(r[iv].replace(',', ''))
    except ValueError:
        continue
    scale = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(r[iu].strip(), 1e-3)
    agg.setdefault(r[ik], []).append(v * scale)
total = sum(sum(v) for v in agg.values())
with open(os.path.join(out, tag + '_launches.md'), 'w') as f:
    f.write('# %s - kernel launch list of `python bench.py --steps 2 --warmup 1 --no-cpu-baseline`\n\n' % tag)
    f.write('`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised launches: '
            'compare SHARES, not absolutes). Raw list: `%s_launches.csv`.\n\n' % tag)
    f.write('| kernel | launches | avg us | total us | share |\n|---|---:|---:|---:|---:|\n')
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        f.write('| `%s` | %d | %.1f | %.1f | %.1f %% |\n' % (k.split('(')[0], len(v), sum(v) / len(v), sum(v), 100 * sum(v) / total))
    f.write('\nThe bench command launches k_accum once per step on the resident batch and eight times per e2e '
            'step (chunked ingest), hence the mixed launch sizes in the average.\n')

raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, units = rr[0], rr[1]
WANT = [('gpu__time_duration.sum', 'duration'), ('dram__bytes_read.sum', 'dram read'), ('dram__bytes_write.sum', 'dram write'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram throughput % of peak'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'achieved occupancy %'),
        ('launch__registers_per_thread', 'registers/thread'), ('launch__grid_size', 'grid'), ('launch__block_size', 'block'),
        ('smsp__inst_executed.sum', 'warp instructions'), ('smsp__thread_inst_executed_per_inst_executed.ratio', 'active threads/instr'),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots busy %'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'stall long scoreboard /issue'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'stall short scoreboard /issue'),
        ('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'stall wait /issue'),
        ('lts__t_sector_hit_rate.pct', 'L2 hit %')]
traffic = None
with open(os.path.join(out, tag + '_kernels.md'), 'w') as f:
    f.write('# %s - `ncu --set full --clock-control none` of the two dominant kernels\n\n' % tag)
    f.write('Command: `python tools/profile_step.py --steps 3` (configs[1] resident in HBM, 28.8 M samples, three key kinds, p95). '
            'Numbers taken under the profiler are evidence for traffic and stall structure only; times come from bench.py.\n\n')
    for r in rr[2:]:
        name = r[h.index('Kernel Name')]
        f.write('## `%s`\n\n| metric | value |\n|---|---:|\n' % name)
        vals = {}
        for m, label in WANT:
            if m in h:
                i = h.index(m)
                vals[m] = (r[i], units[i])
                f.write('| %s | %s %s |\n' % (label, r[i], units[i]))
        f.write('\n')
        if 'k_accum' in name and 'dram__bytes_read.sum' in vals:
            def to_bytes(x):
                v, u = x
                return float(v.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
            traffic = to_bytes(vals['dram__bytes_read.sum']) + to_bytes(vals['dram__bytes_write.sum'])
if traffic:
    json.dump({'dram_bytes_per_launch': traffic, 'source': tag + '_kernels.md (ncu --set full, k_accum<pct>, 28.8 M samples)',
               'algorithmic_bytes_per_launch': 16 * 28800000}, open(os.path.join(out, 'accum_traffic.json'), 'w'), indent=1)
print('profiles written to', out)
