#!/usr/bin/env python
"""Key metrics of every kernel in an `ncu --set full` report as a markdown file.

  python tools/ncu_kernels_md.py "<title line>" <report.ncu-rep> <out.md>"""
import csv
import subprocess
import sys

title, rep, out = sys.argv[1:4]
WANT = [('gpu__time_duration.sum', 'duration'), ('dram__bytes_read.sum', 'dram read'), ('dram__bytes_write.sum', 'dram write'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram throughput % of peak'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'achieved occupancy %'),
        ('launch__registers_per_thread', 'registers/thread'), ('launch__grid_size', 'grid'), ('launch__block_size', 'block'),
        ('smsp__inst_executed.sum', 'warp instructions'), ('smsp__thread_inst_executed_per_inst_executed.ratio', 'active threads/instr'),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue slots busy %'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'stall long scoreboard /issue'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'stall short scoreboard /issue'),
        ('smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'stall math pipe /issue'),
        ('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'stall wait /issue'),
        ('l1tex__t_sector_hit_rate.pct', 'L1 hit %'), ('lts__t_sector_hit_rate.pct', 'L2 hit %')]
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, units = rr[0], rr[1]
with open(out, 'w') as f:
    f.write('# %s\n\nNumbers taken under the profiler are evidence for traffic and stall structure only.\n\n' % title)
    for r in rr[2:]:
        f.write('## `%s`\n\n| metric | value |\n|---|---:|\n' % r[h.index('Kernel Name')].split('(')[0])
        for m, label in WANT:
            if m in h:
                i = h.index(m)
                f.write('| %s | %s %s |\n' % (label, r[i], units[i]))
        f.write('\n')
print('written', out)
