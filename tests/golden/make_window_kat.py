#!/usr/bin/env python
"""Writes tests/golden/window_kat.json: a hand-worked case for rules R1-R6 of oracle/window_stats.py.

Nothing here calls the oracle: the per-cell value lists below were worked out by hand from the
eight samples, and the percentile column is the literal rule R5 applied to those lists.

    gpus_per_node = 2, t0 = 100, step = 10, width = 20 (W = 2), n_windows = 3
    panes 0..3 = [100,110) [110,120) [120,130) [130,140); window j = panes j, j+1
    join pairs, in order: (node0,pid7)->pod5, (node0,pid8)->pod1, (node1,pid7)->pod1, (node0,pid7)->pod0
    (the last pair of a repeated key wins, so (node0,pid7) -> pod0)

    #  ts   gid pid sm   what happens
    1   99   0   7  50   before t0: left out (R3)
    2  100   0   7  10   node0 pod0 pane0
    3  105   1   7  20   node0 (gid 1 // 2) pod0 pane0
    4  110   0   8  30   pod1 pane1
    5  119   2   7  40   node1 pod1 pane1
    6  125   3   9  60   node1, pid 9 has no pair: no pod (R2), still gpu3 / node1, pane2
    7  139   0   7  70   pod0 pane3 (only window 2 sees it)
    8  140   0   7  80   pane 4 does not exist: left out
"""
import json
import math
import os


def r5(vals, q):
    if not vals:
        return None                       # NaN in the tables; JSON null here
    v = sorted(vals)
    n = len(v)
    rank = q * (n - 1)
    lo = int(math.floor(rank))
    hi = min(n - 1, lo + 1)
    w = rank - lo
    return float(v[lo]) * (1.0 - w) + float(v[hi]) * w


SAMPLES = {'ts': [99, 100, 105, 110, 119, 125, 139, 140], 'gpu_gid': [0, 0, 1, 0, 2, 3, 0, 0],
           'pid': [7, 7, 7, 8, 7, 9, 7, 7], 'sm': [50, 10, 20, 30, 40, 60, 70, 80]}
PAIRS = {'node': [0, 0, 1, 0], 'pid': [7, 8, 7, 7], 'pod': [5, 1, 1, 0]}
# hand-worked per-window value lists, cell index = key * 3 + window
CELLS = {
    'pod': {'n_keys': 2, 'values': [[10, 20], [], [70],            # pod0: w0 = panes 0-1, w1 = panes 1-2, w2 = panes 2-3
                                     [30, 40], [30, 40], []]},       # pod1
    'gpu': {'n_keys': 4, 'values': [[10, 30], [30], [70],          # gpu0
                                     [20], [], [],                   # gpu1
                                     [40], [40], [],                 # gpu2
                                     [], [60], [60]]},               # gpu3
    'node': {'n_keys': 2, 'values': [[10, 20, 30], [30], [70],     # node0
                                      [40], [40, 60], [60]]},        # node1
}
QS = [0.0, 0.5, 0.95, 1.0]

out = {'_comment': __doc__, 'gpus_per_node': 2, 't0': 100, 'step': 10, 'width': 20, 'n_windows': 3,
       'samples': SAMPLES, 'pairs': PAIRS, 'qs': QS, 'expect': {}}
for kind, spec in CELLS.items():
    vals = spec['values']
    out['expect'][kind] = {
        'n_keys': spec['n_keys'],
        'count': [len(v) for v in vals],
        'sum': [sum(v) for v in vals],
        'max': [max(v) if v else 0 for v in vals],
        'mean': [float(sum(v)) / float(len(v)) if v else 0.0 for v in vals],
        'pct': [[r5(v, q) for v in vals] for q in QS],
    }
with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'window_kat.json'), 'w') as f:
    json.dump(out, f, indent=1)
print('written')
