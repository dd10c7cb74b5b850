#!/usr/bin/env python
"""Trace size at which the GPU path overtakes the CPU path, transfers included (north_star).

For growing prefixes of a configs[1]-style trace, time (a) the pure-Python oracle (the reference's
style, 1 core) and (b) the GPU end to end through the public API from pinned host columns:
table build + H2D + join + three tables with p95 + D2H of every table.  Prints a markdown table
and the crossover size (first size where the GPU wins, log-interpolated)."""
import math
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200 import synth  # noqa: E402
from kube_gpu_stats_b200.engine import Engine  # noqa: E402
from oracle import window_stats as ws  # noqa: E402

KINDS = ('pod', 'gpu', 'node')


def measure(sizes=(8, 16, 32, 64, 256, 1024, 4096, 16384, 65536, 262144), cpu_cap=70000):
    tr = synth.cluster_trace(10, 8, 1, 3600, n_tenancies=100, seed=0)
    rows = []
    with Engine(0) as eng:
        host = [eng.pinned(max(sizes)) for _ in range(4)]
        # time-major prefix: the first n samples by time across all GPUs (a short live window)
        order = np.argsort(tr.ts, kind='stable')[:max(sizes)]
        for dst, src in zip(host, (tr.ts[order], tr.gpu_gid[order], tr.pid[order], tr.util[order])):
            dst[:] = src
        table = ws.build_table(tr.keys, tr.pod_ids)
        for n in sizes:
            cols = [h[:n] for h in host]
            n_win = int((int(cols[0].max()) - tr.t0) // 60 + 1)
            kw = dict(metric='sm', gpus_per_node=8, t0=tr.t0, step=60, n_windows=n_win, n_gpus=tr.n_gpus,
                      n_nodes=tr.n_nodes, qs=(0.95,), total_samples_hint=n)
            eng.set_run_log_capacity(6 * (n + 8192))

            def gpu_once():
                eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
                eng.ingest(*cols)
                eng.window_stats_device(eng.make_spec(kinds=KINDS, **kw))
                return eng.fetch_all(KINDS)
            for _ in range(3):
                gpu_once()
            reps = 20 if n <= 65536 else 5
            t = time.perf_counter()
            for _ in range(reps):
                gpu_once()
            gpu_s = (time.perf_counter() - t) / reps
            cpu_s = None
            if n <= cpu_cap:
                t = time.perf_counter()
                for kind, nk in ((ws.KEY_POD, tr.n_pods), (ws.KEY_GPU, tr.n_gpus), (ws.KEY_NODE, tr.n_nodes)):
                    ws.window_stats_py(cols[0], cols[1], cols[2], cols[3], table, key_kind=kind, metric=0, gpus_per_node=8,
                                       t0=tr.t0, step=60, width=60, n_windows=n_win, n_keys=nk, qs=(0.95,))
                cpu_s = time.perf_counter() - t
            rows.append((n, gpu_s, cpu_s))
    cross = None
    for (n0, g0, c0), (n1, g1, c1) in zip(rows, rows[1:]):
        if c0 is not None and c1 is not None and c0 < g0 and c1 >= g1:
            # log-log interpolation of cpu/gpu ratio = 1
            r0, r1 = math.log(c0 / g0), math.log(c1 / g1)
            cross = int(round(math.exp(math.log(n0) + (0 - r0) / (r1 - r0) * (math.log(n1) - math.log(n0)))))
    if cross is None and rows[0][2] is not None and rows[0][2] >= rows[0][1]:
        cross = rows[0][0]
    return rows, cross


if __name__ == '__main__':
    rows, cross = measure()
    print('| samples | GPU end to end (ms) | CPU oracle, 1 core (ms) | CPU / GPU |')
    print('|---:|---:|---:|---:|')
    for n, g, c in rows:
        print('| %d | %.3f | %s | %s |' % (n, g * 1e3, '%.3f' % (c * 1e3) if c else 'n/a', '%.1f' % (c / g) if c else 'n/a'))
    print('\ncrossover: the GPU path (transfers included) is faster from about %s samples' % cross)
