#!/usr/bin/env python
"""configs[2]: 100M samples in random order over 1M distinct (node,pid), Zipf pod sizes - join
cardinality stress.  Mean/max per pod over one whole-range window; parity vs the numpy oracle."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200 import synth  # noqa: E402
from kube_gpu_stats_b200.engine import Engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--scale', type=float, default=1.0)
ap.add_argument('--steps', type=int, default=5)
ap.add_argument('--check', action='store_true')
args = ap.parse_args()
t = time.time()
tr = synth.config(3, scale=args.scale)
print('generated %d samples, %d keys, %d pods in %.1f s' % (tr.n, len(tr.keys), tr.n_pods, time.time() - t))
with Engine(0) as eng:
    eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    eng.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    eng.sync()
    eng.set_profiling(True)
    for _ in range(args.steps):
        pods = eng.probe(8)
    spec = eng.make_spec(kinds=('pod',), metric='sm', gpus_per_node=8, t0=tr.t0, step=3600, n_windows=1,
                         n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=(), total_samples_hint=tr.n)
    for _ in range(args.steps):
        eng.window_stats_device(spec)
    eng.sync()
    for name, (ms, n) in eng.kernel_times().items():
        if n:
            print('%-14s %8.3f ms avg over %d launches  (%.1f G samples/s)' % (name, ms / n, n, tr.n / (ms / n) / 1e6))
    got = eng.fetch('pod')
    if args.check:
        from oracle import window_stats as ws
        assert np.array_equal(pods, ws.probe_np(tr.keys, tr.pod_ids, tr.gpu_gid, tr.pid, 8))
        want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=0, metric=0,
                                  gpus_per_node=8, t0=tr.t0, step=3600, width=3600, n_windows=1, n_keys=tr.n_pods)
        for k in ('count', 'sum', 'max'):
            assert np.array_equal(got[k].astype(np.uint64), want[k].astype(np.uint64)), k
        assert np.array_equal(got['mean'].view(np.uint64), want['mean'].view(np.uint64))
        print('parity ok: pod attribution and per-pod count/sum/max/mean identical to the oracle')
    print('largest pod holds %.1f%% of the samples' % (100.0 * got['count'].max() / tr.n))
