#!/usr/bin/env python
"""Short driver for ncu: configs[1] resident in HBM, a few steps of the bench's hot path."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200 import synth  # noqa: E402
from kube_gpu_stats_b200.engine import Engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--steps', type=int, default=3)
ap.add_argument('--nodes', type=int, default=1000)
ap.add_argument('--kinds', default='pod,gpu,node')
ap.add_argument('--qs', default='0.95')
ap.add_argument('--procs', type=int, default=1)
ap.add_argument('--series-major', action='store_true', help='reorder to (gpu, pid, time) before ingest')
ap.add_argument('--reorder', action='store_true', help='kgs_reorder_batch on the GPU after ingest')
ap.add_argument('--time-major', action='store_true', help='deliver the trace second by second')
args = ap.parse_args()
tr = synth.cluster_trace(args.nodes, 8, args.procs, 3600, n_tenancies=args.nodes * 8 * args.procs * 5 // 4, seed=0)
if args.series_major:
    from kube_gpu_stats_b200 import ingest
    o = ingest.series_major_order(tr.ts, tr.gpu_gid, tr.pid)
    tr.ts, tr.gpu_gid, tr.pid, tr.util = tr.ts[o], tr.gpu_gid[o], tr.pid[o], tr.util[o]
if args.time_major:
    import numpy as np
    o = np.argsort(tr.ts, kind='stable')
    tr.ts, tr.gpu_gid, tr.pid, tr.util = tr.ts[o], tr.gpu_gid[o], tr.pid[o], tr.util[o]
kinds = tuple(args.kinds.split(','))
qs = tuple(float(q) for q in args.qs.split(',')) if args.qs else ()
with Engine(0) as eng:
    eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    if (args.procs > 1 or args.time_major) and not (args.series_major or args.reorder):
        eng.set_run_log_capacity(8 * tr.n)
    eng.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    eng.set_profiling(True)
    if args.reorder:
        for _ in range(3):
            eng.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
            eng.reorder()
    spec = eng.make_spec(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=60, n_windows=60,
                         n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=qs, total_samples_hint=tr.n)
    eng.set_profiling(True)
    for _ in range(args.steps):
        eng.window_stats_device(spec)
    eng.sync()
    for name, (ms, n) in eng.kernel_times().items():
        if n:
            print('%-14s %8.3f ms avg over %d launches' % (name, ms / n, n))
    out = eng.fetch(kinds[0])
    print('samples', tr.n, 'cells', out['count'].size, 'count sum', int(out['count'].sum()))
