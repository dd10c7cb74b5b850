#!/usr/bin/env python
"""configs[3]/[4] at scale: a node-sharded 10k-node x 24 h trace (6.9 G samples over 8 GPUs).

Each rank builds its shard of `--nodes` nodes x 8 GPUs x `--seconds` s by replicating a seeded base
trace of `--base-nodes` nodes with shifted node ids (so the shard has distinct nodes, pods and
(node,pid) keys), makes it resident in HBM piece by piece, then runs the job:
1-min tumbling mean/max/p95 for pod + gpu + node tables, followed by the all_gather of the tables.
Parity at this size is checked through size-independent properties and an exact oracle comparison
on the nodes of the base trace.  With --sliding it runs configs[4]: width 300 s, step 10 s, p50/p99
for pod + gpu tables (a slice of `--sliding-windows` windows)."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200 import sharding, synth  # noqa: E402
from kube_gpu_stats_b200.engine import Engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--nodes', type=int, default=1250)
ap.add_argument('--base-nodes', type=int, default=50)
ap.add_argument('--seconds', type=int, default=86400)
ap.add_argument('--steps', type=int, default=3)
ap.add_argument('--sliding', action='store_true')
ap.add_argument('--sliding-windows', type=int, default=360)
ap.add_argument('--check', action='store_true')
args = ap.parse_args()

rank = int(os.environ.get('RANK', '0'))
world = int(os.environ.get('WORLD_SIZE', '1'))
local = int(os.environ.get('LOCAL_RANK', '0'))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))

t = time.time()
base = synth.cluster_trace(args.base_nodes, 8, 1, args.seconds, n_tenancies=args.base_nodes * 10, seed=rank)
reps = args.nodes // args.base_nodes
n = base.n * reps
n_pods, n_nodes, n_gpus = base.n_pods * reps, base.n_nodes * reps, base.n_gpus * reps
keys = np.concatenate([base.keys + (np.uint64(r * base.n_nodes) << np.uint64(32)) for r in range(reps)])
pods = np.concatenate([base.pod_ids + np.uint32(r * base.n_pods) for r in range(reps)])
if rank == 0:
    print('shard per rank: %d nodes, %.1f M samples (%.2f GB), %d pods; base trace generated in %.0f s'
          % (n_nodes, n / 1e6, n * 16 / 1e9, n_pods, time.time() - t), flush=True)

eng = Engine(local)
eng.build_join(keys, pods, n_pods)
eng.batch_reserve(n)
stage = [eng.pinned(base.n) for _ in range(4)]
stage[0][:] = base.ts
stage[2][:] = base.pid
stage[3][:] = base.util
t = time.time()
for r in range(reps):
    eng.sync()                                   # the staging buffer is reused
    np.add(base.gpu_gid, np.uint32(r * base.n_gpus), out=stage[1])
    eng.batch_append(*stage)
eng.sync()
if rank == 0:
    print('resident in HBM after %.1f s' % (time.time() - t), flush=True)

if args.sliding:
    kinds, qs = ('pod', 'gpu'), (0.5, 0.99)
    spec = eng.make_spec(kinds=kinds, metric='sm', gpus_per_node=8, t0=base.t0, step=10, width=300,
                         n_windows=args.sliding_windows, n_gpus=n_gpus, n_nodes=n_nodes, qs=qs, total_samples_hint=n)
else:
    kinds, qs = ('pod', 'gpu', 'node'), (0.95,)
    spec = eng.make_spec(kinds=kinds, metric='sm', gpus_per_node=8, t0=base.t0, step=60, n_windows=args.seconds // 60,
                         n_gpus=n_gpus, n_nodes=n_nodes, qs=qs, total_samples_hint=n)
ap_exchange = os.environ.get('KGS_EXCHANGE', 'peer')
gather = None
stream = torch.cuda.ExternalStream(eng.stream('compute'), device=torch.device('cuda', local))
with torch.cuda.stream(stream):
    if world > 1:
        eng.set_packed_exchange('compact')
    eng.window_stats_device(spec)                 # warm-up (allocates the arena)
    eng.sync()
    if world > 1:
        # the library's own exchange (kgs_exchange_push / kgs_exchange_nccl); not overlapped here: one pass, then its exchange
        slot = eng.exchange_region()[1]
        if ap_exchange == 'peer':
            gather = sharding.PeerExchange(eng, world, rank, slot)
        else:
            ids = [eng.nccl_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            eng.nccl_init(ids[0], world, rank)
            gather = sharding.NcclExchange(eng, world, slot)
        if rank == 0:
            print('exchange: %s, %.1f MB per rank and pass' % (type(gather).__name__, slot / 1e6), flush=True)
        gather.push()
        eng.exchange_join()
    eng.sync(); torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    eng.set_profiling(True); eng.reset_kernel_times()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        eng.window_stats_device(spec)
        if gather:
            gather.push()
            eng.exchange_join()
    e1.record(stream)
    eng.sync(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device='cuda', dtype=torch.float64)
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    kt = eng.kernel_times()
    print('%d GPU(s): %.3f ms per pass over %.2f G samples -> %.1f G samples/s  (%.0f GB/s of 16 B samples per GPU)'
          % (world, float(ms), n * world / 1e9, n * world / float(ms) / 1e6, n * 16 / float(ms) / 1e6))
    print('  kernels per pass (ms): ' + ', '.join('%s %.3f' % (k, v[0] / args.steps) for k, v in kt.items() if v[1]))

if args.check and rank == 0:
    tables = eng.fetch_all(kinds)
    n_win = spec.n_windows
    if not args.sliding:
        assert int(tables['node']['count'].sum()) == n and int(tables['gpu']['count'].sum()) == n
        assert int(tables['pod']['count'].sum()) == n          # every (node,pid) of the trace has a pair
        g = {f: np.asarray(tables['gpu'][f]).reshape(n_nodes, 8, n_win) for f in ('count', 'sum', 'max')}
        assert np.array_equal(g['count'].sum(axis=1).reshape(-1), tables['node']['count'])
        assert np.array_equal(g['sum'].sum(axis=1).reshape(-1), tables['node']['sum'])
        assert np.array_equal(g['max'].max(axis=1).reshape(-1), tables['node']['max'])
        p = tables['node']['pct'][0]
        assert np.all((p[tables['node']['count'] > 0] >= 0) & (p[tables['node']['count'] > 0] <= tables['node']['max'][tables['node']['count'] > 0]))
    # exact: the first replica is the base trace itself -> oracle on the base trace
    from oracle import window_stats as ws
    for kind, kid, nk_base in (('pod', ws.KEY_POD, base.n_pods), ('gpu', ws.KEY_GPU, base.n_gpus), ('node', ws.KEY_NODE, base.n_nodes)):
        if kind not in kinds:
            continue
        want = ws.window_stats_np(base.ts, base.gpu_gid, base.pid, base.util, base.keys, base.pod_ids, key_kind=kid, metric=0,
                                  gpus_per_node=8, t0=base.t0, step=spec.step, width=spec.width, n_windows=n_win,
                                  n_keys=nk_base, qs=qs)
        m = nk_base * n_win
        for f in ('count', 'sum', 'max'):
            assert np.array_equal(np.asarray(tables[kind][f][:m]).astype(np.uint64), want[f].astype(np.uint64)), (kind, f)
        assert np.array_equal(np.asarray(tables[kind]['mean'][:m]).view(np.uint64), want['mean'].view(np.uint64)), kind
        assert np.array_equal(np.ascontiguousarray(tables[kind]['pct'][:, :m]).view(np.uint64), want['pct'].view(np.uint64)), kind
    if rank == 0:
        print('parity ok on rank 0: totals, node = sum of its GPUs, and the base replica bit-identical to the oracle')
if world > 1:
    dist.barrier()
eng.close()
if world > 1:
    dist.destroy_process_group()
