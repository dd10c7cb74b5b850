#!/usr/bin/env python
"""Under torchrun: the two transports of the table exchange side by side.

Every rank runs the configs[1] job on its own shard; the packed tables are gathered (a) by the NCCL
all_gather (sharding.TableGather) and (b) by copy-engine pushes into peer-mapped symmetric memory
(sharding.PeerTableGather).  Checks that both give the same bytes on every rank, for serial and
overlapped use, then times `--steps` overlapped steps with each.  Prints one line per transport."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200 import sharding, synth  # noqa: E402
from kube_gpu_stats_b200.engine import Engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--nodes', type=int, default=1000)
ap.add_argument('--steps', type=int, default=50)
args = ap.parse_args()
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
KINDS = ('pod', 'gpu', 'node')
tr = synth.cluster_trace(args.nodes, 8, 1, 3600, n_tenancies=args.nodes * 10, seed=rank)
eng = Engine(local)
eng.set_packed_exchange(True)
eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
eng.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
eng.sync()
spec = eng.make_spec(kinds=KINDS, metric='sm', gpus_per_node=8, t0=tr.t0, step=60, n_windows=60, n_gpus=tr.n_gpus,
                     n_nodes=tr.n_nodes, qs=(0.95,), total_samples_hint=tr.n)
stream = torch.cuda.ExternalStream(eng.stream('compute'), device=torch.device('cuda', local))
nccl = sharding.TableGather(eng, KINDS, world)
peer = sharding.PeerTableGather(eng, KINDS, world)
with torch.cuda.stream(stream):
    # serial: same job, both transports
    for _ in range(3):                                   # both result regions, both parities
        eng.window_stats_device(spec)
        a = nccl.all_gather().clone()
        b = peer.all_gather()
        eng.sync(); torch.cuda.synchronize()
        assert torch.equal(a, b.reshape(-1)), 'peer push differs from the NCCL all_gather (serial)'
    # overlapped: gather of job i runs beside job i+1
    for _ in range(4):
        peer.before_next_job(stream)
        eng.window_stats_device(spec)
        got = peer.all_gather_async(stream)
    peer.wait(); eng.sync(); torch.cuda.synchronize()
    assert torch.equal(a, got.reshape(-1)), 'peer push differs from the NCCL all_gather (overlapped)'
    dist.barrier()
    for name, g in (('nccl all_gather', nccl), ('copy-engine push', peer), ('nccl all_gather', nccl), ('copy-engine push', peer)):
        for _ in range(5):
            g.before_next_job(stream); eng.window_stats_device(spec); g.all_gather_async(stream)
        g.wait(); dist.barrier(); eng.sync(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            g.before_next_job(stream); eng.window_stats_device(spec); g.all_gather_async(stream)
        g.join(stream)
        e1.record(stream)
        g.wait(); dist.barrier(); eng.sync(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device='cuda', dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            print('%-18s %d GPUs: %.4f ms per step, %.1f G samples/s' % (name, world, float(ms), tr.n * world / float(ms) / 1e6), flush=True)
if rank == 0:
    print('exchange check ok: identical bytes from both transports on %d ranks' % world)
eng.close()
dist.destroy_process_group()
