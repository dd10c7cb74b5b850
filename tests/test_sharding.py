"""CPU, world_size 2 over gloo: node sharding, localisation, the table exchange and the merge.
The per-shard compute is stood in by the oracle (the GPU engine takes its place on the box);
what is under test is the host logic around it."""
import os

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import sharding, synth
from oracle import window_stats as ws

KW = dict(metric=0, gpus_per_node=8, step=300, width=300, n_windows=12, qs=(0.5, 0.95))


def test_shard_ranges_cover_and_balance():
    for n, w in ((10, 3), (8, 8), (7, 2), (3, 4), (10000, 8)):
        r = sharding.shard_ranges(n, w)
        assert r[0][0] == 0 and r[-1][1] == n and all(a[1] == b[0] for a, b in zip(r, r[1:]))
        sizes = [b - a for a, b in r]
        assert max(sizes) - min(sizes) <= 1
        owner = sharding.shard_of_node(np.arange(n), n, w)
        for rank, (a, b) in enumerate(r):
            assert np.all(owner[a:b] == rank)


def test_pack_unpack_round_trip():
    rng = np.random.default_rng(0)
    t = {'count': rng.integers(0, 9, 12).astype(np.uint64), 'sum': rng.integers(0, 900, 12).astype(np.uint64),
         'max': rng.integers(0, 100, 12).astype(np.uint32), 'mean': rng.random(12), 'pct': rng.random((2, 12))}
    t['pct'][1, 3] = np.nan
    assert_stats_equal(sharding.unpack_tables(sharding.pack_tables(t)), t)


def _worker(rank, world, port, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    tr = synth.cluster_trace(5, 8, 2, 3600, n_tenancies=100, pids_per_pod=2, seed=3)     # 5 nodes over 2 ranks: uneven
    rng = sharding.shard_ranges(tr.n_nodes, world)[rank]
    ts, gid, pid, util, keys, pods, n_pods, pod_sel = sharding.localise(
        (tr.ts, tr.gpu_gid, tr.pid, tr.util), tr.keys, tr.pod_ids, tr.pod_node, rng, tr.gpus_per_node)
    n_nodes = rng[1] - rng[0]
    merged = {}
    for kind, kid, nk in (('pod', ws.KEY_POD, n_pods), ('gpu', ws.KEY_GPU, n_nodes * 8), ('node', ws.KEY_NODE, n_nodes)):
        local = ws.window_stats_np(ts, gid, pid, util, keys, pods, key_kind=kid, t0=tr.t0, n_keys=nk, **KW)
        merged[kind] = sharding.merge_tables(sharding.all_gather_tables(local), KW['n_windows'])
    # the reference's one global reduction (who_use_gpu.py:18-20): GPU totals per type over all pods
    mine = [{'P40': 6, 'V100': 2}, {'V100': 5, 'B200': 8, 'P40': 1}][rank]
    total, per_type = sharding.cluster_totals(mine)
    assert total == 22 and list(per_type.items()) == [('P40', 7), ('V100', 7), ('B200', 8)]
    assert sharding.cluster_totals({}) == (0, {})
    if rank == 0:
        np.savez(os.path.join(out_dir, 'merged.npz'), **{'%s_%s' % (k, f): v for k, t in merged.items() for f, v in t.items()})
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_gather_merge_equals_global(tmp_path):
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), 'merged.npz'))
    tr = synth.cluster_trace(5, 8, 2, 3600, n_tenancies=100, pids_per_pod=2, seed=3)
    for kind, kid, nk in (('pod', ws.KEY_POD, tr.n_pods), ('gpu', ws.KEY_GPU, tr.n_gpus), ('node', ws.KEY_NODE, tr.n_nodes)):
        want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=kid, t0=tr.t0,
                                  n_keys=nk, **KW)
        assert_stats_equal({f: got['%s_%s' % (kind, f)] for f in ('count', 'sum', 'max', 'mean', 'pct')}, want, kind)
