"""Two (or more) GPUs, one process each: every rank computes the tables of its node shard with the
CUDA path, the ranks exchange the packed tables GPU to GPU - through the library's C entry points
(kgs_exchange_push: copy-engine pushes into peer memory + barrier kernel; kgs_exchange_nccl: one
ncclAllGather on the library's own and on torch's communicator) and through the Python-driven
transports of round 1 - every transport must deliver identical bytes, and the merged tables are held
to the oracle on the GLOBAL trace, bit for bit.  Skipped on a box with one GPU (tests/test_sharding.py covers the host logic over gloo)."""
import os

import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = ('pod', 'gpu', 'node')
QS = (0.5, 0.95)
STEP, N_WINDOWS = 60, 20


def _n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _trace():
    return synth.cluster_trace(7, 8, 1, 1200, n_tenancies=90, seed=8)      # 7 nodes: uneven shards


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from kube_gpu_stats_b200 import sharding
    from kube_gpu_stats_b200.engine import Engine
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    tr = _trace()
    # equal shard sizes are what the peer exchange needs: every rank is given the table shape of the
    # largest shard (its surplus keys stay empty)
    ranges = sharding.shard_ranges(tr.n_nodes, world)
    nodes_max = max(b - a for a, b in ranges)
    a, b = ranges[rank]
    ts, gid, pid, util, keys, pods, n_pods, pod_sel = sharding.localise(
        (tr.ts, tr.gpu_gid, tr.pid, tr.util), tr.keys, tr.pod_ids, tr.pod_node, (a, b), 8)
    pods_max = max(int(np.count_nonzero((tr.pod_node >= x) & (tr.pod_node < y))) for x, y in ranges)
    eng = Engine(rank)
    eng.set_packed_exchange('compact')
    eng.build_join(keys, pods, pods_max)
    eng.ingest(ts, gid, pid, util)
    spec = eng.make_spec(kinds=KINDS, metric='sm', gpus_per_node=8, t0=tr.t0, step=STEP, n_windows=N_WINDOWS,
                         n_gpus=nodes_max * 8, n_nodes=nodes_max, qs=QS, total_samples_hint=len(ts))
    stream = torch.cuda.ExternalStream(eng.stream('compute'), device=torch.device('cuda', rank))
    results = {}
    with torch.cuda.stream(stream):
        eng.window_stats_device(spec)
        eng.sync()
        nbytes = eng.exchange_region()[1]
        # (1) the Python-driven transports of round 1
        for name, cls in (('peer', sharding.PeerTableGather), ('nccl', sharding.TableGather)):
            gather = cls(eng, KINDS, world)
            dst = gather.all_gather()
            eng.sync()
            torch.cuda.synchronize()
            results[name] = dst.reshape(world, -1)[:, :nbytes].cpu().numpy()
        # (2) the library's own exchange through the C ABI: peer pushes, three jobs in a row (both buffer
        # halves and the reuse of a result region two jobs later), then NCCL on the library's own
        # communicator and on torch's
        px = sharding.PeerExchange(eng, world, rank, nbytes)
        for i in range(3):
            eng.window_stats_device(spec)
            view = px.push()
        eng.exchange_sync()
        results['c_peer'] = view.cpu().numpy()
        # without the device-local copy: the peers' rows arrive as before, the own row is left to the result region
        nx = sharding.PeerExchange(eng, world, rank, nbytes, self_copy=False)
        eng.window_stats_device(spec)
        view = nx.push()
        eng.exchange_sync()
        got = view.cpu().numpy()
        others = [r for r in range(world) if r != rank]
        np.testing.assert_array_equal(got[others], results['peer'][others])
        assert not got[rank].any(), 'the own slot must stay untouched'
        own = torch.as_tensor(sharding._DeviceBytes(eng.exchange_region()[0], nbytes), device='cuda:%d' % rank).cpu().numpy()
        np.testing.assert_array_equal(own, results['peer'][rank])
        eng.exchange_set_self_copy(True)
        if os.environ.get('KGS_TEST_MULTICAST') == '1':     # one copy through the NVSwitch multicast address
            mx = sharding.PeerExchange(eng, world, rank, nbytes, multicast=True)
            for i in range(3):
                eng.window_stats_device(spec)
                view = mx.push()
            eng.exchange_sync()
            results['c_multicast' if mx.multicast else 'c_peer2'] = view.cpu().numpy()
        ids = [eng.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        eng.nccl_init(ids[0], world, rank)
        own = sharding.NcclExchange(eng, world, nbytes)
        eng.window_stats_device(spec)
        view = own.push()
        eng.exchange_sync()
        results['c_nccl_own'] = view.cpu().numpy()
        dist.barrier()
        torch.cuda.synchronize()
        theirs = sharding.NcclExchange(eng, world, nbytes, comm=sharding.NcclExchange.torch_comm(rank))
        eng.window_stats_device(spec)
        view = theirs.push()
        eng.exchange_sync()
        results['c_nccl_torch'] = view.cpu().numpy()
    layout = eng.packed_layout(KINDS)
    for name in results:
        np.testing.assert_array_equal(results['peer'], results[name], err_msg=name)
    if rank == 0:
        print('transports with identical bytes:', sorted(results))
    if rank == 0:
        per_rank = [sharding.decode_packed(results['peer'][r], layout) for r in range(world)]
        np.savez(os.path.join(out_dir, 'tables.npz'),
                 **{'%d_%s_%s' % (r, k, f): v for r, t in enumerate(per_rank) for k, kt in t.items() for f, v in kt.items()})
    dist.barrier()
    eng.close()
    dist.destroy_process_group()


@pytest.mark.skipif(_n_gpus() < 2, reason='needs two GPUs')
def test_shards_on_separate_gpus_exchange_and_merge_to_the_global_tables(tmp_path, built_lib):
    import torch.multiprocessing as mp
    from kube_gpu_stats_b200 import sharding
    world = 2
    port = 29700 + os.getpid() % 200
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    got = np.load(os.path.join(str(tmp_path), 'tables.npz'))
    tr = _trace()
    ranges = sharding.shard_ranges(tr.n_nodes, world)
    nodes_max = max(b - a for a, b in ranges)
    for kind, kid, nk_global in (('pod', ws.KEY_POD, tr.n_pods), ('gpu', ws.KEY_GPU, tr.n_gpus), ('node', ws.KEY_NODE, tr.n_nodes)):
        want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=kid, metric=0,
                                  gpus_per_node=8, t0=tr.t0, step=STEP, width=STEP, n_windows=N_WINDOWS, n_keys=nk_global, qs=QS)
        parts = []
        for r, (a, b) in enumerate(ranges):     # cut every shard's table back to the keys it really owns
            own = {'pod': int(np.count_nonzero((tr.pod_node >= a) & (tr.pod_node < b))), 'gpu': (b - a) * 8, 'node': b - a}[kind]
            t = {f: got['%d_%s_%s' % (r, kind, f)] for f in ('count', 'sum', 'max', 'mean', 'pct')}
            parts.append({f: (v[:, :own * N_WINDOWS] if f == 'pct' else v[:own * N_WINDOWS]) for f, v in t.items()})
            rest = t['count'][own * N_WINDOWS:]
            assert not rest.any(), 'keys beyond the shard must stay empty'
        merged = sharding.merge_tables(parts, N_WINDOWS)
        assert_stats_equal(merged, want, what='2 GPUs merged ' + kind)
