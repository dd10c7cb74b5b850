"""k_finalize_groups (finalize_groups.cuh): the register-resident selection for (gpu, window) cells
of up to 64 samples, their twin (pod, window) cells and the (node, window) cells a group of lanes
holds together - against the oracle, bit for bit, at every edge of its preconditions: group sizes
that are not a power of two or exceed a warp, a GPU outside the per-GPU table whose node is inside
the per-node table, cells of exactly 64 and of 65 samples, values 127 and 128, runs that are not
contiguous in the value log, repeated finalize passes of one job."""
import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = {'pod': ws.KEY_POD, 'gpu': ws.KEY_GPU, 'node': ws.KEY_NODE}
QS = (0.0, 0.5, 0.95, 1.0)


def want_tables(tr, kinds, n_keys, t0, step, n_windows, qs, metric=0, width=None):
    return {k: ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=KINDS[k],
                                  metric=metric, gpus_per_node=tr.gpus_per_node, t0=t0, step=step,
                                  width=step if width is None else width, n_windows=n_windows, n_keys=n_keys[k], qs=qs)
            for k in kinds}


def groups_launches(engine):
    return engine.kernel_times()['finalize_groups'][1]


@pytest.mark.parametrize('gpn', [1, 3, 4, 8, 16, 32, 40])
def test_group_sizes_and_a_gpu_outside_the_gpu_table(engine, gpn):
    """Nodes of `gpn` GPUs; the per-GPU table stops one GPU short of the last node, so that node's
    cells hold samples no lane has (they must fall back to the histogram path and still be exact)."""
    tr = synth.cluster_trace(3, gpn, 1, 260, n_tenancies=3 * gpn + 2 * gpn // 2 + 1, seed=gpn)
    kinds = ('pod', 'gpu', 'node')
    n_keys = {'pod': tr.n_pods, 'gpu': tr.n_gpus - 1, 'node': tr.n_nodes}
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    engine.set_profiling(True)
    got = engine.window_stats(kinds=kinds, metric='sm', gpus_per_node=gpn, t0=tr.t0 + 3, step=20, n_windows=12,
                              n_gpus=n_keys['gpu'], n_nodes=n_keys['node'], qs=QS)
    assert groups_launches(engine) == 1
    want = want_tables(tr, kinds, n_keys, tr.t0 + 3, 20, 12, QS)
    for k in kinds:
        assert_stats_equal(got[k], want[k], what='gpn %d %s' % (gpn, k))


@pytest.mark.parametrize('kinds', [('gpu',), ('gpu', 'node'), ('pod', 'gpu'), ('pod', 'node'), ('pod', 'gpu', 'node')])
def test_every_combination_of_tables(engine, kinds):
    tr = synth.cluster_trace(4, 8, 1, 600, n_tenancies=40, seed=17)
    n_keys = {'pod': tr.n_pods, 'gpu': tr.n_gpus, 'node': tr.n_nodes}
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    got = engine.window_stats(kinds=kinds, metric='mem', gpus_per_node=8, t0=tr.t0, step=60, n_windows=10,
                              n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=(0.95,))
    want = want_tables(tr, kinds, n_keys, tr.t0, 60, 10, (0.95,), metric=1)
    for k in kinds:
        assert_stats_equal(got[k], want[k], what='%s of %s' % (k, kinds))


@pytest.mark.parametrize('step', [63, 64, 65])
def test_cells_of_64_and_65_samples_values_127_and_128(engine, step):
    """64 samples is the most a lane holds, 127 the largest value it can count: one more of either
    sends the cell (and its node cell) to the histogram kernels."""
    rng = np.random.default_rng(step)
    tr = synth.cluster_trace(2, 4, 1, 5 * step, seed=step)
    util = tr.util.copy()
    hot = rng.random(tr.n) < 0.02
    util[hot] = (util[hot] & ~np.uint32(0xFF)) | rng.choice([126, 127, 128, 129, 255], size=int(hot.sum())).astype(np.uint32)
    tr = synth.Trace(tr.ts, tr.gpu_gid, tr.pid, util, tr.keys, tr.pod_ids, tr.n_pods, tr.n_nodes, 4, tr.t0, tr.seconds)
    kinds = ('pod', 'gpu', 'node')
    n_keys = {'pod': tr.n_pods, 'gpu': tr.n_gpus, 'node': tr.n_nodes}
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    got = engine.window_stats(kinds=kinds, metric='sm', gpus_per_node=4, t0=tr.t0, step=step, n_windows=5,
                              n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS)
    want = want_tables(tr, kinds, n_keys, tr.t0, step, 5, QS)
    for k in kinds:
        assert_stats_equal(got[k], want[k], what='step %d %s' % (step, k))


def test_runs_that_are_not_contiguous_and_repeated_finalize(engine):
    """A live feed delivers time slices: a window that straddles two batches has its samples in two
    places of the value log.  The job is finalized after every batch (a running report): each pass
    must be exact for what has arrived, also after the pass counter (a byte) has wrapped."""
    tr = synth.cluster_trace(3, 8, 1, 400, n_tenancies=30, seed=5)
    kinds = ('pod', 'gpu', 'node')
    n_keys = {'pod': tr.n_pods, 'gpu': tr.n_gpus, 'node': tr.n_nodes}
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    spec = engine.make_spec(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=50, n_windows=8,
                            n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=tr.n)
    engine.stats_begin(spec)
    sec = tr.ts - tr.t0
    seen = np.zeros(tr.n, dtype=bool)
    for a, b in ((0, 130), (130, 131), (131, 333), (333, 400)):
        m = (sec >= a) & (sec < b)
        engine.ingest(tr.ts[m], tr.gpu_gid[m], tr.pid[m], tr.util[m])
        engine.stats_accumulate()
        engine.stats_finalize()
        seen |= m
        part = synth.Trace(tr.ts[seen], tr.gpu_gid[seen], tr.pid[seen], tr.util[seen], tr.keys, tr.pod_ids,
                           tr.n_pods, tr.n_nodes, 8, tr.t0, tr.seconds)
        want = want_tables(part, kinds, n_keys, tr.t0, 50, 8, QS)
        for k in kinds:
            assert_stats_equal(engine.fetch(k), want[k], what='after second %d, %s' % (b, k))
    for _ in range(300):                    # the pass stamp is a byte
        engine.stats_finalize()
    for k in kinds:
        assert_stats_equal(engine.fetch(k), want[k], what='after 300 more passes, %s' % k)


def test_groups_path_off_gives_the_same_tables(built_lib, monkeypatch):
    """KGS_FINALIZE_GROUPS=0 leaves every cell to the histogram kernels: same bits."""
    from kube_gpu_stats_b200.engine import Engine
    tr = synth.config(2, scale=0.01)
    kinds = ('pod', 'gpu', 'node')
    kw = dict(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=60, n_windows=60, n_gpus=tr.n_gpus,
              n_nodes=tr.n_nodes, qs=(0.5, 0.95))
    out = []
    for flag in ('1', '0'):
        monkeypatch.setenv('KGS_FINALIZE_GROUPS', flag)
        with Engine(0) as eng:
            eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
            eng.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
            eng.set_profiling(True)
            out.append(eng.window_stats(**kw))
            assert groups_launches(eng) == (1 if flag == '1' else 0)
    for k in kinds:
        assert_stats_equal(out[0][k], out[1][k], what='groups on/off ' + k)
