"""Entry points of include/kgs.h that no other GPU test reaches: batches built piecewise
(kgs_batch_reserve / kgs_batch_append), columns that already live in device memory
(kgs_adopt_device_batch), the device-time marks (kgs_mark / kgs_elapsed_ms), and the second
aggregation Prometheus evaluates for the reference, ``max(...) by (node, pod)``
(gpu_util_stats.py:137), emulated on the GPU."""
import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = {'pod': ws.KEY_POD, 'gpu': ws.KEY_GPU, 'node': ws.KEY_NODE}
QS = (0.5, 0.95)


def _want(tr, kind, nk):
    return ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=KINDS[kind], metric=0,
                              gpus_per_node=8, t0=tr.t0, step=60, width=60, n_windows=20, n_keys=nk, qs=QS)


def _job(engine, tr):
    return engine.window_stats(kinds=('pod', 'gpu', 'node'), metric='sm', gpus_per_node=8, t0=tr.t0, step=60,
                               n_windows=20, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS)


def test_batch_built_piecewise_equals_one_ingest(engine):
    from kube_gpu_stats_b200 import _ffi
    from kube_gpu_stats_b200.engine import KgsError
    tr = synth.cluster_trace(3, 8, 1, 1200, n_tenancies=30, seed=2)
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.batch_reserve(tr.n)
    for a, b in ((0, 1), (1, 4097), (4097, 4097), (4097, 20000), (20000, tr.n)):      # ragged pieces, one empty
        engine.batch_append(tr.ts[a:b], tr.gpu_gid[a:b], tr.pid[a:b], tr.util[a:b])
    assert engine.n_resident == tr.n
    for got, src in zip(engine.fetch_batch(), (tr.ts, tr.gpu_gid, tr.pid, tr.util)):
        np.testing.assert_array_equal(got, src)
    got = _job(engine, tr)
    for k, nk in (('pod', tr.n_pods), ('gpu', tr.n_gpus), ('node', tr.n_nodes)):
        assert_stats_equal(got[k], _want(tr, k, nk), what='piecewise ' + k)
    with pytest.raises(KgsError) as e:                                               # beyond the reserved capacity
        engine.batch_append(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    assert e.value.code == _ffi.KGS_ERR_CAPACITY


def test_adopted_device_columns(engine):
    """The caller's own device arrays (torch tensors here) as the resident batch, without a copy."""
    import torch
    from kube_gpu_stats_b200 import _ffi
    from kube_gpu_stats_b200.engine import KgsError
    tr = synth.cluster_trace(2, 8, 1, 1200, n_tenancies=20, seed=4)
    dev = torch.device('cuda', 0)
    cols = [torch.from_numpy(c.view(np.int32)).to(dev) for c in (tr.ts, tr.gpu_gid, tr.pid, tr.util)]
    torch.cuda.synchronize()
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.adopt_device_batch(*[c.data_ptr() for c in cols], tr.n)
    np.testing.assert_array_equal(engine.probe(8), ws.probe_np(tr.keys, tr.pod_ids, tr.gpu_gid, tr.pid, 8))
    got = _job(engine, tr)
    for k, nk in (('pod', tr.n_pods), ('gpu', tr.n_gpus), ('node', tr.n_nodes)):
        assert_stats_equal(got[k], _want(tr, k, nk), what='adopted ' + k)
    with pytest.raises(KgsError) as e:                                               # columns must be 32-byte aligned
        engine.adopt_device_batch(cols[0].data_ptr() + 4, cols[1].data_ptr(), cols[2].data_ptr(), cols[3].data_ptr(), 8)
    assert e.value.code == _ffi.KGS_ERR_ARG


def test_marks_bracket_device_time(engine):
    tr = synth.cluster_trace(2, 8, 1, 1200, n_tenancies=20, seed=4)
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    engine.sync()
    spec = engine.make_spec(kinds=('pod', 'gpu', 'node'), metric='sm', gpus_per_node=8, t0=tr.t0, step=60,
                            n_windows=20, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=tr.n)
    engine.mark(0)
    engine.mark(1)
    empty = engine.elapsed_ms()
    engine.mark(0)
    for _ in range(20):
        engine.window_stats_device(spec)
    engine.mark(1)
    busy = engine.elapsed_ms()
    assert 0.0 <= empty < 0.5 and busy > empty and busy < 1000.0


def test_promql_max_by_node_pod(engine):
    """``max(kube_pod_container_resource_requests...) by (node, pod)`` of gpu_util_stats.py:137, taken
    at the instants of a range query: sampled windows of one second, statistic max, rendered as the
    matrix JSON the script reads (first sample only, gpu_util_stats.py:142)."""
    from kube_gpu_stats_b200 import promql
    tr = synth.cluster_trace(3, 8, 2, 1800, n_tenancies=60, pids_per_pod=2, seed=6)    # pods of two processes
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    step, n_steps = 300, 6
    table = promql.instant_by_pod(engine, tr.t0, step, n_steps)
    labels = [{'node': 'n%d' % tr.pod_node[p], 'pod': 'p%d' % p} for p in range(tr.n_pods)]
    got = promql.matrix_json(table, tr.t0, step, n_steps, labels, stat='max')
    # the same, by hand: per pod and instant the max over the pod's processes of the sample AT that second
    pod_of = ws.probe_np(tr.keys, tr.pod_ids, tr.gpu_gid, tr.pid, 8)
    sm = (tr.util & 0xFF).astype(np.int64)
    want = []
    for p in range(tr.n_pods):
        pts = []
        for k in range(n_steps):
            m = (pod_of == p) & (tr.ts == tr.t0 + k * step)
            if m.any():
                pts.append([int(tr.t0 + k * step), repr(float(sm[m].max()))])
        want.append({'metric': labels[p], 'values': pts})
    assert got['data']['result'] == want
    assert any(len(r['values']) for r in want)
