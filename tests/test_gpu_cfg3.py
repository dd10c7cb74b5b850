"""BASELINE configs[2]: a flat stream in random order over many distinct (node, pid) pairs with
Zipf-sized pods - the join-cardinality stress.  The reference's counterpart is the dict probe loop
of gpu_util_stats.py:83-92; here every sample is its own run, the hot pods' cells would serialise
on one L2 address each, and k_accum folds runs in a per-CTA accumulator cache first.  The tests
hold the tables to the oracle bit for bit AND check that the cache path was the one that ran."""
import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = {'pod': ws.KEY_POD, 'gpu': ws.KEY_GPU, 'node': ws.KEY_NODE}


def test_cfg3_scaled_all_kinds_p95(engine):
    """configs[2] at 1/50: 2 M samples, 20 k keys, 2 k pods, 250 nodes; pod + gpu + node, 6 windows of
    10 minutes, p95 (every cell far beyond a thread's histogram: the warp-per-cell path)."""
    tr = synth.config(3, scale=0.02)
    kinds = ('pod', 'gpu', 'node')
    engine.set_auto_reorder(False)          # the stream as it comes: one run per sample, accumulator cache, run log retry
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    np.testing.assert_array_equal(engine.probe(8), ws.probe_np(tr.keys, tr.pod_ids, tr.gpu_gid, tr.pid, 8))
    got = engine.window_stats(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=600, n_windows=6,
                              n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=(0.95,))
    assert engine.job_counter('cache_ctas') > 0, 'the per-CTA accumulator cache was not used'
    assert engine.job_counter('big_cells') > 0, 'no cell took the warp-per-cell path'
    for k, nk in (('pod', tr.n_pods), ('gpu', tr.n_gpus), ('node', tr.n_nodes)):
        want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=KINDS[k], metric=0,
                                  gpus_per_node=8, t0=tr.t0, step=600, width=600, n_windows=6, n_keys=nk, qs=(0.95,))
        assert_stats_equal(got[k], want, what='cfg3 x0.02 ' + k)
    # the default: a disordered batch of a percentile job is accumulated from a series-major copy - same tables
    engine.set_auto_reorder(True)
    engine.set_profiling(True)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    again = engine.window_stats(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=600, n_windows=6,
                                n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=(0.95,))
    assert engine.kernel_times()['reorder'][1] > 0, 'the sorted copy was not made'
    for k in kinds:
        assert_stats_equal(again[k], got[k], what='cfg3 x0.02 sorted copy ' + k)


def test_cfg3_scaled_no_percentiles_uses_the_cache(engine):
    """The shape tools/run_cfg3.py times: mean / max per pod over the whole range, no value log."""
    tr = synth.config(3, scale=0.01)
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    got = engine.window_stats(kinds=('pod',), metric='sm', gpus_per_node=8, t0=tr.t0, step=3600, n_windows=1,
                              n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=())
    assert engine.job_counter('cache_ctas') > 0
    want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=ws.KEY_POD, metric=0,
                              gpus_per_node=8, t0=tr.t0, step=3600, width=3600, n_windows=1, n_keys=tr.n_pods)
    assert_stats_equal(got['pod'], want, what='cfg3 x0.01 pod')


def test_cfg3_full_size(engine):
    """configs[2] at FULL size: 100 M samples, 1 M distinct (node, pid), 100 k pods.  Pod attribution of
    every sample and the per-pod count / sum / max / mean against the numpy oracle (exact), plus
    size-independent properties of the per-GPU and per-node tables."""
    tr = synth.config(3)
    assert tr.n == 100_000_000 and len(tr.keys) == 1_000_000
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    pods = engine.probe(8)
    want_pods = ws.probe_np(tr.keys, tr.pod_ids, tr.gpu_gid, tr.pid, 8)
    np.testing.assert_array_equal(pods, want_pods)
    del pods
    got = engine.window_stats(kinds=('pod', 'gpu', 'node'), metric='sm', gpus_per_node=8, t0=tr.t0, step=3600,
                              n_windows=1, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=())
    assert engine.job_counter('cache_ctas') > 0
    v = (tr.util & np.uint32(0xFF)).astype(np.int64)
    cnt = np.bincount(want_pods, minlength=tr.n_pods).astype(np.uint64)
    tot = np.bincount(want_pods, weights=v, minlength=tr.n_pods).astype(np.uint64)
    np.testing.assert_array_equal(got['pod']['count'], cnt)
    np.testing.assert_array_equal(got['pod']['sum'], tot)
    mx = np.zeros(tr.n_pods, dtype=np.int64)
    np.maximum.at(mx, want_pods, v)
    np.testing.assert_array_equal(got['pod']['max'].astype(np.int64), mx)
    mean = np.divide(tot.astype(np.float64), cnt.astype(np.float64), out=np.zeros(tr.n_pods), where=cnt > 0)
    np.testing.assert_array_equal(got['pod']['mean'].view(np.uint64), mean.view(np.uint64))
    assert int(got['gpu']['count'].sum()) == tr.n and int(got['node']['count'].sum()) == tr.n
    assert int(got['gpu']['sum'].sum()) == int(v.sum()) == int(got['node']['sum'].sum())
    g = got['gpu']['count'].reshape(tr.n_nodes, 8)
    np.testing.assert_array_equal(g.sum(axis=1), got['node']['count'])
    print('cfg3 full size: largest pod holds %.1f%% of the samples' % (100.0 * float(cnt.max()) / tr.n))


def test_streamed_big_cells_with_no_hint(engine):
    """Day-wide cells fed by many batches with total_samples_hint = 0 and no resident batch at begin:
    nothing about the job's size is known up front, every per-GPU cell ends far above 65 535 samples.
    The lists of cells left to the slow kernels have one entry per window cell, so none can be dropped."""
    tr = synth.cluster_trace(3, 8, 1, 125_000, n_tenancies=30, seed=21)        # 24 GPUs x 125 000 s = 3 M samples
    kinds = ('pod', 'gpu', 'node')
    qs = (0.5, 0.95, 0.99)
    from kube_gpu_stats_b200.engine import Engine
    with Engine(0) as eng:                                                      # a fresh handle: no resident batch
        eng.build_join(tr.keys, tr.pod_ids, tr.n_pods)
        spec = eng.make_spec(kinds=kinds, metric='sm', gpus_per_node=8, t0=tr.t0, step=125_000, n_windows=1,
                             n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=qs, total_samples_hint=0)
        eng.stats_begin(spec)
        order = np.argsort(tr.ts, kind='stable')                                # a live feed: time slices
        for a in range(0, tr.n, 200_000):
            p = np.sort(order[a:a + 200_000])                                   # each slice in collector order
            eng.ingest(tr.ts[p], tr.gpu_gid[p], tr.pid[p], tr.util[p])
            eng.stats_accumulate()
        eng.stats_finalize()
        assert eng.job_counter('big_cells') >= 24 + 3
        for k, nk in (('pod', tr.n_pods), ('gpu', tr.n_gpus), ('node', tr.n_nodes)):
            want = ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids, key_kind=KINDS[k], metric=0,
                                      gpus_per_node=8, t0=tr.t0, step=125_000, width=125_000, n_windows=1, n_keys=nk, qs=qs)
            assert_stats_equal(eng.fetch(k), want, what='no-hint big cells ' + k)
