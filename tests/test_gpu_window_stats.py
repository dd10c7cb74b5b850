"""Windowed statistics through the C ABI vs the oracle (rules R1-R6).  Integer results and
percentiles are bit-exact; mean is an exact integer sum with one correctly rounded divide, so
it is compared bit for bit as well (tolerance 0, tighter than north_star's 1e-12)."""
import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = {'pod': ws.KEY_POD, 'gpu': ws.KEY_GPU, 'node': ws.KEY_NODE}
METRICS = {'sm': 0, 'mem': 1, 'fb': 2}
QS = (0.5, 0.95, 0.99)


def n_keys_of(tr, kind):
    return {'pod': tr.n_pods, 'gpu': tr.n_gpus, 'node': tr.n_nodes}[kind]


def oracle(tr, kind, metric, t0, step, width, n_windows, qs):
    return ws.window_stats_np(tr.ts, tr.gpu_gid, tr.pid, tr.util, tr.keys, tr.pod_ids,
                              key_kind=KINDS[kind], metric=METRICS[metric], gpus_per_node=tr.gpus_per_node,
                              t0=t0, step=step, width=width, n_windows=n_windows,
                              n_keys=n_keys_of(tr, kind), qs=qs)


def run(engine, tr, kinds, metric, t0, step, width, n_windows, qs, ingest=True):
    if ingest:
        engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
        engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    return engine.window_stats(kinds=kinds, metric=metric, gpus_per_node=tr.gpus_per_node, t0=t0, step=step,
                               width=width, n_windows=n_windows, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=qs)


def check(engine, tr, kinds, metric, t0, step, width, n_windows, qs, ingest=True):
    got = run(engine, tr, kinds, metric, t0, step, width, n_windows, qs, ingest)
    for k in kinds:
        assert_stats_equal(got[k], oracle(tr, k, metric, t0, step, width, n_windows, qs), what='%s/%s' % (k, metric))


@pytest.fixture(scope='module')
def cfg1():
    return synth.config(1)


def test_cfg1_tumbling_all_kinds(engine, cfg1):
    """BASELINE configs[0]: 1 node x 8 GPUs x 8 procs, 32 pods, 1 h at 1 s; 1-min windows."""
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0, 60, 60, 60, QS)


def test_cfg1_mem_metric_single_kinds(engine, cfg1):
    for kind in ('pod', 'gpu', 'node'):
        check(engine, cfg1, (kind,), 'mem', cfg1.t0, 60, 60, 60, (0.95,), ingest=(kind == 'pod'))


def test_cfg1_small_python_statement(engine, cfg1):
    """The loop statement of the rules (window_stats_py) on a slice small enough for it."""
    n = 20000
    tr = synth.Trace(cfg1.ts[:n], cfg1.gpu_gid[:n], cfg1.pid[:n], cfg1.util[:n], cfg1.keys, cfg1.pod_ids,
                     cfg1.n_pods, cfg1.n_nodes, 8, cfg1.t0, 3600)
    got = run(engine, tr, ('pod',), 'sm', tr.t0, 60, 60, 6, QS)['pod']
    table = ws.build_table(tr.keys, tr.pod_ids)
    want = ws.window_stats_py(tr.ts, tr.gpu_gid, tr.pid, tr.util, table, key_kind=ws.KEY_POD, metric=0,
                              gpus_per_node=8, t0=tr.t0, step=60, width=60, n_windows=6, n_keys=tr.n_pods, qs=QS)
    assert_stats_equal(got, {k: np.array(v) for k, v in want.items()}, what='py statement')


def test_cfg2_slice(engine):
    """BASELINE configs[1] shape at 1/50 scale: 20 nodes x 8 GPUs x 1 h, 200 pods (tenancy churn)."""
    tr = synth.config(2, scale=0.02)
    check(engine, tr, ('pod', 'gpu', 'node'), 'sm', tr.t0, 60, 60, 60, (0.95,))


def test_mean_max_only_and_fb_metric(engine, cfg1):
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0, 60, 60, 60, ())
    check(engine, cfg1, ('pod', 'node'), 'fb', cfg1.t0, 300, 300, 12, (), ingest=False)


def test_fb_mib_percentiles_two_digit_radix_select(engine, cfg1):
    """16-bit metric: percentiles by the two-pass (high byte, low byte) radix select; bit-exact."""
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'fb', cfg1.t0, 300, 300, 12, QS)
    check(engine, cfg1, ('gpu',), 'fb', cfg1.t0, 10, 50, 100, (0.0, 0.5, 1.0), ingest=False)    # sliding
    rng = np.random.default_rng(9)                                                               # full 16-bit domain
    n = 5000
    tr = synth.Trace(np.sort(rng.integers(100, 400, n)).astype(np.uint32), rng.integers(0, 4, n).astype(np.uint32),
                     rng.integers(1, 3, n).astype(np.uint32), rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32),
                     np.array([1, 2], np.uint64), np.array([0, 1], np.uint32), 2, 1, 8, 100, 300)
    check(engine, tr, ('pod', 'gpu', 'node'), 'fb', 100, 60, 60, 5, (0.01, 0.5, 0.95, 0.99))


def test_window_edges_and_odd_step(engine, cfg1):
    """t0 inside the trace, a step that divides nothing, fewer windows than the trace covers:
    samples before t0 and after the last window are left out (R3)."""
    check(engine, cfg1, ('pod', 'gpu'), 'sm', cfg1.t0 + 123, 7, 7, 100, QS)
    check(engine, cfg1, ('gpu',), 'sm', cfg1.t0 + 5000, 60, 60, 4, QS, ingest=False)      # nothing in range
    check(engine, cfg1, ('node',), 'sm', cfg1.t0 - 100, 1, 1, 200, (0.5,), ingest=False)  # step 1, empty head


def test_sliding_windows(engine, cfg1):
    """configs[4] shape in small: width 5 panes, step 10 s."""
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0, 10, 50, 356, (0.5, 0.99))


def test_sampled_windows_and_promql_instants(engine, cfg1):
    """width < step (R3): the instants a PromQL range query evaluates.  Also the avg-by-pod emulation
    of gpu_util_stats.py:159 against the host rendering used for the replay fixtures."""
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0 + 7, 60, 5, 59, (0.5, 0.95))
    from kube_gpu_stats_b200 import promql, replay
    tr = synth.config(2, scale=0.01)
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    step, n_steps = 300, 12
    table = promql.instant_by_pod(engine, tr.t0, step, n_steps)
    node_name, pod_name = replay.names(tr)
    labels = [{'kubernetes_io_hostname': node_name[int(tr.pod_node[p])], 'nvidia_gpu_type': 'P40', 'pod_name': pod_name[p]}
              for p in range(tr.n_pods)]
    got = promql.matrix_json(table, tr.t0, step, n_steps, labels)
    want = replay.prometheus_answers(tr, step)['sm_util']
    assert got == want


def test_whole_range_big_cells(engine, cfg1):
    """One window over the hour: the node cell holds 230 400 samples (> 65 535: the u32-bin path)."""
    check(engine, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0, 3600, 3600, 1, QS)


def test_shuffled_order(engine, cfg1):
    """Any input order must give the same tables (runs of length ~1: ragged path, run-log retry)."""
    rng = np.random.default_rng(5)
    n = 60000
    idx = rng.permutation(cfg1.n)[:n]
    tr = synth.Trace(cfg1.ts[idx], cfg1.gpu_gid[idx], cfg1.pid[idx], cfg1.util[idx], cfg1.keys, cfg1.pod_ids,
                     cfg1.n_pods, cfg1.n_nodes, 8, cfg1.t0, 3600)
    check(engine, tr, ('pod', 'gpu', 'node'), 'sm', tr.t0, 60, 60, 60, QS)


@pytest.mark.parametrize('n', [0, 1, 3, 4, 5, 127, 128, 129, 1000])
def test_ragged_sizes_full_u8_domain(engine, n):
    """Tiny and ragged batches, values over the whole 0..255 byte domain, unmatched pids."""
    rng = np.random.default_rng(n)
    gid = np.sort(rng.integers(0, 16, size=n)).astype(np.uint32)
    ts = (1000 + np.arange(n) // 3).astype(np.uint32)
    pid = rng.integers(1, 4, size=n).astype(np.uint32)
    util = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    keys = np.array([(0 << 32) | 1, (0 << 32) | 2, (1 << 32) | 1, (1 << 32) | 3], np.uint64)
    pods = np.array([0, 1, 2, 1], np.uint32)
    tr = synth.Trace(ts, gid, pid, util, keys, pods, 3, 2, 8, 1000, 400)
    check(engine, tr, ('pod', 'gpu', 'node'), 'mem', 1000, 5, 10, 20, (0.0, 0.5, 1.0))


def test_streamed_batches_equal_one_batch(engine, cfg1):
    """begin / (ingest, accumulate)* / finalize over 7 uneven batches == the one-shot job."""
    tr = cfg1
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.set_run_log_capacity(2 * 3 * (tr.n + 8 * 4096))   # 8 interleaved processes per GPU: every sample is its own run
    spec = engine.make_spec(kinds=('pod', 'gpu', 'node'), metric='sm', gpus_per_node=8, t0=tr.t0, step=60,
                            n_windows=60, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=tr.n)
    engine.stats_begin(spec)
    cuts = [0, 1001, 50000, 50003, 120000, 200001, 230399, tr.n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        engine.ingest(tr.ts[a:b], tr.gpu_gid[a:b], tr.pid[a:b], tr.util[a:b])
        engine.stats_accumulate()
    engine.stats_finalize()
    for k in ('pod', 'gpu', 'node'):
        assert_stats_equal(engine.fetch(k), oracle(tr, k, 'sm', tr.t0, 60, 60, 60, QS), what='streamed ' + k)


def test_fetch_all_single_copy_equals_per_table_fetch(engine, cfg1):
    """The one-transfer fetch of the result region (pinned staging) against the per-table fetch."""
    tr = cfg1
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    engine.set_run_log_capacity(2 * 3 * (tr.n + 8 * 4096))
    spec = engine.make_spec(kinds=('pod', 'gpu', 'node'), metric='mem', gpus_per_node=8, t0=tr.t0, step=60,
                            n_windows=60, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=tr.n)
    engine.window_stats_device(spec)
    for with_mean in (True, False):
        both = engine.fetch_all(('pod', 'gpu', 'node'), with_mean=with_mean)
        for k in ('pod', 'gpu', 'node'):
            assert_stats_equal({n: np.array(v) for n, v in both[k].items()}, engine.fetch(k), what='fetch_all ' + k)
            assert_stats_equal(engine.fetch(k), oracle(tr, k, 'mem', tr.t0, 60, 60, 60, QS), what='mem ' + k)


@pytest.mark.parametrize('mode', [1, 2])
@pytest.mark.parametrize('metric,step,width', [('sm', 60, 60), ('mem', 10, 50), ('fb', 300, 300), ('sm', 1800, 1800)])
def test_packed_exchange_tables_decode_to_the_full_tables(engine, cfg1, metric, step, width, mode):
    """What a peer GPU receives (packed: 10 + 2*n_q bytes per cell; compact: 6 + 2*n_q, with the cells
    above 65 534 samples - the node cells of the half-hour windows here - on the escape list) must
    rebuild the sender's tables bit for bit."""
    import torch
    from kube_gpu_stats_b200 import sharding
    tr = cfg1
    kinds = ('pod', 'gpu', 'node')
    engine.set_packed_exchange(mode)
    engine.set_run_log_capacity(2 * 3 * (tr.n + 8 * 4096))
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    engine.ingest(tr.ts, tr.gpu_gid, tr.pid, tr.util)
    spec = engine.make_spec(kinds=kinds, metric=metric, gpus_per_node=8, t0=tr.t0, step=step, width=width,
                            n_windows=40, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=tr.n)
    engine.window_stats_device(spec)
    full = {k: engine.fetch(k) for k in kinds}
    ptr, nbytes, packed = engine.exchange_region()
    # fb sums may pass 2^32: u64; the compact tables are for the 8-bit metrics, an fb job falls back to packed
    assert packed == (2 if metric == 'fb' else (3 if mode == 2 else 1)) and nbytes < engine.result_region()[1]
    if step == 1800:
        assert int(full['node']['count'].max()) > 65534
    src = torch.as_tensor(sharding._DeviceBytes(ptr, nbytes), device='cuda:0')
    got = sharding.decode_packed(src.cpu().numpy(), engine.packed_layout(kinds))
    for k in kinds:
        assert_stats_equal(got[k], full[k], what='packed %s %s' % (metric, k))
        assert_stats_equal(full[k], oracle(tr, k, metric, tr.t0, step, width, 40, QS), what='oracle %s %s' % (metric, k))


@pytest.mark.parametrize('order', ['collector', 'time_major'])
def test_many_tiny_batches_grow_the_logs(engine, order):
    """A live feed: hundreds of small batches and a total_samples_hint that is far too small.  The run
    log and the value log have to grow (entries keep their indices) and the tables stay exact.
    Time-major delivery (all GPUs of one second together) makes every sample its own run: such a
    feed declares it with kgs_set_run_log_capacity, which sizes the per-warp slabs for the worst case."""
    tr = synth.cluster_trace(2, 8, 1, 3000, n_tenancies=20, seed=11)
    ts, gid, pid, util = tr.ts, tr.gpu_gid, tr.pid, tr.util
    if order == 'time_major':
        p = np.argsort(tr.ts, kind='stable')
        ts, gid, pid, util = ts[p], gid[p], pid[p], util[p]
        engine.set_run_log_capacity(4096)                        # any value: slabs for the worst case, the log grows
    live = synth.Trace(ts, gid, pid, util, tr.keys, tr.pod_ids, tr.n_pods, tr.n_nodes, 8, tr.t0, 3000)
    engine.build_join(tr.keys, tr.pod_ids, tr.n_pods)
    spec = engine.make_spec(kinds=('pod', 'gpu', 'node'), metric='sm', gpus_per_node=8, t0=tr.t0, step=60,
                            n_windows=50, n_gpus=tr.n_gpus, n_nodes=tr.n_nodes, qs=QS, total_samples_hint=1000)
    engine.stats_begin(spec)
    for a in range(0, live.n, 160):                              # 300 batches
        b = min(live.n, a + 160)
        engine.ingest(ts[a:b], gid[a:b], pid[a:b], util[a:b])
        engine.stats_accumulate()
    engine.stats_finalize()
    for k in ('pod', 'gpu', 'node'):
        assert_stats_equal(engine.fetch(k), oracle(live, k, 'sm', tr.t0, 60, 60, 50, QS), what='live feed ' + k)


def test_state_and_argument_errors(engine, cfg1):
    from kube_gpu_stats_b200 import _ffi
    from kube_gpu_stats_b200.engine import KgsError
    with pytest.raises(KgsError) as e:
        engine.stats_accumulate()
    assert e.value.code == _ffi.KGS_ERR_STATE
    engine.ingest(cfg1.ts[:8], cfg1.gpu_gid[:8], cfg1.pid[:8], cfg1.util[:8])
    with pytest.raises(KgsError) as e:      # per-pod job without a join table
        engine.window_stats(kinds=('pod',), t0=cfg1.t0, step=60, n_windows=1)
    assert e.value.code == _ffi.KGS_ERR_STATE
    with pytest.raises(KgsError) as e:      # width not a multiple of step
        engine.window_stats(kinds=('gpu',), t0=cfg1.t0, step=60, width=90, n_windows=1, n_gpus=8)
    assert e.value.code == _ffi.KGS_ERR_ARG
    with pytest.raises(KgsError) as e:
        engine.window_stats(kinds=('gpu',), t0=cfg1.t0, step=60, n_windows=1, n_gpus=8, qs=(1.5,))
    assert e.value.code == _ffi.KGS_ERR_ARG


def test_tma_staged_accumulate_gives_the_same_tables(built_lib, cfg1, monkeypatch):
    """k_accum's optional TMA path (cp.async.bulk + mbarrier tile ring, KGS_ACCUM_TMA=1) against the oracle."""
    from kube_gpu_stats_b200.engine import Engine
    monkeypatch.setenv('KGS_ACCUM_TMA', '1')
    tr = synth.config(2, scale=0.02)
    with Engine(0) as eng:
        check(eng, tr, ('pod', 'gpu', 'node'), 'sm', tr.t0, 60, 60, 60, QS)
        check(eng, cfg1, ('pod', 'gpu', 'node'), 'mem', cfg1.t0 + 3, 7, 21, 90, (0.5,))


def test_sixteen_samples_per_lane_accumulate_gives_the_same_tables(built_lib, cfg1, monkeypatch):
    """k_accum's optional two-tiles-at-once path (KGS_DENSE16=1: lane summaries paired by shuffles, one pass of the
    chain machinery per 512 samples) against the oracle: dense input with tenancy changes, a step that is not a
    multiple of 16, sliding windows, and input where it never applies."""
    from kube_gpu_stats_b200.engine import Engine
    monkeypatch.setenv('KGS_DENSE16', '1')
    tr = synth.config(2, scale=0.03)
    with Engine(0) as eng:
        check(eng, tr, ('pod', 'gpu', 'node'), 'sm', tr.t0, 60, 60, 60, QS)
        check(eng, tr, ('pod', 'gpu', 'node'), 'mem', tr.t0 + 5, 17, 17, 200, (0.5,), ingest=False)
        check(eng, tr, ('gpu', 'node'), 'sm', tr.t0, 20, 100, 170, (0.95,), ingest=False)
        check(eng, cfg1, ('pod', 'gpu', 'node'), 'sm', cfg1.t0, 60, 60, 60, (0.5,))


def test_cfg2_full_size_properties_and_exact_slice(engine):
    """BASELINE configs[1] at FULL size (1 000 nodes, 28.8 M samples, 10 000 pods): size-independent
    properties on the whole result and an exact oracle comparison on the first 20 nodes."""
    tr = synth.config(2)
    kinds = ('pod', 'gpu', 'node')
    got = run(engine, tr, kinds, 'sm', tr.t0, 60, 60, 60, (0.95,))
    n_win = 60
    for k in kinds:
        assert int(got[k]['count'].sum()) == tr.n, k                    # every sample lands in exactly one cell per kind
        nz = got[k]['count'] > 0
        p95 = got[k]['pct'][0]
        assert np.all(np.isnan(p95[~nz])) and np.all(p95[nz] <= got[k]['max'][nz]) and np.all(p95[nz] >= 0)
        assert np.all(got[k]['mean'][nz] <= got[k]['max'][nz])
    g = {f: np.asarray(got['gpu'][f]).reshape(tr.n_nodes, 8, n_win) for f in ('count', 'sum', 'max')}
    np.testing.assert_array_equal(g['count'].sum(axis=1).reshape(-1), got['node']['count'])   # node = sum of its GPUs
    np.testing.assert_array_equal(g['sum'].sum(axis=1).reshape(-1), got['node']['sum'])
    np.testing.assert_array_equal(g['max'].max(axis=1).reshape(-1), got['node']['max'])
    assert int(got['pod']['sum'].sum()) == int((tr.util & 0xFF).astype(np.uint64).sum())      # a checksum of checksums
    # exact: the first 20 nodes own the first samples, GPUs and pods of the trace
    nodes = 20
    m = nodes * 8 * 3600
    pods = int(np.searchsorted(tr.pod_node, nodes))
    part = synth.Trace(tr.ts[:m], tr.gpu_gid[:m], tr.pid[:m], tr.util[:m], tr.keys, tr.pod_ids, pods, nodes, 8, tr.t0, 3600)
    for k, nk in (('pod', pods), ('gpu', nodes * 8), ('node', nodes)):
        want = oracle(part, k, 'sm', tr.t0, 60, 60, n_win, (0.95,))
        head = {f: (np.asarray(v)[:, :nk * n_win] if f == 'pct' else np.asarray(v)[:nk * n_win]) for f, v in got[k].items()}
        assert_stats_equal(head, want, what='full-size slice ' + k)
