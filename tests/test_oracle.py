"""CPU: the oracle against the hand-written known-answer vectors, and its two statements of the
declared rules against each other."""
import io
import json
import math
import os
from contextlib import redirect_stdout
from datetime import datetime

import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import synth
from oracle import ref_gpu_util_stats as ref
from oracle import ref_who_use_gpu as wug
from oracle import window_stats as ws

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def gus_fetch(answers):
    from kube_gpu_stats_b200 import replay
    return replay.fetch_for_oracle(answers)


def test_ref_gus_known_answers():
    g = load('ref_rules_gus.json')
    fetch, exp = gus_fetch(g['answers']), g['expect']
    inv = ref.node_inventory(fetch)
    assert {k: list(v) for k, v in inv.items()} == exp['get_gpu_servers']['value']
    table, dump = ref.pod_requests(fetch, datetime(2023, 11, 14), datetime(2023, 11, 15))
    assert table == exp['get_pod_by_servers']['value']
    assert dump == exp['get_pod_by_servers']['dump']
    rows, dump2 = ref.stats_pod_results(g['answers']['sm_util'], datetime(2023, 11, 14), datetime(2023, 11, 15), fetch)
    assert rows == exp['stats_pod_results']['rows'] and dump2 == dump
    assert [ref.py2_repr(r) for r in rows] == exp['stats_pod_results']['printed']
    srv = ref.stats_server_results(g['answers']['sm_util'], 86400, 3600, fetch)
    assert srv == exp['stats_server_results']['rows']
    assert type(srv[1][2]) is int and type(srv[0][2]) is float      # GUS:56 int 0 vs GUS:51 float
    text = ref.main_text(fetch, datetime(2023, 11, 14), datetime(2023, 11, 15))
    assert text == dump + '\n' + '\n'.join(exp['stats_pod_results']['printed']) + '\n'


def test_ref_gus_missing_label_is_keyerror():
    """GUS:70-71 index the metric dict directly."""
    bad = {'data': {'result': [{'metric': {'kubernetes_io_hostname': 'n'}, 'values': []}]}}
    with pytest.raises(KeyError):
        ref.pod_means(bad)


def test_ref_naive_sum_is_left_to_right():
    xs = [0.1] * 10 + [1e16, -1e16, 0.3]
    acc = 0.0
    for x in xs:
        acc += x
    assert ref.naive_sum(xs) == acc
    assert ref.naive_sum([]) == 0 and type(ref.naive_sum([])) is int      # Py2 sum([]) is the int 0


def test_ref_wug_known_answers():
    g = load('ref_rules_wug.json')
    rows, total, per_type, printed = wug.ownership(g['pods'])
    assert rows == g['expect']['rows'] and total == g['expect']['total']
    assert per_type == g['expect']['per_type'] and printed == g['expect']['printed_by_checkPod']
    assert wug.main_text(g['pods']) == g['expect']['stdout']


def test_window_rules_known_answers():
    k = load('window_kat.json')
    s, p = k['samples'], k['pairs']
    util = np.array(s['sm'], dtype=np.uint32) | (np.uint32(7) << np.uint32(8)) | (np.uint32(999) << np.uint32(16))
    keys = (np.array(p['node'], dtype=np.uint64) << np.uint64(32)) | np.array(p['pid'], dtype=np.uint64)
    pods = np.array(p['pod'], dtype=np.uint32)
    table = ws.build_table(keys, pods)
    assert table[(0 << 32) | 7] == 0                                  # last pair wins (R2)
    for kind, kid in (('pod', ws.KEY_POD), ('gpu', ws.KEY_GPU), ('node', ws.KEY_NODE)):
        e = k['expect'][kind]
        want = {'count': e['count'], 'sum': e['sum'], 'max': e['max'], 'mean': e['mean'],
                'pct': [[float('nan') if v is None else v for v in row] for row in e['pct']]}
        kw = dict(key_kind=kid, metric=ws.METRIC_SM, gpus_per_node=k['gpus_per_node'], t0=k['t0'], step=k['step'],
                  width=k['width'], n_windows=k['n_windows'], n_keys=e['n_keys'], qs=tuple(k['qs']))
        got_py = ws.window_stats_py(s['ts'], s['gpu_gid'], s['pid'], util, table, **kw)
        got_np = ws.window_stats_np(s['ts'], s['gpu_gid'], s['pid'], util, keys, pods, **kw)
        assert_stats_equal({n: np.array(v) for n, v in got_py.items()}, {n: np.array(v) for n, v in want.items()}, 'py ' + kind)
        assert_stats_equal(got_np, {n: np.array(v) for n, v in want.items()}, 'np ' + kind)


def test_percentile_rule_examples():
    """R5 spelled out: rank = q(n-1), two separately rounded products, one sum."""
    assert ws.percentile([10, 20], 0.5) == 15.0
    assert ws.percentile([7], 0.99) == 7.0
    assert ws.percentile([1, 2, 3, 4], 1.0) == 4.0 and ws.percentile([1, 2, 3, 4], 0.0) == 1.0
    v = list(range(60))
    rank = 0.95 * 59
    lo = math.floor(rank)
    assert ws.percentile(v, 0.95) == v[lo] * (1.0 - (rank - lo)) + v[lo + 1] * (rank - lo)


@pytest.mark.parametrize('seed', [0, 1, 2])
def test_numpy_statement_equals_loop_statement(seed):
    """Random traces incl. unmatched pids, duplicate pairs, out-of-range timestamps, sliding windows."""
    rng = np.random.default_rng(seed)
    n, gpn = 5000, 4
    gid = rng.integers(0, 12, size=n).astype(np.uint32)
    pid = rng.integers(1, 6, size=n).astype(np.uint32)
    ts = rng.integers(950, 1300, size=n).astype(np.uint32)
    util = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    m = 14
    keys = (rng.integers(0, 3, size=m).astype(np.uint64) << np.uint64(32)) | rng.integers(1, 6, size=m).astype(np.uint64)
    pods = rng.integers(0, 5, size=m).astype(np.uint32)
    table = ws.build_table(keys, pods)
    np.testing.assert_array_equal(ws.probe_np(keys, pods, gid, pid, gpn), np.array(ws.probe_py(table, gid, pid, gpn), dtype=np.uint32))
    for kind, nk in ((ws.KEY_POD, 5), (ws.KEY_GPU, 10), (ws.KEY_NODE, 3)):       # gpu 10,11 out of domain (R6)
        for metric in (0, 1, 2):
            kw = dict(key_kind=kind, metric=metric, gpus_per_node=gpn, t0=1000, step=7, width=21, n_windows=30, n_keys=nk,
                      qs=(0.0, 0.5, 0.95, 0.99, 1.0))
            a = ws.window_stats_py(ts, gid, pid, util, table, **kw)
            b = ws.window_stats_np(ts, gid, pid, util, keys, pods, **kw)
            assert_stats_equal({k: np.array(v) for k, v in a.items()}, b, 'kind %d metric %d' % (kind, metric))


def test_synth_configs_shape():
    c1 = synth.config(1)
    assert c1.n == 230400 and c1.n_pods == 32 and len(c1.keys) == 64 and c1.n_gpus == 8
    assert np.all(np.diff(c1.gpu_gid.astype(np.int64)) >= 0)                   # collector order
    c2 = synth.config(2, scale=0.01)
    assert c2.n == 10 * 8 * 3600 and c2.n_pods == 100 and len(np.unique(c2.keys)) == 100
    assert (c2.util & 0xFF).max() <= 100 and ((c2.util >> 8) & 0xFF).max() <= 100
    # pods are node-local (a pod is bound to one node, who_use_gpu.py:33)
    node_of_key = (c2.keys >> np.uint64(32)).astype(np.int64)
    assert np.array_equal(c2.pod_node[c2.pod_ids], node_of_key)


def test_sampled_windows_rule():
    """R3, width < step: only the first `width` seconds of every step count (PromQL instants)."""
    ts = np.array([100, 101, 104, 105, 109, 110, 111, 115], dtype=np.uint32)
    gid = np.zeros(8, dtype=np.uint32)
    util = np.array([1, 2, 3, 4, 5, 6, 7, 8], dtype=np.uint32)
    kw = dict(key_kind=ws.KEY_GPU, metric=0, gpus_per_node=8, t0=100, step=5, width=2, n_windows=3, n_keys=1, qs=(0.5,))
    got = ws.window_stats_np(ts, gid, gid, util, np.zeros(0, np.uint64), np.zeros(0, np.uint32), **kw)
    # windows [100,102) [105,107) [110,112): values {1,2} {4} {6,7}
    assert got['count'].tolist() == [2, 1, 2] and got['sum'].tolist() == [3, 4, 13] and got['max'].tolist() == [2, 4, 7]
    assert got['pct'][0].tolist() == [1.5, 4.0, 6.5]
    py = ws.window_stats_py(ts, gid, gid, util, {}, **kw)
    assert_stats_equal({k: np.array(v) for k, v in py.items()}, got, 'sampled')
