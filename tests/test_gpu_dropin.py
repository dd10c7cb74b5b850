"""GPU: the reference-shaped functions of the drop-in module (means and join on the device)
against the hand-written golden vectors and against the oracle restatement on a synthetic cluster."""
import io
import json
import os
from contextlib import redirect_stdout
from datetime import datetime

import numpy as np
import pytest

from conftest import bits
from kube_gpu_stats_b200 import gpu_util_stats as gus
from kube_gpu_stats_b200 import replay, synth
from oracle import ref_gpu_util_stats as ref

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
T0, T1 = datetime(2023, 11, 14), datetime(2023, 11, 15)


def quiet(fn, *a):
    buf = io.StringIO()
    with redirect_stdout(buf):
        out = fn(*a)
    return out, buf.getvalue()


def same_rows(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g[:3] == w[:3] and type(g[3]) is type(w[3])
        assert bits([g[3]])[0] == bits([w[3]])[0], (g, w)


def test_golden_rules_on_gpu(built_lib):
    g = json.load(open(os.path.join(GOLDEN, 'ref_rules_gus.json')))
    gus.set_transport(replay.transport(g['answers']))
    try:
        rows, out = quiet(gus.stats_pod_results, g['answers']['sm_util'], T0, T1)
        assert rows == g['expect']['stats_pod_results']['rows']
        assert out == g['expect']['get_pod_by_servers']['dump'] + '\n'
        srv, _ = quiet(gus.stats_server_results, g['answers']['sm_util'], 86400, 3600)
        assert srv == g['expect']['stats_server_results']['rows']
        assert [type(r[2]) for r in srv] == [float, int, int]
    finally:
        gus.set_transport(None)


@pytest.mark.parametrize('step', [3600, 60])
def test_synthetic_cluster_equals_oracle(built_lib, step):
    """configs[0]-style replay: the fixture cluster through Prometheus-shaped JSON.  step 3600 is the
    reference's own query (<= 25 points per series: bit-exact naive sums); step 60 gives 60 points."""
    tr = synth.config(2, scale=0.02)
    req = np.random.default_rng(1).integers(1, 5, size=tr.n_pods)
    live = [p for p in range(tr.n_pods) if p % 7 != 3]                     # some pods have ended
    ans = replay.prometheus_answers(tr, step, gpus_requested=req, namespace_live=live)
    ans['sm_util']['data']['result'].append(dict(ans['sm_util']['data']['result'][5]))   # duplicate key: last wins
    ans['sm_util']['data']['result'][10]['values'] = []                                  # empty series
    gus.set_transport(replay.transport(ans))
    try:
        got, out = quiet(gus.stats_pod_results, ans['sm_util'], T0, T1)
        want, dump = ref.stats_pod_results(ans['sm_util'], T0, T1, replay.fetch_for_oracle(ans))
        same_rows(got, want)
        assert out == dump + '\n'
        got_s, _ = quiet(gus.stats_server_results, ans['sm_util'], 86400, step)
        want_s = ref.stats_server_results(ans['sm_util'], 86400, step, replay.fetch_for_oracle(ans))
        assert [r[:2] + r[3:] for r in got_s] == [r[:2] + r[3:] for r in want_s]
        np.testing.assert_array_equal(bits([r[2] for r in got_s]), bits([r[2] for r in want_s]))
    finally:
        gus.set_transport(None)


def test_main_prints_reference_format(built_lib, capsys):
    g = json.load(open(os.path.join(GOLDEN, 'ref_rules_gus.json')))
    gus.set_transport(replay.transport(g['answers']))
    try:
        gus.main()
    finally:
        gus.set_transport(None)
    out = capsys.readouterr().out
    assert out == g['expect']['get_pod_by_servers']['dump'] + '\n' + '\n'.join(g['expect']['stats_pod_results']['printed']) + '\n'


def test_pmon_text_to_pinned_columns_to_tables(engine):
    """Collector text -> pinned SoA columns -> join + per-GPU table, against the oracle."""
    from kube_gpu_stats_b200 import ingest
    from oracle import window_stats as ws
    rng = np.random.default_rng(4)
    cols = ingest.SampleColumns(engine, 4096)
    for t in range(100):
        rows = ['# gpu pid type sm mem enc dec fb command']
        for g in range(4):
            rows.append('%5d %8d C %5d %5d - - %6d python' % (g, 500 + g, rng.integers(0, 101), rng.integers(0, 101), rng.integers(0, 40000)))
        cols.append_pmon('\n'.join(rows), ts=1000 + t, node_id=3, gpus_per_node=8)
    ts, gid, pid, util = cols.views()
    assert cols.n == 400 and gid.min() == 24 and gid.max() == 27
    keys = np.array([(3 << 32) | 500, (3 << 32) | 502], dtype=np.uint64)
    pods = np.array([0, 1], dtype=np.uint32)
    engine.build_join(keys, pods, 2)
    engine.ingest(ts, gid, pid, util)
    got = engine.window_stats(kinds=('pod', 'gpu'), metric='sm', t0=1000, step=20, n_windows=5, n_gpus=32, n_nodes=4, qs=(0.5,))
    for kind, kid, nk in (('pod', ws.KEY_POD, 2), ('gpu', ws.KEY_GPU, 32)):
        want = ws.window_stats_np(ts, gid, pid, util, keys, pods, key_kind=kid, metric=0, gpus_per_node=8, t0=1000,
                                  step=20, width=20, n_windows=5, n_keys=nk, qs=(0.5,))
        for f in ('count', 'sum', 'max'):
            np.testing.assert_array_equal(np.asarray(got[kind][f]).astype(np.uint64), want[f].astype(np.uint64))
        np.testing.assert_array_equal(bits(got[kind]['mean']), bits(want['mean']))
        np.testing.assert_array_equal(bits(got[kind]['pct']), bits(want['pct']))
