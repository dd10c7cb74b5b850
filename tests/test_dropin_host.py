"""CPU: the host half of the drop-in modules - function surface, transport seam, stdout formats -
against the hand-written golden vectors and the oracle."""
import inspect
import io
import json
import os
from contextlib import redirect_stdout
from datetime import datetime

import pytest

from kube_gpu_stats_b200 import gpu_util_stats as gus
from kube_gpu_stats_b200 import py2fmt, replay, synth
from kube_gpu_stats_b200 import who_use_gpu as wug
from oracle import ref_who_use_gpu as ref_wug

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture()
def golden_transport():
    g = load('ref_rules_gus.json')
    log = []
    gus.set_transport(replay.transport(g['answers'], log))
    yield g, log
    gus.set_transport(None)


def test_reference_function_surface():
    """Names and arity of the reference."""
    sig = lambda f: list(inspect.signature(f).parameters)   # noqa: E731
    assert sig(gus.query_prom) == ['query', 'start_time', 'end_time', 'step_sec']
    assert sig(gus.query_prom_instant) == ['query']
    assert sig(gus.stats_server_results) == ['json_result', 'time_range_sec', 'step_sec']
    assert sig(gus.stats_pod_results) == ['json_result', 'start_time', 'end_time']
    assert sig(gus.get_gpu_servers) == [] and sig(gus.get_pod_by_servers) == ['start_time', 'end_time']
    assert sig(gus.main) == [] and sig(wug.main) == []
    assert sig(wug.checkPod) == ['pod_info'] and sig(wug.getGPUType) == ['pod_info']
    assert gus.prom_base_url == 'http://prometheus.ke-xs-sys.qiniu.io/api/v1'
    assert gus.proxies == {'http': 'http://10.34.33.80'}


def test_inventory_and_requests_match_golden(golden_transport):
    g, log = golden_transport
    inv = gus.get_gpu_servers()
    assert {k: list(v) for k, v in inv.items()} == g['expect']['get_gpu_servers']['value']
    buf = io.StringIO()
    with redirect_stdout(buf):
        table = gus.get_pod_by_servers(datetime(2023, 11, 14), datetime(2023, 11, 15))
    assert table == g['expect']['get_pod_by_servers']['value']
    assert buf.getvalue() == g['expect']['get_pod_by_servers']['dump'] + '\n'        # GUS:150
    # transport details of the reference: timeout 5 everywhere, proxies except on the instant query
    assert all(t == 5 for _, _, t, _ in log)
    by_kind = {replay.classify(p): (u, p, px) for u, p, _, px in log}
    assert by_kind['live_pods'][2] is None and by_kind['total_cards'][2] == gus.proxies
    assert by_kind['pod_requests'][0].endswith('/query_range') and by_kind['pod_requests'][1]['step'] == 3600
    assert by_kind['pod_requests'][1]['end'] - by_kind['pod_requests'][1]['start'] == 86400


def test_py2_formats():
    assert py2fmt.py2_repr(['node-1', 'pod-a', '4', 37.5]) == "[u'node-1', u'pod-a', u'4', 37.5]"
    assert py2fmt.py2_repr(['n', '', 0, 0, '8']) == "[u'n', u'', 0, 0, u'8']"
    assert py2fmt.py2_json_dump({'a': {'b': '1', 'c': '2'}}) == '{\n  "a": {\n    "b": "1", \n    "c": "2"\n  }\n}'
    assert py2fmt.py2_json_dump({}) == '{}'


def test_who_use_gpu_matches_golden(tmp_path, monkeypatch):
    g = load('ref_rules_wug.json')
    path = tmp_path / 'pods.json'
    path.write_text(json.dumps(g['pods']))
    monkeypatch.setenv('KGS_KUBECTL_JSON', str(path))
    buf = io.StringIO()
    with redirect_stdout(buf):
        wug.main()
    assert buf.getvalue() == g['expect']['stdout']
    with redirect_stdout(io.StringIO()):
        rows = [wug.checkPod(p) for p in g['pods']['items']]
    assert [r for r in rows if r] == g['expect']['rows']
    assert wug.getGPUType({}) == '<unspecified>'


def test_who_use_gpu_equals_oracle_on_synthetic_cluster(tmp_path, monkeypatch):
    tr = synth.config(2, scale=0.005)
    pods = replay.kubectl_pods(tr)
    pods['items'][3]['spec'].pop('nodeName')                       # an unscheduled pod
    pods['items'][5]['spec']['containers'][0]['resources']['limits'] = {}
    path = tmp_path / 'pods.json'
    path.write_text(json.dumps(pods))
    monkeypatch.setenv('KGS_KUBECTL_JSON', str(path))
    buf = io.StringIO()
    with redirect_stdout(buf):
        wug.main()
    assert buf.getvalue() == ref_wug.main_text(pods)


def test_join_pairs_from_identity_rows():
    rows = [['ava', 'node-a', 'train-1', 'P40', 3], ['ava', 'node-b', 'infer-2', '<unspecified>', 1]]
    owner = [('node-a', 100, 'ava', 'train-1'), ('node-a', 101, 'ava', 'train-1'), ('node-b', 100, 'ava', 'infer-2'),
             ('node-b', 7, 'default', 'web'), ('node-q', 9, 'ava', 'train-1')]
    keys, pods, n = wug.join_pairs(rows, {'node-a': 0, 'node-b': 1}, owner)
    assert keys.tolist() == [100, 101, (1 << 32) | 100] and pods.tolist() == [0, 0, 1] and n == 2


def test_replay_dir_round_trip(tmp_path, monkeypatch):
    g = load('ref_rules_gus.json')
    replay.write_replay_dir(g['answers'], str(tmp_path), None, None)
    monkeypatch.setenv('KGS_REPLAY_DIR', str(tmp_path))
    gus.set_transport(None)
    assert {k: list(v) for k, v in gus.get_gpu_servers().items()} == g['expect']['get_gpu_servers']['value']


def test_answer_skeleton_keeps_everything_but_the_samples():
    """Host half of MatrixAnswer: the text minus the inside of every values array decodes to the
    answer with empty ``values`` lists (the byte spans come from the GPU; here from a plain scan)."""
    import json
    import numpy as np
    from kube_gpu_stats_b200 import gpu_util_stats as gus
    answer = {'status': 'success', 'data': {'resultType': 'matrix', 'result': [
        {'metric': {'pod_name': 'a]"[', 'kubernetes_io_hostname': 'n1'}, 'values': [[1, '2'], [3, '4.5']]},
        {'metric': {'pod_name': 'b'}, 'values': []},
        {'metric': {'pod_name': 'c'}},
        {'values': [[5, '6']], 'metric': {}}]}}
    for dump in ({}, {'separators': (',', ':')}, {'indent': 1}):
        raw = json.dumps(answer, **dump).encode()
        spans, at = [], raw.index(b'"result"')
        for res in answer['data']['result']:
            body = json.dumps(res, **dump)
            if dump.get('indent'):
                body = body.replace('\n', '\n' + ' ' * 3)      # nested three levels deep
            so = raw.index(body.encode(), at)
            sc = so + len(body.encode()) - 1
            none = 2 ** 64 - 1
            if 'values' in res:
                vo = raw.index(b'"values"', so) + raw[raw.index(b'"values"', so):].index(b'[')
                depth, vc = 0, vo
                while True:
                    depth += {ord('['): 1, ord(']'): -1}.get(raw[vc], 0)
                    if depth == 0:
                        break
                    vc += 1
            else:
                vo = vc = none
            spans.append((so, vo, vc, sc))
            at = sc
        got = json.loads(gus.answer_skeleton(raw, np.array(spans, dtype=np.uint64)))
        want = json.loads(raw)
        for res in want['data']['result']:
            if 'values' in res:
                res['values'] = []
        assert got == want



def test_answer_scan_oracle_and_skeleton_on_random_answers():
    """oracle/prom_answer.py (the structure kgs_matrix_parse reports) against json.loads, and the host
    half of MatrixAnswer on its spans, for labels full of quotes, backslashes and brackets."""
    import json
    import random
    import numpy as np
    from kube_gpu_stats_b200 import gpu_util_stats as gus
    from oracle import prom_answer
    rng = random.Random(17)
    alphabet = 'ab\\"[]{},: é\n'
    for trial in range(200):
        series = []
        for _ in range(rng.randint(0, 6)):
            metric = {''.join(rng.choice(alphabet) for _ in range(rng.randint(0, 9))): ''.join(rng.choice(alphabet) for _ in range(rng.randint(0, 30)))
                      for _ in range(rng.randint(0, 3))}
            values = [[1700000000 + j, repr(rng.random() * 100)] for j in range(rng.randint(0, 5))]
            entry = {'metric': metric, 'values': values}
            if rng.random() < 0.2:
                del entry['values']
            if rng.random() < 0.3:
                entry = dict(reversed(list(entry.items())))
            series.append(entry)
        answer = {'status': 'success', 'data': {'resultType': 'matrix', 'result': series}}
        dump = rng.choice(({}, {'separators': (',', ':')}, {'indent': 2}, {'ensure_ascii': False}))
        raw = json.dumps(answer, **dump).encode()
        pos, offsets, spans = prom_answer.scan_answer(raw)
        assert len(spans) == len(series) and len(offsets) == len(series) + 1
        for s, entry in enumerate(series):
            so, vo, vc, sc = spans[s]
            assert json.loads(raw[so:sc + 1]) == entry
            vals = entry.get('values', [])
            assert offsets[s + 1] - offsets[s] == len(vals)
            if 'values' in entry:
                assert json.loads(raw[vo:vc + 1]) == vals
            else:
                assert vo == prom_answer.NONE and vc == prom_answer.NONE
            for k, p in enumerate(pos[offsets[s]:offsets[s + 1]]):
                assert raw[p:p + 1] == b'[' and json.loads(raw[p:raw.index(b']', p) + 1]) == vals[k]
        got = json.loads(gus.answer_skeleton(raw, np.array(spans, dtype=np.uint64).reshape(-1, 4)))
        for entry in series:
            if 'values' in entry:
                entry['values'] = []
        assert got == answer
    import pytest
    with pytest.raises(ValueError):
        prom_answer.scan_answer(b'{"data":{"result":[{"metric":{"a":"b}')
