"""GPU: kgs_matrix_parse against the reference's own decode - ``json.loads`` of the answer
(gpu_util_stats.py:25) and ``float(val[1])`` per sample (:47-48, :73-74) - bit for bit, through the C ABI."""
import io
import json
import random
import struct
from contextlib import redirect_stdout
from datetime import datetime

import numpy as np
import pytest

from conftest import bits
from kube_gpu_stats_b200 import gpu_util_stats as gus
from kube_gpu_stats_b200 import replay, synth
from oracle import ref_gpu_util_stats as ref

pytestmark = pytest.mark.gpu
T0, T1 = datetime(2023, 11, 14), datetime(2023, 11, 15)


def reference_decode(text, vector=False):
    """What the reference's loops see: per series the list of float(val[1]) (and of the timestamps)."""
    ans = json.loads(text)
    out = []
    for res in ans['data']['result']:
        pairs = [res['value']] if vector else res.get('values', [])
        out.append(([float(p[0]) for p in pairs], [float(p[1]) for p in pairs], res['metric']))
    return out


def check(engine, text, vector=False):
    if isinstance(text, str):
        text = text.encode()
    want = reference_decode(text, vector)
    n_series, n_samples = engine.matrix_parse(text, 'vector' if vector else 'matrix')
    assert n_series == len(want) and n_samples == sum(len(w[1]) for w in want)
    ts, val, off, spans = engine.matrix_fetch()
    assert off[0] == 0 and off[-1] == n_samples
    for s, (w_ts, w_val, w_metric) in enumerate(want):
        a, b = int(off[s]), int(off[s + 1])
        assert b - a == len(w_val), s
        np.testing.assert_array_equal(bits(val[a:b]), bits(w_val), err_msg='series %d values' % s)
        np.testing.assert_array_equal(bits(ts[a:b]), bits(w_ts), err_msg='series %d timestamps' % s)
        so, vo, vc, sc = (int(x) for x in spans[s])
        assert text[so:so + 1] == b'{' and text[sc:sc + 1] == b'}'
        obj = json.loads(text[so:sc + 1])
        assert obj['metric'] == w_metric
        if not vector and 'values' in obj:
            assert json.loads(text[vo:vc + 1]) == obj['values']
    return n_series, n_samples


def matrix(series, **dump):
    return json.dumps({'status': 'success', 'data': {'resultType': 'matrix', 'result': [
        {'metric': m, 'values': v} for m, v in series]}}, **dump)


def test_small_answers_and_layout_variants(engine):
    series = [({'pod_name': 'a', 'kubernetes_io_hostname': 'n1'}, [[1700000000, '57.5'], [1700003600.123, '0'], [1700007200, '100']]),
              ({'pod_name': 'b'}, []),
              ({'pod_name': 'c'}, [[1.7e9, '1e-7']])]
    for dump in ({}, {'separators': (',', ':')}, {'indent': 2}, {'indent': '\t', 'separators': (' , ', ' : ')}):
        check(engine, matrix(series, **dump))
    check(engine, matrix([]))
    check(engine, '{"status":"success","data":{"resultType":"matrix","result":[]}}')
    check(engine, '{"status":"error","errorType":"bad_data","error":"x","data":{"result":[]}}')
    # key order inside a series does not matter; bare numbers as values are what float() accepts too
    check(engine, '{"data":{"result":[{"values":[[1,"2"],[3, 4.5 ]],"metric":{"a":"b"}},{"metric":{},"values":[[5,"6"]]}]}}')


def test_strings_that_look_like_structure(engine):
    """Quotes, backslashes and brackets inside label values; escapes that sit on chunk borders."""
    nasty = ['a"b', 'back\\slash', 'ends with \\', '\\\\', '[[1,"2"]]', '{"values":[[1,"9"]]}', ']}]}', '\\"', 'é \x07',
             '"', '\\' * 63, '\\' * 64, '\\' * 65 + '"', '"' * 70]
    series = [({'pod_name': s, 'x' + s: s}, [[i, str(i)], [i + 1, '2.5']]) for i, s in enumerate(nasty)]
    check(engine, matrix(series))
    check(engine, matrix(series, ensure_ascii=False))
    rng = random.Random(3)
    for pad in range(0, 130, 3):      # slide everything across the 64-byte chunk grid
        lab = ''.join(rng.choice('\\"[]{}a,') for _ in range(rng.randint(1, 90)))
        check(engine, matrix([({'p': 'x' * pad}, [[1, '1']]), ({'pod_name': lab}, [[2, '3'], [4, '5']]), ({'q': lab[::-1]}, [])],
                             separators=(',', ':')))


def test_value_grammar_is_floats(engine):
    toks = ['NaN', '+Inf', '-Inf', 'inf', 'nan', 'Infinity', '1e5', '1E-5', '.5', '5.', ' 7 ', '+3', '-0', '0.1', '0.30000000000000004',
            '999999999999999', '9999999999999999', '-.5', '00012.50', '0.000000000000001', '123456789.123456', '1234567890.123456', '2.675', '-0.0',
            '4.9406564584124654e-324', '2.4703282292062328e-324', '1.7976931348623157e308', '1.7976931348623159e308', '1e400',
            '9007199254740993', '9007199254740992.5', '9007199254740993.0000000000000000000000001', '123456789012345678901234567890',
            '1000000000000000000000', '0.000001', '1' + '0' * 400, '0.' + '0' * 400 + '1']
    rng = random.Random(9)
    for _ in range(20000):
        x = struct.unpack('<d', struct.pack('<Q', rng.getrandbits(64)))[0]
        if x == x and abs(x) != float('inf'):
            toks.append(repr(x))
    for _ in range(5000):
        toks.append('0.' + ''.join(rng.choice('0123456789') for _ in range(rng.randint(17, 30))) + 'e%d' % rng.randint(-330, 310))
    for _ in range(20000):     # short decimals: the one-division path of k_mx_parse
        toks.append('%d.%s' % (rng.randint(0, 10 ** rng.randint(0, 9)), ''.join(rng.choice('0123456789') for _ in range(rng.randint(0, 6)))))
    series = [({'i': str(k)}, [[1700000000 + j if j % 3 else 1700000000.0 + j / 8, t] for j, t in enumerate(toks[k::7])]) for k in range(7)]
    check(engine, matrix(series, separators=(',', ':')))


@pytest.mark.parametrize('bad,where', [
    ('{"data":{"result":[{"metric":{},"values":[[1,"abc"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1,"1__0"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1,"1_000"]]}]}}', 'malformed sample'),     # Python 2's float() has no underscores
    ('{"data":{"result":[{"metric":{},"values":[[+1,"2"]]}]}}', 'malformed sample'),        # bare numbers follow json.loads, not float()
    ('{"data":{"result":[{"metric":{},"values":[[.5,"2"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[5.,"2"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[007,"2"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[nan,"2"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1, 0x10]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1,"1\\u0030"]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1,null]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1]]}]}}', 'malformed sample'),
    ('{"data":{"result":[{"metric":{},"values":[[1,"2"]', 'unbalanced'),
    ('{"data":{"result":[{"metric":{"a":"b},"values":[[1,"2"]]}]}}', 'inside a string'),
])
def test_errors_are_loud(engine, bad, where):
    with pytest.raises(ValueError, match=where):
        engine.matrix_parse(bad.encode())
    with pytest.raises(Exception, match='no decoded answer'):
        engine.matrix_mean()


def test_vector_answer(engine):
    text = json.dumps({'status': 'success', 'data': {'resultType': 'vector', 'result': [
        {'metric': {'node': 'n%d' % i, 'label_nvidia_gpu_type': 'P40'}, 'value': [1700000000.5, str(i % 9)]} for i in range(300)]}})
    assert check(engine, text, vector=True) == (300, 300)


@pytest.mark.parametrize('piece_kb', [None, 16, 112, 4096])
def test_large_answer_crosses_every_boundary(built_lib, monkeypatch, piece_kb):
    """~2.5 M samples / 50 MB of text: thousands of CTAs, series of ragged length, labels full of escapes.
    The text is decoded in pieces while later pieces are still being copied (default 32 MB); small
    pieces put piece borders everywhere and make the result buffers grow while they fill."""
    from kube_gpu_stats_b200.engine import Engine
    if piece_kb:
        monkeypatch.setenv('KGS_PARSE_PIECE_KB', str(piece_kb))
    with Engine(0) as engine:
        _large_answer(engine, 3000 if piece_kb != 16 else 400)


def _large_answer(engine, n_series_made):
    rng = np.random.default_rng(12)
    parts, want_val, lens = [], [], []
    for s in range(n_series_made):
        n = int(rng.integers(0, 1700))
        v = np.round(rng.random(n) * 100, int(rng.integers(0, 6)))
        strs = [repr(float(x)) if i % 3 else ('%g' % x) for i, x in enumerate(v)]
        want_val.append(np.array([float(t) for t in strs], dtype=np.float64))
        lens.append(n)
        lab = json.dumps({'pod_name': 'pod-%d-%s' % (s, '\\"' * int(rng.integers(0, 40))), 'kubernetes_io_hostname': 'n%d' % (s % 50)})
        vals = ','.join('[%d,"%s"]' % (1700000000 + i, t) for i, t in enumerate(strs))
        parts.append('{"metric":%s,"values":[%s]}' % (lab, vals))
    text = ('{"status":"success","data":{"resultType":"matrix","result":[' + ','.join(parts) + ']}}').encode()
    n_series, n_samples = engine.matrix_parse(text)
    assert n_series == n_series_made and n_samples == sum(lens)
    ts, val, off, spans = engine.matrix_fetch()
    np.testing.assert_array_equal(np.diff(off.astype(np.int64)), lens)
    np.testing.assert_array_equal(bits(val), bits(np.concatenate(want_val)))
    assert ts[0] == 1700000000.0 and np.all(np.diff(ts)[np.diff(ts) != 1.0] <= 0)
    for s in (0, 1, n_series_made // 2, n_series_made - 1):
        so, vo, vc, sc = (int(x) for x in spans[s])
        assert json.loads(text[so:sc + 1])['metric']['pod_name'].startswith('pod-%d-' % s)
    # means of the decoded answer where it lies == means of the same CSR sent from the host
    np.testing.assert_array_equal(bits(engine.matrix_mean()), bits(engine.segmented_mean(val, off)))
    np.testing.assert_array_equal(bits(engine.matrix_mean(60, 86400)), bits(engine.segmented_mean(val, off, 60, 86400)))


def quiet(fn, *a):
    buf = io.StringIO()
    with redirect_stdout(buf):
        out = fn(*a)
    return out, buf.getvalue()


def test_dropin_on_raw_bytes_equals_oracle(built_lib):
    """query_prom hands the answer's bytes to the GPU (MatrixAnswer); rows equal the oracle's."""
    tr = synth.config(2, scale=0.02)
    req = np.random.default_rng(1).integers(1, 5, size=tr.n_pods)
    live = [p for p in range(tr.n_pods) if p % 7 != 3]
    ans = replay.prometheus_answers(tr, 3600, gpus_requested=req, namespace_live=live)
    ans['sm_util']['data']['result'].append(dict(ans['sm_util']['data']['result'][5]))
    ans['sm_util']['data']['result'][10]['values'] = []
    as_dicts = replay.transport(ans)

    def as_bytes(url, params, timeout, proxies):
        body = as_dicts(url, params, timeout, proxies)
        return json.dumps(body).encode() if url.endswith('/query_range') else body
    gus.set_transport(as_bytes)
    try:
        answer = gus.query_prom('avg(container_gpu_sm_util) by (kubernetes_io_hostname, nvidia_gpu_type, pod_name)', T0, T1, 3600)
        assert isinstance(answer, gus.MatrixAnswer) and answer['status'] == 'success'
        assert answer['data']['result'][0]['values'] == ans['sm_util']['data']['result'][0]['values']
        assert answer['data']['result'][10]['values'] == []
        got, out = quiet(gus.stats_pod_results, answer, T0, T1)
        want, dump = ref.stats_pod_results(ans['sm_util'], T0, T1, replay.fetch_for_oracle(ans))
        assert len(got) == len(want)
        for g, w in zip(got, want):
            assert g[:3] == w[:3] and bits([g[3]])[0] == bits([w[3]])[0], (g, w)
        assert out == dump + '\n'
        got_s, _ = quiet(gus.stats_server_results, answer, 86400, 3600)
        want_s = ref.stats_server_results(ans['sm_util'], 86400, 3600, replay.fetch_for_oracle(ans))
        assert [r[:2] + r[3:] for r in got_s] == [r[:2] + r[3:] for r in want_s]
        np.testing.assert_array_equal(bits([r[2] for r in got_s]), bits([r[2] for r in want_s]))
    finally:
        gus.set_transport(None)
