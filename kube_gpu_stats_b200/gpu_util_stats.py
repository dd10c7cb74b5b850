"""Drop-in for the reference's ``gpu_util_stats/gpu_util_stats.py`` (Python 3, GPU arithmetic).

Same module globals, function names, arguments, return shapes and stdout as the reference
(cited as GUS:line).  What changed underneath:

* the per-sample ``float()`` + ``sum()/len()`` loops (GUS:43-52, GUS:68-80) run on the GPU:
  ``query_prom`` hands the BYTES of the answer to ``kgs_matrix_parse`` (every ``[ts, "v"]`` pair
  found and converted to fp64 on the GPU, bit-identical to ``float()``) and returns a
  ``MatrixAnswer``: a dict with the same shape as ``.json()`` whose sample values live in device
  memory; ``stats_*`` reduce them there (``kgs_matrix_mean``).  A plain dict (a caller's own
  decoded JSON) still works: its values go through ``series_to_csr`` + ``kgs_segmented_mean_f64``;
* the (server, pod) join (GUS:82-92) goes through libkgs' open-addressing hash table
  (``kgs_build_join`` / ``kgs_probe``) with dictionary-encoded server and pod ids - last
  series of a key wins, exactly like the reference's dict assignment (GUS:78);
* HTTP goes through ``_http_get`` so a fixture directory can stand in for Prometheus
  (``KGS_REPLAY_DIR`` or ``set_transport``): the GPU box has no network.

There is no CPU fallback: ``stats_*`` raise if libkgs.so or a CUDA device is missing.
Unspecified in the reference and fixed here: pods of one server are emitted sorted by name
(GUS:86 iterates a Py2 set).
"""
import hashlib
import json
import os
from datetime import datetime, timedelta
from time import mktime

import numpy as np

from .py2fmt import py2_json_dump, py2_repr

prom_base_url = 'http://prometheus.ke-xs-sys.qiniu.io/api/v1'     # GUS:9
proxies = {'http': 'http://10.34.33.80'}                            # GUS:10-12

_transport = None
_engine = None


def set_transport(fn):
    """fn(url, params, timeout, proxies) -> decoded JSON; None restores HTTP / KGS_REPLAY_DIR."""
    global _transport
    _transport = fn


def replay_name(url, params):
    """File name of a recorded answer: endpoint + a digest of the PromQL string."""
    return '%s_%s.json' % (url.rsplit('/', 1)[-1], hashlib.sha1(params['query'].encode()).hexdigest()[:12])


def _http_get(url, params, timeout, use_proxies):
    if _transport is not None:
        return _transport(url, params, timeout, use_proxies)
    replay = os.environ.get('KGS_REPLAY_DIR')
    if replay:
        with open(os.path.join(replay, replay_name(url, params))) as f:
            return json.load(f)
    import requests
    return requests.get(url, params=params, timeout=timeout, proxies=use_proxies).json()


def _http_get_raw(url, params, timeout, use_proxies):
    """Body of the answer as bytes (or whatever a custom transport returns)."""
    if _transport is not None:
        return _transport(url, params, timeout, use_proxies)
    replay = os.environ.get('KGS_REPLAY_DIR')
    if replay:
        with open(os.path.join(replay, replay_name(url, params)), 'rb') as f:
            return f.read()
    import requests
    return requests.get(url, params=params, timeout=timeout, proxies=use_proxies).content


def engine():
    global _engine
    if _engine is None:
        from .engine import Engine
        _engine = Engine(int(os.environ.get('KGS_DEVICE', '0')))
    return _engine


def query_prom(query, start_time, end_time, step_sec):
    """GUS:14-27 (local-timezone epoch seconds, timeout 5, proxies)."""
    payload = {
        'query': query,
        'start': int(mktime(start_time.timetuple())),
        'end': int(mktime(end_time.timetuple())),
        'step': step_sec,
    }
    body = _http_get_raw(prom_base_url + '/query_range', payload, 5, proxies)
    if isinstance(body, (bytes, bytearray, memoryview)):
        return MatrixAnswer(body)
    return body


def query_prom_instant(query):
    """GUS:29-34 (no proxies, unlike query_prom)."""
    return _http_get(prom_base_url + '/query', {'query': query}, 5, None)


class _Series(dict):
    """One entry of ``data.result``: ``metric`` decoded on the host; ``['values']`` is produced on
    demand from the answer's text (the bulk of it is only ever read by the GPU)."""

    def __init__(self, fields, owner, index):
        super().__init__(fields)
        self._owner, self._index = owner, index

    def __missing__(self, key):
        if key != 'values':
            raise KeyError(key)
        return self._owner.values_of(self._index)

    # the reference's decoded JSON has 'values' as an ordinary key: answer the usual dict questions the same way
    def __contains__(self, key):
        return key == 'values' or super().__contains__(key)

    def get(self, key, default=None):
        return self['values'] if key == 'values' else super().get(key, default)

    def keys(self):
        return list(super().keys()) + ['values']

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return super().__len__() + 1

    def items(self):
        return [(k, self[k]) for k in self.keys()]


def answer_skeleton(raw, spans):
    """The answer's text with the inside of every ``values`` array cut out: what is left (status,
    resultType, the label objects, ``"values":[]``) is small and is decoded on the host.
    ``spans[s]`` = byte index of series s' ``{``, of the ``[`` and ``]`` of its values array
    (2**64-1 when it has none) and of its ``}``, as ``kgs_matrix_fetch`` reports them."""
    none = 0xFFFFFFFFFFFFFFFF
    pieces, at = [], 0
    for so, vo, vc, sc in np.asarray(spans, dtype=np.uint64).reshape(-1, 4).tolist():
        if vo != none and vc != none:
            pieces.append(raw[at:vo + 1])
            at = vc
    pieces.append(raw[at:])
    return b''.join(pieces)


class MatrixAnswer(dict):
    """A /query_range answer decoded by ``kgs_matrix_parse`` (GUS:25 ``.json()`` + GUS:47-48,
    GUS:73-74 ``float(val[1])``).  Looks like the reference's decoded JSON; the per-sample numbers
    stay on the GPU as a CSR in series order."""

    def __init__(self, body):
        self.raw = bytes(body)
        self._parse()
        _, _, _, spans = engine().matrix_fetch(ts=False, values=False)
        self.spans = spans
        super().__init__(json.loads(answer_skeleton(self.raw, spans)))
        result = self.get('data', {}).get('result', []) if isinstance(self.get('data'), dict) else []
        if len(result) != len(spans):
            raise ValueError('answer is not a Prometheus matrix: %d series objects, %d entries in data.result'
                             % (len(spans), len(result)))
        for i, fields in enumerate(result):
            had = 'values' in fields
            fields.pop('values', None)
            result[i] = _Series(fields, self, i) if had else fields

    def _parse(self):
        eng = engine()
        self.n_series, self.n_samples = eng.matrix_parse(self.raw, 'matrix')
        self._mine = (id(eng), eng.matrix_generation)      # the engine counts every decode, whoever asked for it

    def _resident(self):
        eng = engine()
        if self._mine != (id(eng), eng.matrix_generation):   # another answer was decoded since (or the engine is new)
            self._parse()
        return eng

    def values_of(self, index):
        so, vo, vc, sc = (int(x) for x in self.spans[index])
        return json.loads(self.raw[vo:vc + 1])

    def series_means(self, mul=0.0, div=0.0):
        return self._resident().matrix_mean(mul, div)

    def series_lengths(self):
        _, _, offsets, _ = self._resident().matrix_fetch(ts=False, values=False, spans=False)
        return np.diff(offsets.astype(np.int64))


def _series_stat(json_result, results, mul=0.0, div=0.0):
    """(statistic per series, samples per series) - on the decoded answer where it lies, else via
    the host-side decode of a caller-supplied dict."""
    if isinstance(json_result, MatrixAnswer):
        return json_result.series_means(mul, div), json_result.series_lengths()
    values, offsets = series_to_csr(results)
    stat = engine().segmented_mean(values, offsets, mul=mul, div=div) if results else np.zeros(0)
    return stat, np.diff(offsets.astype(np.int64))


def series_to_csr(results):
    """GUS:47-48 / GUS:73-74 decode: every ``float(val[1])`` of every series, once, into one
    fp64 array + offsets (the layout ``kgs_segmented_mean_f64`` reduces)."""
    lens = np.fromiter((len(r['values']) for r in results), dtype=np.uint64, count=len(results))
    offsets = np.zeros(len(results) + 1, dtype=np.uint64)
    np.cumsum(lens, out=offsets[1:])
    flat = [v[1] for r in results for v in r['values']]
    values = np.array(flat, dtype=np.float64) if flat else np.zeros(0, dtype=np.float64)
    return values, offsets


def _encode(names, table):
    """Dictionary-encode strings to dense ids (first-seen order)."""
    out = np.empty(len(names), dtype=np.uint64)
    for i, s in enumerate(names):
        out[i] = table.setdefault(s, len(table))
    return out


def stats_server_results(json_result, time_range_sec, step_sec):
    """GUS:36-60.  Keyed by hostname only: the LAST series of a server replaces earlier ones
    (GUS:49); value = sum * step_sec / time_range_sec (GUS:51); one row per inventory server."""
    stat = _series_stat(json_result, json_result['data']['result'], mul=step_sec, div=time_range_sec) \
        if isinstance(json_result, MatrixAnswer) else None
    server_cards = get_gpu_servers()
    results = json_result['data']['result']
    last = {}
    for i, res in enumerate(results):
        last[res['metric']['kubernetes_io_hostname']] = i
    val, lens = stat if stat is not None else _series_stat(json_result, results, mul=step_sec, div=time_range_sec)
    lines = []
    for server in sorted(server_cards.keys()):
        total, used, card = server_cards[server]
        v = float(val[last[server]]) if server in last else 0
        if server in last and lens[last[server]] == 0:
            v = 0                                  # Py2: sum([]) * step / range is the int 0
        lines.append([server, card, v, used, total])
    return lines


def stats_pod_results(json_result, start_time, end_time):
    """GUS:62-94.  Returns ``[[server, pod, card_num, util], ...]``."""
    # GUS:68-80: mean per series on the GPU.  A MatrixAnswer is reduced before the two inventory
    # queries below decode their own (small) answers over it in device memory.
    means = json_result.series_means() if isinstance(json_result, MatrixAnswer) else None
    server_cards = get_gpu_servers()
    server_pods = get_pod_by_servers(start_time, end_time)
    results = json_result['data']['result']
    eng = engine()
    if means is None:
        means, _ = _series_stat(json_result, results)
    # GUS:78 keeps the last series of a (server, pod): build the join table series-key -> index
    sid, pid = {}, {}
    s_ids = _encode([r['metric']['kubernetes_io_hostname'] for r in results], sid)
    p_ids = _encode([r['metric']['pod_name'] for r in results], pid)
    eng.build_join((s_ids << np.uint64(32)) | p_ids, np.arange(len(results), dtype=np.uint32), len(results))
    # GUS:83-92: candidates are the request records of inventory servers (pods without a request
    # record are dropped, GUS:87-90); probe them against the series table
    cand = [(server, pod) for server in sorted(server_cards.keys()) for pod in sorted(server_pods.get(server, {}))]
    if not cand:
        return []
    c_s = _encode([c[0] for c in cand], sid).astype(np.uint32)
    c_p = _encode([c[1] for c in cand], pid).astype(np.uint32)
    zero = np.zeros(len(cand), dtype=np.uint32)
    eng.ingest(zero, c_s, c_p, zero)
    hit = eng.probe(gpus_per_node=1)
    lines = []
    for (server, pod), idx in zip(cand, hit):
        util = float(means[idx]) if idx != 0xFFFFFFFF else 0.0        # GUS:91 default
        lines.append([server, pod, server_pods[server][pod], util])
    return lines


def get_gpu_servers():
    """GUS:96-127: node -> (total cards, used cards or 0, gpu type or '')."""
    server_card, server_total, server_used = {}, {}, {}
    q_total = ('max(kube_node_status_allocatable_nvidia_gpu_cards * on (instance, node) '
               'group_left(label_nvidia_gpu_type) kube_node_labels) by (node,label_nvidia_gpu_type)')
    for res in _http_get(prom_base_url + '/query', {'query': q_total}, 5, proxies)['data']['result']:
        server_card[res['metric']['node']] = res['metric']['label_nvidia_gpu_type']
        server_total[res['metric']['node']] = res['value'][1]
    q_used = ('(sum(max(kube_pod_container_resource_requests_nvidia_gpu_devices{node!=""} * on (instance,pod)  '
              'group_right(node) kube_pod_status_phase{phase="Running"}) by (node,pod)) by (node) > 0) '
              '* on (node) group_right kube_node_labels')
    for res in _http_get(prom_base_url + '/query', {'query': q_used}, 5, proxies)['data']['result']:
        server_used[res['metric']['node']] = res['value'][1]
    return {s: (server_total[s], server_used.get(s, 0), server_card.get(s, '')) for s in server_total}


def get_pod_by_servers(start_time, end_time):
    """GUS:129-151, including the JSON dump on stdout (GUS:150)."""
    result_dict = {}
    live = set()
    status = query_prom_instant('max(kube_pod_status_phase{namespace="ava",phase=~"Running|Pending"}) by (pod) > 0')
    for m in status['data']['result']:
        live.add(m['metric']['pod'])
    answer = query_prom('max(kube_pod_container_resource_requests_nvidia_gpu_devices) by (node, pod)',
                        start_time, end_time, 3600)
    for m in answer['data']['result']:
        server = m['metric'].get('node', '<unknown>')
        pod = m['metric'].get('pod', '<unknown>')
        val = m['values'][0][1]
        if pod not in live:
            continue
        result_dict.setdefault(server, {})[pod] = val
    print(py2_json_dump(result_dict))
    return result_dict


def main():
    """GUS:154-163."""
    step = 3600
    end_time = datetime.now()
    start_time = end_time - timedelta(days=1)
    results = query_prom('avg(container_gpu_sm_util) by (kubernetes_io_hostname, nvidia_gpu_type, pod_name)',
                         start_time, end_time, step)
    for line in stats_pod_results(results, start_time, end_time):
        print(py2_repr(line))


if __name__ == '__main__':
    main()
