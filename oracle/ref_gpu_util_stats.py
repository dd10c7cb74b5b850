"""Oracle restatement of the reference's ``gpu_util_stats/gpu_util_stats.py``.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Pure Python, no numpy, float
arithmetic in the reference's own style so that it also serves as the timed CPU
baseline.  The reference is Python 2 and cannot execute here; **parity unpinned**
(no reference tests exist), rules pinned by ``tests/golden/ref_rules_gus.json``.

Every function cites the reference lines it follows (``GUS:a-b`` =
``/root/reference/gpu_util_stats/gpu_util_stats.py`` lines a..b).  The HTTP
transport (GUS:14-34, GUS:107, GUS:119) is replaced by a ``fetch`` callable so
fixtures can be replayed: ``fetch(endpoint, params) -> decoded JSON`` where
``endpoint`` is ``'query'`` or ``'query_range'``.

Differences from the Python-2 source that are unavoidable and deliberate:
* Py2 ``sum()`` is a naive left-to-right fp64 sum; Py3.12 ``sum()`` is
  compensated.  ``naive_sum`` below restates the Py2 behaviour.
* The order of pods inside one server (GUS:86 iterates a Py2 ``set``) is
  unspecified in the reference; the oracle emits them sorted by pod name.
"""
import json
from time import mktime

# PromQL strings sent by the reference (GUS:105, GUS:117, GUS:133, GUS:137, GUS:159).
Q_TOTAL_CARDS = ('max(kube_node_status_allocatable_nvidia_gpu_cards * on (instance, node) '
                 'group_left(label_nvidia_gpu_type) kube_node_labels) by (node,label_nvidia_gpu_type)')
Q_USED_CARDS = ('(sum(max(kube_pod_container_resource_requests_nvidia_gpu_devices{node!=""} * on (instance,pod)  '
                'group_right(node) kube_pod_status_phase{phase="Running"}) by (node,pod)) by (node) > 0) '
                '* on (node) group_right kube_node_labels')
Q_LIVE_PODS = 'max(kube_pod_status_phase{namespace="ava",phase=~"Running|Pending"}) by (pod) > 0'
Q_POD_REQUESTS = 'max(kube_pod_container_resource_requests_nvidia_gpu_devices) by (node, pod)'
Q_SM_UTIL = 'avg(container_gpu_sm_util) by (kubernetes_io_hostname, nvidia_gpu_type, pod_name)'


def naive_sum(xs):
    """Left-to-right fp64 sum: what Py2 ``sum(list_of_floats)`` does (GUS:51, GUS:78).
    Like ``sum`` it starts from the int 0, so an empty list gives the int 0."""
    acc = 0
    for x in xs:
        acc += x
    return acc


def range_params(query, start_time, end_time, step_sec):
    """GUS:16-23: local-timezone epoch seconds, microseconds dropped."""
    return {
        'query': query,
        'start': int(mktime(start_time.timetuple())),
        'end': int(mktime(end_time.timetuple())),
        'step': step_sec,
    }


def decode_series(series):
    """GUS:47-48 / GUS:73-74: ``float(val[1])`` for each [ts, "str"] pair."""
    return [float(pair[1]) for pair in series['values']]


def node_inventory(fetch):
    """GUS:96-127 ``get_gpu_servers``: node -> (total, used or 0, gpu type or '').

    Counts stay the strings Prometheus delivered (GUS:110, GUS:122).  A node that
    only appears in the 'used' answer is not reported (GUS:125 iterates totals).
    """
    kind, total, used = {}, {}, {}
    for item in fetch('query', {'query': Q_TOTAL_CARDS})['data']['result']:
        node = item['metric']['node']
        kind[node] = item['metric']['label_nvidia_gpu_type']   # KeyError if absent (GUS:111)
        total[node] = item['value'][1]
    for item in fetch('query', {'query': Q_USED_CARDS})['data']['result']:
        used[item['metric']['node']] = item['value'][1]
    return {node: (total[node], used.get(node, 0), kind.get(node, '')) for node in total}


def pod_requests(fetch, start_time, end_time):
    """GUS:129-151 ``get_pod_by_servers``: server -> pod -> first sample string.

    Returns ``(table, dump_text)`` where ``dump_text`` is what GUS:150 prints
    (Py2 ``json.dumps(indent=2)`` uses ', ' as item separator, so every line that
    ends in a comma carries a trailing blank).
    """
    live = set()
    for item in fetch('query', {'query': Q_LIVE_PODS})['data']['result']:
        live.add(item['metric']['pod'])
    table = {}
    answer = fetch('query_range', range_params(Q_POD_REQUESTS, start_time, end_time, 3600))
    for item in answer['data']['result']:
        server = item['metric'].get('node', '<unknown>')
        pod = item['metric'].get('pod', '<unknown>')
        first = item['values'][0][1]                   # GUS:142: first sample only
        if pod in live:                                # GUS:144-145
            table.setdefault(server, {})[pod] = first
    return table, json.dumps(table, indent=2, separators=(', ', ': '))


def pod_means(json_result):
    """GUS:65-80: (server, pod) -> mean; empty series -> 0.0; the last series wins."""
    cells = {}
    for series in json_result['data']['result']:
        server = series['metric']['kubernetes_io_hostname']
        pod = series['metric']['pod_name']
        vals = decode_series(series)
        per_server = cells.setdefault(server, {})
        per_server[pod] = naive_sum(vals) / len(vals) if vals else 0.0
    return cells


def join_pod_rows(inventory, requests, means):
    """GUS:82-94: sorted inventory servers x (request pods U util pods).

    Pods with utilisation but no request record are dropped (GUS:87-90); pods
    with a request record but no series report 0.0 (GUS:91).
    """
    rows = []
    for server in sorted(inventory):
        if server not in requests and server not in means:
            continue
        asked = requests.get(server, {})
        seen = means.get(server, {})
        for pod in sorted(set(asked) | set(seen)):
            cards = asked.get(pod)
            if cards is None:
                continue
            rows.append([server, pod, cards, seen.get(pod, 0.0)])
    return rows


def stats_pod_results(json_result, start_time, end_time, fetch):
    """GUS:62-94.  Inventory and request look-ups come first (GUS:63-64), so the
    JSON dump of GUS:150 precedes any row.  Returns ``(rows, dump_text)``."""
    inventory = node_inventory(fetch)
    requests, dump_text = pod_requests(fetch, start_time, end_time)
    return join_pod_rows(inventory, requests, pod_means(json_result)), dump_text


def server_weighted_means(json_result, time_range_sec, step_sec):
    """GUS:43-52: keyed by hostname only, so a later series of the same server
    replaces an earlier one (GUS:49); value = sum * step / range (GUS:51)."""
    last_vals = {}
    for series in json_result['data']['result']:
        last_vals[series['metric']['kubernetes_io_hostname']] = decode_series(series)
    out = {}
    for server, vals in last_vals.items():
        total = naive_sum(vals)
        if isinstance(total, int) and isinstance(step_sec, int) and isinstance(time_range_sec, int):
            out[server] = total * step_sec // time_range_sec      # Py2 int / int (only the empty series: 0)
        else:
            out[server] = total * step_sec / time_range_sec
    return out


def stats_server_results(json_result, time_range_sec, step_sec, fetch):
    """GUS:36-60 (dead code in the reference: its only call, GUS:160, is commented
    out).  One row per inventory server; int 0 when the server has no series."""
    inventory = node_inventory(fetch)
    value = server_weighted_means(json_result, time_range_sec, step_sec)
    rows = []
    for server in sorted(inventory):
        total, used, card = inventory[server]
        rows.append([server, card, value[server] if server in value else 0, used, total])
    return rows


def py2_repr(obj):
    """Python-2 ``repr`` of the row lists printed at GUS:163: text that came out of
    ``json`` is unicode there, hence the ``u'..'`` prefix."""
    if isinstance(obj, str):
        return 'u' + repr(obj)
    if isinstance(obj, (list, tuple)):
        body = ', '.join(py2_repr(x) for x in obj)
        if isinstance(obj, list):
            return '[' + body + ']'
        return '(' + body + (',)' if len(obj) == 1 else ')')
    return repr(obj)


def main_text(fetch, start_time, end_time, step=3600):
    """GUS:154-163 as text: the JSON dump, then one Py2 list repr per row."""
    series = fetch('query_range', range_params(Q_SM_UTIL, start_time, end_time, step))
    rows, dump_text = stats_pod_results(series, start_time, end_time, fetch)
    return '\n'.join([dump_text] + [py2_repr(r) for r in rows]) + '\n'
