"""Replay fixtures: a synthetic cluster rendered in the two wire formats the reference consumes
 - Prometheus matrix / vector JSON and ``kubectl get pods -o json`` - so the
reference-shaped functions can run with no network and no kubectl.

``prometheus_answers`` evaluates, on the host with numpy, the PromQL the reference sends:
``avg(container_gpu_sm_util) by (host, type, pod)`` at every step (gpu_util_stats.py:159) is the
mean over the pod's processes of the sample AT that second, rendered with ``repr(float)`` the
way Prometheus renders sample values.
"""
import json
import os

import numpy as np

from . import gpu_util_stats as gus


def names(trace):
    node = ['node-%04d' % i for i in range(trace.n_nodes)]
    pod = ['pod-%05d' % i for i in range(trace.n_pods)]
    return node, pod


def pod_of_sample(trace):
    """numpy restatement of the join for fixture rendering only (host side, not the product path)."""
    order = np.argsort(trace.keys, kind='stable')
    ks, vs = trace.keys[order], trace.pod_ids[order]
    probe = ((trace.gpu_gid.astype(np.uint64) // np.uint64(trace.gpus_per_node)) << np.uint64(32)) | trace.pid.astype(np.uint64)
    pos = np.minimum(np.searchsorted(ks, probe), len(ks) - 1)
    return np.where(ks[pos] == probe, vs[pos], 0xFFFFFFFF).astype(np.uint32)


def prometheus_answers(trace, step, gpu_type='P40', gpus_requested=None, namespace_live=None):
    """{answer name: decoded JSON} for the five queries of gpu_util_stats.py."""
    node_name, pod_name = names(trace)
    pod = pod_of_sample(trace)
    ok = pod != 0xFFFFFFFF
    rel = trace.ts.astype(np.int64) - trace.t0
    at_step = ok & (rel % step == 0)
    k = rel[at_step] // step
    n_steps = int(trace.seconds // step) + (1 if trace.seconds % step else 0)
    cell = pod[at_step].astype(np.int64) * n_steps + k
    sm = (trace.util[at_step] & 0xFF).astype(np.float64)
    cnt = np.bincount(cell, minlength=trace.n_pods * n_steps)
    tot = np.bincount(cell, weights=sm, minlength=trace.n_pods * n_steps)
    series = []
    for p in range(trace.n_pods):
        vals = []
        for s in range(n_steps):
            c = p * n_steps + s
            if cnt[c]:
                vals.append([int(trace.t0 + s * step), repr(float(tot[c] / cnt[c]))])
        series.append({'metric': {'kubernetes_io_hostname': node_name[int(trace.pod_node[p])],
                                  'nvidia_gpu_type': gpu_type, 'pod_name': pod_name[p]}, 'values': vals})
    req = gpus_requested if gpus_requested is not None else np.ones(trace.n_pods, dtype=np.int64)
    used = np.zeros(trace.n_nodes, dtype=np.int64)
    np.add.at(used, trace.pod_node, req)
    live = namespace_live if namespace_live is not None else range(trace.n_pods)
    t = int(trace.t0)
    return {
        'sm_util': {'status': 'success', 'data': {'resultType': 'matrix', 'result': series}},
        'total_cards': {'status': 'success', 'data': {'resultType': 'vector', 'result': [
            {'metric': {'node': n, 'label_nvidia_gpu_type': gpu_type}, 'value': [t, str(trace.gpus_per_node)]}
            for n in node_name]}},
        'used_cards': {'status': 'success', 'data': {'resultType': 'vector', 'result': [
            {'metric': {'node': node_name[i]}, 'value': [t, str(int(used[i]))]} for i in range(trace.n_nodes) if used[i]]}},
        'live_pods': {'status': 'success', 'data': {'resultType': 'vector', 'result': [
            {'metric': {'pod': pod_name[p]}, 'value': [t, '1']} for p in live]}},
        'pod_requests': {'status': 'success', 'data': {'resultType': 'matrix', 'result': [
            {'metric': {'node': node_name[int(trace.pod_node[p])], 'pod': pod_name[p]},
             'values': [[t, str(int(req[p]))]]} for p in range(trace.n_pods)]}},
    }


def classify(params):
    """Which of the five answers a request of the reference is asking for."""
    q = params['query']
    if q.startswith('avg(container_gpu_sm_util)'):
        return 'sm_util'
    if 'kube_node_status_allocatable_nvidia_gpu_cards' in q:
        return 'total_cards'
    if q.startswith('(sum(max(kube_pod_container_resource_requests_nvidia_gpu_devices'):
        return 'used_cards'
    if 'kube_pod_status_phase{namespace="ava"' in q:
        return 'live_pods'
    if q.startswith('max(kube_pod_container_resource_requests_nvidia_gpu_devices)'):
        return 'pod_requests'
    raise KeyError('no recorded answer for query: ' + q)


def transport(answers, log=None):
    """A ``set_transport`` callable serving ``answers``; ``log`` collects (url, params, timeout, proxies)."""
    def fn(url, params, timeout, proxies):
        if log is not None:
            log.append((url, dict(params), timeout, proxies))
        return answers[classify(params)]
    return fn


def fetch_for_oracle(answers):
    """The ``fetch(endpoint, params)`` shape oracle/ref_gpu_util_stats.py takes."""
    return lambda endpoint, params: answers[classify(params)]


def write_replay_dir(answers, path, start_time, end_time):
    """Record the answers under the file names ``gpu_util_stats._http_get`` looks for (KGS_REPLAY_DIR)."""
    os.makedirs(path, exist_ok=True)
    asked = {
        'sm_util': ('/query_range', 'avg(container_gpu_sm_util) by (kubernetes_io_hostname, nvidia_gpu_type, pod_name)'),
        'pod_requests': ('/query_range', 'max(kube_pod_container_resource_requests_nvidia_gpu_devices) by (node, pod)'),
        'live_pods': ('/query', 'max(kube_pod_status_phase{namespace="ava",phase=~"Running|Pending"}) by (pod) > 0'),
        'total_cards': ('/query', 'max(kube_node_status_allocatable_nvidia_gpu_cards * on (instance, node) '
                                  'group_left(label_nvidia_gpu_type) kube_node_labels) by (node,label_nvidia_gpu_type)'),
        'used_cards': ('/query', '(sum(max(kube_pod_container_resource_requests_nvidia_gpu_devices{node!=""} * on (instance,pod)  '
                                 'group_right(node) kube_pod_status_phase{phase="Running"}) by (node,pod)) by (node) > 0) '
                                 '* on (node) group_right kube_node_labels'),
    }
    for name, (endpoint, query) in asked.items():
        with open(os.path.join(path, gus.replay_name(gus.prom_base_url + endpoint, {'query': query})), 'w') as f:
            json.dump(answers[name], f)


def kubectl_pods(trace, gpu_type='P40', gpus_requested=None, namespace='ava'):
    """``kubectl get pods --all-namespaces -o json`` for the trace's pods (who_use_gpu.py:14,29-40,54)."""
    node_name, pod_name = names(trace)
    req = gpus_requested if gpus_requested is not None else np.ones(trace.n_pods, dtype=np.int64)
    items = []
    for p in range(trace.n_pods):
        items.append({
            'metadata': {'name': pod_name[p], 'namespace': namespace, 'uid': 'uid-%05d' % p},
            'spec': {'nodeName': node_name[int(trace.pod_node[p])],
                     'containers': [{'name': 'main', 'resources': {'limits': {'alpha.kubernetes.io/nvidia-gpu': str(int(req[p]))}}}],
                     'affinity': {'nodeAffinity': {'requiredDuringSchedulingIgnoredDuringExecution': {'nodeSelectorTerms': [
                         {'matchExpressions': [{'key': 'nvidia_gpu_type', 'operator': 'In', 'values': [gpu_type]}]}]}}}}})
    return {'apiVersion': 'v1', 'kind': 'List', 'items': items}
