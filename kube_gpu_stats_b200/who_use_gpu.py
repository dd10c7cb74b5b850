"""Drop-in for the reference's ``who_use_gpu/who_use_gpu.py`` (cited as WUG:line).

Same function names, arguments, return shapes and stdout.  This script is the BUILD side of
the join: its rows are the pod identity records, and ``join_pairs`` turns a PID->pod resolution
(build-defined: the reference has none) into the (node << 32 | pid, pod_id)
pairs ``kgs_build_join`` takes.  The JSON walk itself stays on the host (a few thousand pods).
``kubectl`` is replaced by ``KGS_KUBECTL_JSON`` (a file) when set: the GPU box has no kubectl.
"""
import json
import os
import subprocess

import numpy as np

from .py2fmt import pretty_table


def _kubectl_pods():
    path = os.environ.get('KGS_KUBECTL_JSON')
    if path:
        with open(path) as f:
            return f.read()
    proc = subprocess.Popen(['kubectl get pods --all-namespaces -o json'], shell=True, stdout=subprocess.PIPE)
    return proc.stdout.read()


LEGACY_GPU_KEY = 'alpha.kubernetes.io/nvidia-gpu'        # the only resource key the reference counts (WUG:40)
TABLE_HEADER = ['Namespace', 'Node', 'Pod', 'GPU Type', 'GPU Cores']
_AFFINITY_PATH = ('spec', 'affinity', 'nodeAffinity', 'requiredDuringSchedulingIgnoredDuringExecution',
                  'nodeSelectorTerms', 0, 'matchExpressions', 0, 'values', 0)


def _gpu_limits(containers):
    """Sum of the containers' legacy GPU limits; containers without resources/limits count 0."""
    asked = (c['resources']['limits'].get(LEGACY_GPU_KEY, 0)
             for c in containers if 'limits' in c.get('resources', ()))
    return sum(int(v) for v in asked)


def ownership_rows(items):
    """(rows, grand total, totals per GPU type in first-seen order) over the pods that own GPUs."""
    rows, per_type = [], {}
    for item in items:
        rec = checkPod(item)
        if rec:
            rows.append(rec)
            per_type[rec[3]] = per_type.get(rec[3], 0) + rec[4]
    return rows, sum(r[4] for r in rows), per_type


def main():
    """WUG:7-25: table, grand total, one line per GPU type."""
    rows, total, per_type = ownership_rows(json.loads(_kubectl_pods())['items'])
    lines = [pretty_table(TABLE_HEADER, rows), 'Total GPU: %d' % total]
    lines += ['%s\t%d' % kv for kv in per_type.items()]
    print('\n'.join(lines))


def checkPod(pod_info):
    """WUG:27-49: ``[namespace, node, pod, gpu_type, sum of container GPU limits]`` for a pod that
    owns GPUs, else ``[]``.  Any missing field prints the exception and yields ``[]`` (WUG:46-48);
    the node name is looked up before the GPUs are counted, as in the reference (WUG:33)."""
    try:
        meta, spec = pod_info['metadata'], pod_info['spec']
        ident = (meta['name'], meta['namespace'])
        containers, node = spec['containers'], spec['nodeName']
        gpus = _gpu_limits(containers)
    except Exception as err:
        print(err)
        return []
    if gpus <= 0:
        return []
    return [ident[1], node, ident[0], getGPUType(pod_info), gpus]


def getGPUType(pod_info):
    """WUG:51-58: first value of the first required node-affinity expression, else '<unspecified>'."""
    here = pod_info
    try:
        for step in _AFFINITY_PATH:
            here = here[step]
    except Exception:
        return '<unspecified>'
    return here


def join_pairs(rows, node_ids, pid_owner):
    """Pod identity rows -> the build side of the GPU join.

    rows      : checkPod rows ([namespace, node, pod, gpu_type, n]); row index = pod_id
    node_ids  : {node name: node_id}
    pid_owner : iterable of (node name, pid, namespace, pod name) from a PID->cgroup->pod
                resolver (``ingest.pod_of_cgroup``)
    Returns (keys u64, pod_ids u32, n_pods).  PIDs whose pod has no GPU row are left out, which
    makes their samples unmatched, i.e. dropped from per-pod tables (GUS:87-90 semantics)."""
    index = {(r[0], r[2]): i for i, r in enumerate(rows)}
    keys, pods = [], []
    for node, pid, namespace, pod in pid_owner:
        pod_id = index.get((namespace, pod))
        if pod_id is None or node not in node_ids:
            continue
        keys.append((int(node_ids[node]) << 32) | int(pid))
        pods.append(pod_id)
    return np.array(keys, dtype=np.uint64), np.array(pods, dtype=np.uint32), len(rows)


if __name__ == '__main__':
    main()
