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


def main():
    """WUG:7-25."""
    info = json.loads(_kubectl_pods())
    rows = []
    type_gpu_map = {}
    total_gpu_num = 0
    for pod_info in info['items']:
        row = checkPod(pod_info)
        if len(row) == 0:
            continue
        gpu_type, gpu_num = row[3], row[4]
        total_gpu_num += gpu_num
        type_gpu_map[gpu_type] = type_gpu_map.get(gpu_type, 0) + gpu_num
        rows.append(row)
    print(pretty_table(['Namespace', 'Node', 'Pod', 'GPU Type', 'GPU Cores'], rows))
    print('Total GPU: %d' % total_gpu_num)
    for gpu_type, gpu_num in type_gpu_map.items():
        print('%s\t%d' % (gpu_type, gpu_num))


def checkPod(pod_info):
    """WUG:27-49: [namespace, node, pod, gpu_type, sum of container GPU limits] or []."""
    try:
        metadata = pod_info['metadata']
        pod_name = metadata['name']
        namespace = metadata['namespace']
        containers = pod_info['spec']['containers']
        node_name = pod_info['spec']['nodeName']
        total_gpu_num = 0
        for container in containers:
            if 'resources' not in container:
                continue
            if 'limits' not in container['resources']:
                continue
            total_gpu_num += int(container['resources']['limits'].get('alpha.kubernetes.io/nvidia-gpu', 0))
        if total_gpu_num > 0:
            return [namespace, node_name, pod_name, getGPUType(pod_info), total_gpu_num]
    except Exception as e:
        print(e)
        return []
    return []


def getGPUType(pod_info):
    """WUG:51-58."""
    try:
        return (pod_info['spec']['affinity']['nodeAffinity']['requiredDuringSchedulingIgnoredDuringExecution']
                ['nodeSelectorTerms'][0]['matchExpressions'][0]['values'][0])
    except Exception:
        return '<unspecified>'


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
