"""The two aggregations Prometheus evaluates for the reference before the scripts see any data
(SURVEY.md section 8f item 2), done on the GPU from raw per-process samples:

* ``avg(container_gpu_sm_util) by (kubernetes_io_hostname, nvidia_gpu_type, pod_name)`` evaluated
  every ``step`` seconds (gpu_util_stats.py:159) - the per-pod mean of the samples AT each
  evaluation instant, i.e. windowed statistics with sampled windows of width 1 s;
* ``max(...) by (node, pod)`` (gpu_util_stats.py:137) - the same with max instead of mean.

Both are one ``Engine.window_stats`` job (the fused join + accumulate kernel); this module only
shapes the result as the Prometheus matrix JSON of SURVEY.md appendix B.
"""
import numpy as np


def instant_by_pod(engine, t0, step, n_steps, gpus_per_node=8, metric='sm', width=1):
    """Per-pod count / sum / max / mean of the resident batch at t0, t0+step, ... (n_steps instants)."""
    return engine.window_stats(kinds=('pod',), metric=metric, gpus_per_node=gpus_per_node, t0=t0, step=step,
                               width=width, n_windows=n_steps)['pod']


def matrix_json(table, t0, step, n_steps, pod_labels, stat='mean'):
    """Prometheus range-query answer (resultType matrix): one series per pod, instants with no
    sample are absent, values rendered with repr(float) like Prometheus renders samples.
    pod_labels[p] = the metric dict of pod p."""
    vals = np.asarray(table[stat], dtype=np.float64).reshape(len(pod_labels), n_steps)
    cnt = np.asarray(table['count']).reshape(len(pod_labels), n_steps)
    result = []
    for p, labels in enumerate(pod_labels):
        pts = [[int(t0 + k * step), repr(float(vals[p, k]))] for k in range(n_steps) if cnt[p, k]]
        result.append({'metric': dict(labels), 'values': pts})
    return {'status': 'success', 'data': {'resultType': 'matrix', 'result': result}}
