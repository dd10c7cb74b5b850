#!/usr/bin/env python
"""Value decode of a 1 Hz replay answer: the reference's
``.json()`` + ``float(val[1])`` + ``sum()/len()`` on one host core against kgs_matrix_parse +
kgs_matrix_mean from a pinned host buffer (H2D copy inside the timed region).

The answer text is Prometheus-shaped: ``series`` series of ``points`` [ts,"v"] pairs each
(configs[0]: 64 x 3600; configs[1]-scale: 8000 x 3600 = 28.8 M samples, ~0.6 GB of text).
Prints one JSON object."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kube_gpu_stats_b200.engine import Engine  # noqa: E402


def answer_text(series, points, distinct=64, seed=0):
    """Bytes of a /query_range answer; `distinct` different series bodies are cycled."""
    rng = np.random.default_rng(seed)
    bodies = []
    for s in range(min(distinct, series)):
        v = np.clip(np.round(50 + np.cumsum(rng.normal(0, 2, points)), 1), 0, 100)
        bodies.append(','.join('[%d,"%s"]' % (1700000000 + i, repr(float(x)).rstrip('0').rstrip('.') or '0')
                               for i, x in enumerate(v)))
    parts = []
    for s in range(series):
        parts.append('{"metric":{"kubernetes_io_hostname":"node-%d","nvidia_gpu_type":"B200","pod_name":"pod-%d"},"values":[%s]}'
                     % (s // 8, s, bodies[s % len(bodies)]))
    return ('{"status":"success","data":{"resultType":"matrix","result":[' + ','.join(parts) + ']}}').encode()


def reference_decode(text):
    """gpu_util_stats.py:25 + :68-80 on the host: decode, float() every value, mean per series."""
    res = json.loads(text)['data']['result']
    out = []
    for r in res:
        vals = []
        for v in r['values']:
            vals.append(float(v[1]))
        out.append(sum(vals) / len(vals) if vals else 0.0)
    return out


def measure(series=8000, points=3600, cpu_series=256, reps=5):
    text = answer_text(series, points)
    n_samples = series * points
    with Engine(0) as eng:
        pin = eng.pinned(len(text), dtype=np.uint8)
        pin[:] = np.frombuffer(text, dtype=np.uint8)
        where = (pin.ctypes.data, len(text))
        for _ in range(2):
            eng.matrix_parse(where)
            means = eng.matrix_mean()
        launches0 = eng.launch_count()
        t0 = time.perf_counter()
        for _ in range(reps):
            ns, nv = eng.matrix_parse(where)
            means = eng.matrix_mean()
        wall = (time.perf_counter() - t0) / reps
        launches = (eng.launch_count() - launches0) // reps
        assert (ns, nv) == (series, n_samples)
        # pageable bytes object instead of the pinned buffer (what requests hands over); the first call pins the
        # library's three staging pieces (once per handle), so it is not the one that is timed
        eng.matrix_parse(text)
        t0 = time.perf_counter()
        for _ in range(2):
            eng.matrix_parse(text)
            eng.matrix_mean()
        wall_pageable = (time.perf_counter() - t0) / 2
    small = answer_text(cpu_series, points)
    t0 = time.perf_counter()
    ref = reference_decode(small)
    cpu = time.perf_counter() - t0
    same = bool(np.array_equal(np.asarray(means[:cpu_series]), np.asarray(ref))) or \
        bool(np.allclose(means[:cpu_series], ref, rtol=1e-12, atol=0))
    return {
        'what': 'value decode of a /query_range answer (SURVEY 8 row a1): bytes -> fp64 CSR -> per-series mean',
        'workload': '%d series x %d points, %.1f MB of answer text' % (series, points, len(text) / 1e6),
        'gpu_ms_per_answer': round(wall * 1e3, 3), 'gpu_samples_per_s': round(n_samples / wall),
        'gpu_text_GBps': round(len(text) / wall / 1e9, 2), 'h2d_bytes_per_answer': len(text),
        'gpu_launches_per_answer': int(launches),
        'timing': 'host wall clock around kgs_matrix_parse + kgs_matrix_mean from a pinned host buffer (copy inside)',
        'gpu_ms_from_pageable_bytes': round(wall_pageable * 1e3, 3),
        'cpu_reference_samples_per_s': round(cpu_series * points / cpu),
        'cpu_sample': 'json.loads + float() + sum()/len() on %d series x %d points (%.2f s, 1 core)' % (cpu_series, points, cpu),
        'means_agree_with_cpu': same}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--series', type=int, default=8000)
    ap.add_argument('--points', type=int, default=3600)
    ap.add_argument('--cpu-series', type=int, default=256, help='series in the bounded CPU sample')
    ap.add_argument('--reps', type=int, default=5)
    a = ap.parse_args()
    print(json.dumps(measure(a.series, a.points, a.cpu_series, a.reps)))


if __name__ == '__main__':
    main()
