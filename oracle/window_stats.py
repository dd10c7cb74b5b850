"""Oracle for the *declared* extensions: (node,pid)->pod join and windowed statistics.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  The reference computes only a
whole-range mean per (server, pod) (``gpu_util_stats.py:77-80``); max, percentiles,
windows, per-GPU / per-node keys and the PID join are named by BASELINE.json's
north_star but have **no reference counterpart** (SURVEY.md section 8a row a9), so
this module is their specification.  **Parity unpinned** against the reference;
pinned by ``tests/golden/window_kat.json``.

Declared rules
--------------
R1  Sample = (ts:u32 seconds, gpu_gid:u32 = node*gpus_per_node + gpu, pid:u32,
    util:u32 = sm | mem<<8 | fb_mib<<16).  Metric 0/1/2 selects sm / mem / fb_mib.
R2  Join key = (gpu_gid // gpus_per_node, pid).  The table is built from
    (key, pod_id) pairs in order; a repeated key keeps the LAST pair (the
    reference's dict assignment, gpu_util_stats.py:78).  No entry -> POD_NONE;
    such samples are left out of per-pod cells (gpu_util_stats.py:87-90) but still
    count for per-GPU and per-node cells.
R3  Pane p = (ts - t0) // step for t0 <= ts < t0 + P*step, P = n_windows + W - 1,
    W = width // step (width is a multiple of step; W == 1 is a tumbling window).
    Window j covers panes j .. j+W-1, i.e. [t0 + j*step, t0 + j*step + width).
    A width SHORTER than step gives sampled windows with gaps (the instant vectors a PromQL
    range query evaluates every step, gpu_util_stats.py:159): W = 1, window j =
    [t0 + j*step, t0 + j*step + width), samples in the gaps are left out.
R4  Cell (key, window): count, integer sum, max (0 when empty),
    mean = float(sum)/float(count), 0.0 when empty (gpu_util_stats.py:77-80).
R5  Percentile q in [0,1] over the cell's sorted values v[0..n-1] (Hyndman-Fan
    type 7 in the PromQL ``quantile`` form): rank = q*(n-1); lo = floor(rank);
    hi = min(n-1, lo+1); w = rank - lo; result = v[lo]*(1-w) + v[hi]*w, every
    operation a separately rounded fp64 operation (no FMA).  NaN when empty.
R6  Keys outside their domain (pod_id >= n_pods, gpu_gid >= n_gpus,
    node >= n_nodes) are ignored for that key kind.
"""
import math

import numpy as np

POD_NONE = 0xFFFFFFFF
KEY_POD, KEY_GPU, KEY_NODE = 0, 1, 2
METRIC_SM, METRIC_MEM, METRIC_FB = 0, 1, 2


def metric_of(util, metric):
    """R1: unpack one metric from the packed util word (works on ints and arrays)."""
    if metric == METRIC_SM:
        return util & 0xFF
    if metric == METRIC_MEM:
        return (util >> 8) & 0xFF
    return (util >> 16) & 0xFFFF


def build_table(keys, pod_ids):
    """R2: dict keyed by the 64-bit (node<<32 | pid); the last duplicate wins."""
    table = {}
    for k, p in zip(keys, pod_ids):
        table[int(k)] = int(p)
    return table


def percentile(sorted_vals, q):
    """R5 on an already sorted, non-empty sequence of numbers."""
    n = len(sorted_vals)
    rank = q * (n - 1)
    lo = int(math.floor(rank))
    hi = min(n - 1, lo + 1)
    w = rank - lo
    return float(sorted_vals[lo]) * (1.0 - w) + float(sorted_vals[hi]) * w


# ----------------------------------------------------------------------------
# Pure-Python statement (small cases, and the timed CPU baseline: it is written in
# the reference's own style - dict probes, list appends, sum()/len()).
# ----------------------------------------------------------------------------

def probe_py(table, gpu_gid, pid, gpus_per_node):
    return [table.get(((int(g) // gpus_per_node) << 32) | int(p), POD_NONE) for g, p in zip(gpu_gid, pid)]


def window_stats_py(ts, gpu_gid, pid, util, table, *, key_kind, metric, gpus_per_node,
                    t0, step, width, n_windows, n_keys, qs=()):
    """R1-R6 with plain loops.  Returns dict of lists indexed [key*n_windows + j]."""
    sparse = width < step
    W = 1 if sparse else width // step
    assert sparse or W * step == width
    n_panes = n_windows + W - 1
    panes = {}                                   # (key, pane) -> list of values
    for i in range(len(ts)):
        t = int(ts[i])
        if t < t0:
            continue
        p = (t - t0) // step
        if p >= n_panes:
            continue
        if sparse and (t - t0) - p * step >= width:
            continue
        g = int(gpu_gid[i])
        if key_kind == KEY_GPU:
            key = g
        elif key_kind == KEY_NODE:
            key = g // gpus_per_node
        else:
            key = table.get(((g // gpus_per_node) << 32) | int(pid[i]), POD_NONE)
        if key >= n_keys:                        # R6 (POD_NONE is >= any n_keys)
            continue
        panes.setdefault((key, p), []).append(int(metric_of(int(util[i]), metric)))
    ncell = n_keys * n_windows
    out = {'count': [0] * ncell, 'sum': [0] * ncell, 'max': [0] * ncell, 'mean': [0.0] * ncell,
           'pct': [[float('nan')] * ncell for _ in qs]}
    touched = {}
    for (key, p) in panes:
        for j in range(max(0, p - W + 1), min(n_windows - 1, p) + 1):
            touched.setdefault((key, j), []).append(p)
    for (key, j), plist in touched.items():
        vals = []
        for p in plist:
            vals.extend(panes[(key, p)])
        c = key * n_windows + j
        out['count'][c] = len(vals)
        out['sum'][c] = sum(vals)
        out['max'][c] = max(vals)
        out['mean'][c] = float(sum(vals)) / float(len(vals))
        if qs:
            vals.sort()
            for iq, q in enumerate(qs):
                out['pct'][iq][c] = percentile(vals, q)
    return out


# ----------------------------------------------------------------------------
# numpy statement of the same rules (so parity tests at 10^6..10^8 samples finish in
# seconds).  tests/test_oracle.py checks it against window_stats_py.
# ----------------------------------------------------------------------------

def probe_np(keys, pod_ids, gpu_gid, pid, gpus_per_node):
    """R2 vectorised: sort the (deduplicated, last wins) keys and binary-search them."""
    keys = np.asarray(keys, dtype=np.uint64)
    pod_ids = np.asarray(pod_ids, dtype=np.uint32)
    out = np.full(len(gpu_gid), POD_NONE, dtype=np.uint32)
    if len(keys) == 0 or len(gpu_gid) == 0:
        return out
    # last occurrence of each key: stable sort, then keep the final element of each group
    order = np.argsort(keys, kind='stable')
    ks = keys[order]
    last = np.ones(len(ks), dtype=bool)
    last[:-1] = ks[1:] != ks[:-1]
    ks, vs = ks[last], pod_ids[order][last]
    probe = ((np.asarray(gpu_gid, dtype=np.uint64) // np.uint64(gpus_per_node)) << np.uint64(32)) \
        | np.asarray(pid, dtype=np.uint64)
    pos = np.searchsorted(ks, probe)
    pos_c = np.minimum(pos, len(ks) - 1)
    hit = ks[pos_c] == probe
    out[hit] = vs[pos_c[hit]]
    return out


def percentile_sorted_groups(vals_sorted, starts, counts, q):
    """R5 for many cells at once: ``vals_sorted`` holds each cell's values ascending and
    contiguous from ``starts``; empty cells give NaN.  Same fp64 operation order as
    ``percentile`` (numpy evaluates each elementwise op with one rounding, no FMA)."""
    res = np.full(len(counts), np.nan, dtype=np.float64)
    nz = counts > 0
    n = counts[nz].astype(np.float64)
    rank = np.float64(q) * (n - 1.0)
    lo = np.floor(rank)
    hi = np.minimum(n - 1.0, lo + 1.0)
    w = rank - lo
    base = starts[nz]
    vlo = vals_sorted[base + lo.astype(np.int64)].astype(np.float64)
    vhi = vals_sorted[base + hi.astype(np.int64)].astype(np.float64)
    res[nz] = vlo * (1.0 - w) + vhi * w
    return res


def window_stats_np(ts, gpu_gid, pid, util, keys, pod_ids, *, key_kind, metric, gpus_per_node,
                    t0, step, width, n_windows, n_keys, qs=()):
    """R1-R6 vectorised.  Returns dict of numpy arrays indexed [key*n_windows + j];
    'pct' has shape (len(qs), n_keys*n_windows)."""
    ts = np.asarray(ts, dtype=np.uint32)
    gpu_gid = np.asarray(gpu_gid, dtype=np.uint32)
    util = np.asarray(util, dtype=np.uint32)
    sparse = width < step
    W = 1 if sparse else width // step
    assert sparse or W * step == width
    n_panes = n_windows + W - 1
    ncell = n_keys * n_windows
    if key_kind == KEY_GPU:
        key = gpu_gid.astype(np.int64)
    elif key_kind == KEY_NODE:
        key = (gpu_gid // np.uint32(gpus_per_node)).astype(np.int64)
    else:
        key = probe_np(keys, pod_ids, gpu_gid, pid, gpus_per_node).astype(np.int64)
    t = ts.astype(np.int64) - int(t0)
    pane = np.floor_divide(t, int(step))
    ok = (t >= 0) & (pane < n_panes) & (key < n_keys)
    if sparse:
        ok &= (t - pane * int(step)) < int(width)
    key, pane = key[ok], pane[ok]
    val = metric_of(util[ok], metric).astype(np.int64)
    count = np.zeros(ncell, dtype=np.int64)
    total = np.zeros(ncell, dtype=np.int64)
    vmax = np.zeros(ncell, dtype=np.int64)
    cell_chunks, val_chunks = [], []
    for d in range(W):                           # a pane feeds windows pane-d, d = 0..W-1
        j = pane - d
        m = (j >= 0) & (j < n_windows)
        cell = key[m] * n_windows + j[m]
        count += np.bincount(cell, minlength=ncell)
        total += np.bincount(cell, weights=val[m], minlength=ncell).astype(np.int64)
        np.maximum.at(vmax, cell, val[m])
        if qs:
            cell_chunks.append(cell)
            val_chunks.append(val[m])
    mean = np.zeros(ncell, dtype=np.float64)
    nz = count > 0
    mean[nz] = total[nz].astype(np.float64) / count[nz].astype(np.float64)
    out = {'count': count.astype(np.uint64), 'sum': total.astype(np.uint64),
           'max': vmax.astype(np.uint32), 'mean': mean,
           'pct': np.full((len(qs), ncell), np.nan, dtype=np.float64)}
    if qs:
        cell = np.concatenate(cell_chunks)
        v = np.concatenate(val_chunks)
        order = np.lexsort((v, cell))
        vs = v[order]
        starts = np.zeros(ncell, dtype=np.int64)
        starts[1:] = np.cumsum(count)[:-1]
        for iq, q in enumerate(qs):
            out['pct'][iq] = percentile_sorted_groups(vs, starts, count, q)
    return out


def segmented_mean_py(values, offsets):
    """The reference's per-series mean (gpu_util_stats.py:77-80) over a CSR layout:
    series s owns values[offsets[s]:offsets[s+1]]; naive left-to-right fp64 sum."""
    out = []
    for s in range(len(offsets) - 1):
        a, b = int(offsets[s]), int(offsets[s + 1])
        acc = 0.0
        for x in values[a:b]:
            acc += float(x)
        out.append(acc / (b - a) if b > a else 0.0)
    return out
