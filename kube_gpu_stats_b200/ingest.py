"""Ingestion: collector text -> the 16 B/sample SoA columns in pinned memory (north_star
subsystem 1).

The reference has no sample source of its own: it reads pre-aggregated
series from Prometheus.  The per-process path named by north_star needs two collector outputs,
whose formats are therefore fixed here:

* ``nvidia-smi pmon -s um -c 1`` - one row per (gpu, process): sm %, mem %, frame buffer MB;
* ``/proc/<pid>/cgroup``         - whose kubepods path carries the pod UID, which
  ``kubectl get pods -o json`` (``metadata.uid``) maps to the pod row of who_use_gpu.

``SampleColumns`` appends parsed rows straight into page-locked columns so ``Engine.ingest``
can hand them to ``cudaMemcpyAsync`` without another copy.
"""
import re

import numpy as np

_POD_UID = re.compile(r'pod([0-9a-fA-F]{8}[-_][0-9a-fA-F]{4}[-_][0-9a-fA-F]{4}[-_][0-9a-fA-F]{4}[-_][0-9a-fA-F]{12})')


def pod_of_cgroup(text):
    """Pod UID from the contents of /proc/<pid>/cgroup, or None for a process outside kubepods.
    Handles the cgroupfs layout (``/kubepods/burstable/pod<uid>/<ctr>``) and the systemd layout
    (``kubepods-burstable-pod<uid with _>.slice``)."""
    for line in text.splitlines():
        if 'kubepods' not in line:
            continue
        m = _POD_UID.search(line)
        if m:
            return m.group(1).replace('_', '-').lower()
    return None


def parse_pmon(text):
    """``nvidia-smi pmon -s um`` rows -> (gpu, pid, sm, mem, fb_mib) int arrays.  Comment lines,
    idle GPUs (pid '-') are skipped; '-' in a utilisation column counts as 0."""
    gpu, pid, sm, mem, fb = [], [], [], [], []
    for line in text.splitlines():
        f = line.split()
        if not f or f[0].startswith('#') or len(f) < 8 or f[1] == '-':
            continue
        num = lambda s: 0 if s == '-' else int(s)   # noqa: E731
        gpu.append(int(f[0]))
        pid.append(int(f[1]))
        sm.append(min(num(f[3]), 255))
        mem.append(min(num(f[4]), 255))
        fb.append(min(num(f[7]), 65535))
    a = lambda x: np.array(x, dtype=np.uint32)      # noqa: E731
    return a(gpu), a(pid), a(sm), a(mem), a(fb)


def pack_util(sm, mem, fb):
    return (np.asarray(sm, np.uint32) & 0xFF) | ((np.asarray(mem, np.uint32) & 0xFF) << 8) | \
        ((np.asarray(fb, np.uint32) & 0xFFFF) << 16)


def series_major_order(ts, gpu_gid, pid):
    """Permutation that puts a batch into series-major collector order: by GPU, then process, then
    time (for one process per GPU this is the plain (gpu, time) order).  The kernels accept any
    order and give the same tables; this order is the fast one, because consecutive samples then
    share (gpu, pid, window) and fold into long runs - with several processes per GPU a plain
    (gpu, time) order interleaves them and every sample becomes its own run."""
    return np.lexsort((np.asarray(ts), np.asarray(pid), np.asarray(gpu_gid)))


class SampleColumns:
    """Growing SoA batch (ts, gpu_gid, pid, util) over pinned memory of a fixed capacity."""

    def __init__(self, engine, capacity):
        self.cols = [engine.pinned(capacity) for _ in range(4)]
        self.n = 0
        self.capacity = capacity

    def append_pmon(self, text, ts, node_id, gpus_per_node=8):
        gpu, pid, sm, mem, fb = parse_pmon(text)
        k = len(gpu)
        if self.n + k > self.capacity:
            raise ValueError('SampleColumns is full')
        s = slice(self.n, self.n + k)
        self.cols[0][s] = ts
        self.cols[1][s] = node_id * gpus_per_node + gpu
        self.cols[2][s] = pid
        self.cols[3][s] = pack_util(sm, mem, fb)
        self.n += k
        return k

    def views(self):
        return [c[:self.n] for c in self.cols]

    def sort_series_major(self):
        """Reorder the batch in place into (gpu, pid, time) order (see series_major_order)."""
        order = series_major_order(*[c[:self.n] for c in self.cols[:3]])
        for c in self.cols:
            c[:self.n] = c[:self.n][order]

    def clear(self):
        self.n = 0
