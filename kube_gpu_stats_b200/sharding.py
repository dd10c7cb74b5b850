"""Node sharding across the GPUs of one box and the exchange of per-shard result tables.

A Kubernetes pod is bound to one node (who_use_gpu.py:33 reads ``spec.nodeName``) and the
reference keys its statistic by (server, pod) (gpu_util_stats.py:70-71), so with nodes
partitioned into contiguous ranges every per-pod, per-GPU and per-node cell lives on exactly
one shard: there is no cross-shard arithmetic.  The only exchange step is the all_gather of
the finished tables (north_star: "per-shard partial aggregates are merged with NCCL
allgather"), after which ``merge_tables`` is a concatenation along the key axis.

``torch.distributed`` is the plumbing (rendezvous + NCCL); the tables themselves are the
library's device buffers, aliased as tensors without a copy.
"""
import numpy as np


def shard_ranges(n_items, world):
    """Contiguous, near-equal ranges [start, end) of n_items for each of ``world`` ranks."""
    base, extra = divmod(int(n_items), int(world))
    out, start = [], 0
    for r in range(world):
        end = start + base + (1 if r < extra else 0)
        out.append((start, end))
        start = end
    return out


def shard_of_node(node, n_nodes, world):
    """Rank that owns ``node`` under shard_ranges(n_nodes, world) (vectorised)."""
    node = np.asarray(node)
    base, extra = divmod(int(n_nodes), int(world))
    split = extra * (base + 1)
    lo = node // (base + 1) if base + 1 else np.zeros_like(node)
    hi = extra + (node - split) // max(base, 1)
    return np.where(node < split, lo, hi)


def localise(trace_cols, keys, pod_ids, pod_node, node_range, gpus_per_node):
    """Cut one shard out of a global trace: samples, join pairs and pods whose node lies in
    ``node_range``, with node / gpu / pod ids renumbered from 0.  Returns
    (ts, gid, pid, util, keys, pod_ids, n_pods_local, pod_global_ids)."""
    ts, gid, pid, util = trace_cols
    a, b = node_range
    node = gid // gpus_per_node
    m = (node >= a) & (node < b)
    pod_sel = np.nonzero((pod_node >= a) & (pod_node < b))[0]
    remap = np.full(len(pod_node), 0xFFFFFFFF, dtype=np.uint32)
    remap[pod_sel] = np.arange(len(pod_sel), dtype=np.uint32)
    knode = (keys >> np.uint64(32)).astype(np.int64)
    km = (knode >= a) & (knode < b)
    lkeys = ((knode[km] - a).astype(np.uint64) << np.uint64(32)) | (keys[km] & np.uint64(0xFFFFFFFF))
    return (ts[m], (gid[m] - a * gpus_per_node).astype(np.uint32), pid[m], util[m], lkeys,
            remap[pod_ids[km]], len(pod_sel), pod_sel)


def merge_tables(per_rank, n_windows):
    """Concatenate the shards' tables along the key axis (rank order = node order).
    per_rank: list of {count,sum,max,mean,pct} dicts with [n_keys_r * n_windows] arrays."""
    out = {}
    for name in ('count', 'sum', 'max', 'mean'):
        out[name] = np.concatenate([t[name] for t in per_rank])
    out['pct'] = np.concatenate([np.asarray(t['pct']).reshape(len(t['pct']), -1) for t in per_rank], axis=1)
    return out


class _DeviceBytes:
    """Zero-copy view of library-owned device memory for torch.as_tensor."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {'shape': (int(nbytes),), 'typestr': '|u1', 'data': (int(ptr), False),
                                         'version': 2}


class TableGather:
    """all_gather of the job's whole result region (every table of every key kind) in ONE NCCL
    collective, issued on the stream that is current when all_gather() is called."""

    def __init__(self, engine, kinds, world):
        self.eng, self.kinds, self.world = engine, kinds, world
        self._src_ptr = None
        self._src = None
        self._dst = None

    def all_gather(self):
        import torch
        import torch.distributed as dist
        ptr, nbytes = self.eng.result_region()
        if ptr != self._src_ptr or self._src is None or self._src.numel() != nbytes:
            self._src = torch.as_tensor(_DeviceBytes(ptr, nbytes), device=torch.device('cuda', self.eng.device))
            self._dst = torch.empty(nbytes * self.world, dtype=torch.uint8, device=self._src.device)
            self._src_ptr = ptr
        dist.all_gather_into_tensor(self._dst, self._src)
        return self._dst
