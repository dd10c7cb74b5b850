"""Seeded synthetic telemetry traces shaped like BASELINE.json's configs.

The reference has no sample source of its own (it asks Prometheus for
``container_gpu_sm_util``, gpu_util_stats.py:159), so the per-process record is
defined here: 16 bytes per sample in four u32 columns

    ts       seconds
    gpu_gid  node * gpus_per_node + gpu
    pid      process id (unique per node only)
    util     sm | mem << 8 | fb_mib << 16

stored in collector order, i.e. sorted by (gpu_gid, ts) with the processes of one
GPU in a fixed slot order inside each second.  The PID->pod table is the list of
``(node << 32 | pid, pod_id)`` pairs a ``who_use_gpu``-style resolver would emit.
"""
from dataclasses import dataclass, field

import numpy as np

T0_DEFAULT = 1_700_000_000


@dataclass
class Trace:
    ts: np.ndarray
    gpu_gid: np.ndarray
    pid: np.ndarray
    util: np.ndarray
    keys: np.ndarray          # u64 (node << 32 | pid), one per tenancy
    pod_ids: np.ndarray       # u32, same length as keys
    n_pods: int
    n_nodes: int
    gpus_per_node: int
    t0: int
    seconds: int
    pod_node: np.ndarray = field(default=None)   # u32 node of each pod (pods are node-local here)
    name: str = ''

    @property
    def n(self):
        return int(self.ts.shape[0])

    @property
    def n_gpus(self):
        return self.n_nodes * self.gpus_per_node


def _ar1(rng, shape, phi, sigma):
    """AR(1) noise along the last axis: x[t] = phi*x[t-1] + e[t]."""
    from scipy.signal import lfilter
    e = rng.standard_normal(shape, dtype=np.float32) * np.float32(sigma)
    return lfilter([1.0], [1.0, -phi], e, axis=-1).astype(np.float32)


def cluster_trace(n_nodes, gpus_per_node=8, procs_per_gpu=1, seconds=3600, n_tenancies=None,
                  pids_per_pod=1, seed=0, t0=T0_DEFAULT, name=''):
    """A cluster where every (gpu, slot) runs one process at a time, 1 Hz sampling.

    ``n_tenancies`` >= number of streams: the surplus streams are cut once at a
    random second into two consecutive tenancies (new pid, new pod) - the churn
    that turns cfg2's 8 000 GPU streams into 10 000 pods.  Tenancies of one node
    are grouped ``pids_per_pod`` at a time into pods, so a pod never spans nodes
    (a Kubernetes pod is bound to one node, who_use_gpu.py:33).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    G = n_nodes * gpus_per_node
    S = G * procs_per_gpu                       # streams
    T = seconds
    n_tenancies = S if n_tenancies is None else n_tenancies
    n_split = n_tenancies - S
    assert 0 <= n_split <= S
    split_at = np.full(S, T, dtype=np.int64)    # second at which the 2nd tenancy starts
    if n_split:
        chosen = rng.choice(S, size=n_split, replace=False)
        split_at[chosen] = rng.integers(1, T, size=n_split)
    # tenancy ids: stream s owns tenancy first[s] and (if split) first[s]+1
    n_ten_stream = 1 + (split_at < T)
    first = np.zeros(S, dtype=np.int64)
    first[1:] = np.cumsum(n_ten_stream)[:-1]
    ten_stream = np.repeat(np.arange(S), n_ten_stream)
    ten_node = ten_stream // (gpus_per_node * procs_per_gpu)
    # pids: unique inside a node (a random permutation slice per node), deliberately
    # reused across nodes so the join key has to carry the node.
    node_start = np.searchsorted(ten_node, np.arange(n_nodes))
    rank_in_node = np.arange(n_tenancies) - node_start[ten_node]
    pool = 4 * int(rank_in_node.max() + 1)
    perm = rng.permutation(pool)                # injective on the rank, rotated per node
    ten_pid = 1000 + (perm[rank_in_node] + 13 * ten_node) % pool
    # pods: consecutive tenancies of a node, pids_per_pod at a time
    pod_local = rank_in_node // pids_per_pod
    pods_in_node = np.zeros(n_nodes, dtype=np.int64)
    np.maximum.at(pods_in_node, ten_node, pod_local + 1)
    pod_base = np.zeros(n_nodes, dtype=np.int64)
    pod_base[1:] = np.cumsum(pods_in_node)[:-1]
    ten_pod = pod_base[ten_node] + pod_local
    n_pods = int(pods_in_node.sum())
    pod_node = np.repeat(np.arange(n_nodes), pods_in_node).astype(np.uint32)

    # per-sample tenancy, laid out [G, T, P]
    sec = np.arange(T, dtype=np.int64)
    second_half = sec[None, :] >= split_at[:, None]                 # [S, T]
    ten_of = first[:, None] + second_half                           # [S, T]
    ten_of = ten_of.reshape(G, procs_per_gpu, T).transpose(0, 2, 1)  # [G, T, P]
    # utilisation: AR(1) around a per-pod base level, clipped to 0..100
    pod_level = rng.uniform(5.0, 95.0, size=n_pods).astype(np.float32)
    base = pod_level[ten_pod][ten_of]
    wobble = _ar1(rng, (S, T), 0.9, 6.0).reshape(G, procs_per_gpu, T).transpose(0, 2, 1)
    sm = np.clip(np.rint(base + wobble), 0, 100).astype(np.uint32)
    wobble2 = _ar1(rng, (S, T), 0.8, 4.0).reshape(G, procs_per_gpu, T).transpose(0, 2, 1)
    mem = np.clip(np.rint(0.6 * base + wobble2), 0, 100).astype(np.uint32)
    fb_level = rng.integers(256, 60000, size=n_tenancies).astype(np.float32)
    drift = _ar1(rng, (S, T), 0.999, 2.0).reshape(G, procs_per_gpu, T).transpose(0, 2, 1)
    fb = np.clip(np.rint(fb_level[ten_of] + drift), 0, 65535).astype(np.uint32)
    util = (sm | (mem << 8) | (fb << 16)).astype(np.uint32).reshape(-1)
    ts = np.broadcast_to((t0 + sec).astype(np.uint32)[None, :, None], (G, T, procs_per_gpu)).reshape(-1).copy()
    gid = np.broadcast_to(np.arange(G, dtype=np.uint32)[:, None, None], (G, T, procs_per_gpu)).reshape(-1).copy()
    pid = ten_pid[ten_of].astype(np.uint32).reshape(-1)
    keys = (ten_node.astype(np.uint64) << np.uint64(32)) | ten_pid.astype(np.uint64)
    assert len(np.unique(keys)) == n_tenancies
    return Trace(ts=ts, gpu_gid=gid, pid=pid, util=util, keys=keys, pod_ids=ten_pod.astype(np.uint32),
                 n_pods=n_pods, n_nodes=n_nodes, gpus_per_node=gpus_per_node, t0=t0, seconds=T,
                 pod_node=pod_node, name=name)


def flat_trace(n_samples, n_keys=1_000_000, n_pods=100_000, n_nodes=12_500, gpus_per_node=8,
               seconds=3600, zipf_s=1.1, seed=0, t0=T0_DEFAULT, name=''):
    """cfg3: a flat stream in random order over ``n_keys`` distinct (node,pid) pairs;
    the number of pids per pod follows a Zipf law, so a few pods own most samples."""
    rng = np.random.Generator(np.random.PCG64(seed))
    per_node = -(-n_keys // n_nodes)
    k = np.arange(n_keys, dtype=np.int64)
    key_node = k // per_node
    key_pid = 1000 + rng.permutation(per_node * 4)[k % per_node] + (key_node % 3)
    keys = (key_node.astype(np.uint64) << np.uint64(32)) | key_pid.astype(np.uint64)
    assert len(np.unique(keys)) == n_keys
    weight = 1.0 / np.arange(1, n_pods + 1, dtype=np.float64) ** zipf_s
    cdf = np.cumsum(weight) / weight.sum()
    pod_of_key = np.searchsorted(cdf, rng.random(n_keys)).astype(np.uint32)
    pod_of_key = np.minimum(pod_of_key, n_pods - 1)
    pick = rng.integers(0, n_keys, size=n_samples)
    gid = (key_node[pick] * gpus_per_node + rng.integers(0, gpus_per_node, size=n_samples)).astype(np.uint32)
    pid = key_pid[pick].astype(np.uint32)
    ts = (t0 + rng.integers(0, seconds, size=n_samples)).astype(np.uint32)
    sm = rng.integers(0, 101, size=n_samples).astype(np.uint32)
    mem = rng.integers(0, 101, size=n_samples).astype(np.uint32)
    fb = rng.integers(0, 65536, size=n_samples).astype(np.uint32)
    util = sm | (mem << 8) | (fb << 16)
    return Trace(ts=ts, gpu_gid=gid, pid=pid, util=util.astype(np.uint32), keys=keys, pod_ids=pod_of_key,
                 n_pods=n_pods, n_nodes=n_nodes, gpus_per_node=gpus_per_node, t0=t0, seconds=seconds,
                 name=name)


def config(idx, seed=0, scale=1.0):
    """BASELINE.json ``configs[idx-1]`` (1-based like SURVEY.md 8d).  ``scale`` < 1 shrinks
    the node count / sample count for parity tests; the bench uses scale 1."""
    if idx == 1:
        return cluster_trace(1, 8, 8, 3600, n_tenancies=64, pids_per_pod=2, seed=seed, name='cfg1')
    if idx == 2:
        nodes = max(1, int(round(1000 * scale)))
        return cluster_trace(nodes, 8, 1, 3600, n_tenancies=nodes * 10, pids_per_pod=1, seed=seed,
                             name='cfg2' if scale == 1.0 else 'cfg2x%g' % scale)
    if idx == 3:
        n = int(100_000_000 * scale)
        return flat_trace(n, n_keys=max(1000, int(1_000_000 * scale)), n_pods=max(100, int(100_000 * scale)),
                          n_nodes=max(13, int(12_500 * scale)), seed=seed,
                          name='cfg3' if scale == 1.0 else 'cfg3x%g' % scale)
    raise ValueError('configs 4 and 5 are multi-GPU shards of cluster_trace; see sharding.py')
