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
    """all_gather of the job's exchange region (sum, pct, count, max of every key kind) in ONE NCCL
    collective.

    ``all_gather()`` issues it on the stream that is current (serial with the kernels).
    ``all_gather_async(stream)`` issues it on a side stream once the library's compute stream has
    finished the job, so it overlaps the NEXT job's kernels: consecutive jobs write alternate result
    regions (include/kgs.h), and a region is only reused two jobs later, by which time the compute
    stream has been made to wait for the collective that read it.  ``wait()`` joins the side stream.
    """

    def __init__(self, engine, kinds, world):
        self.eng, self.kinds, self.world = engine, kinds, world
        self._views = {}          # region address -> (src alias, dst buffer, done event)
        self._side = None
        self._last = None         # region of the most recent job

    def _buffers(self):
        import torch
        ptr, nbytes, _ = self.eng.exchange_region()    # packed tables when on, else 16 + 8*n_q bytes per cell
        v = self._views.get(ptr)
        if v is None or v[0].numel() != nbytes:
            dev = torch.device('cuda', self.eng.device)
            src = torch.as_tensor(_DeviceBytes(ptr, nbytes), device=dev)
            v = (src, torch.empty(nbytes * self.world, dtype=torch.uint8, device=dev), torch.cuda.Event())
            self._views[ptr] = v
        self._last = ptr
        return v

    def all_gather(self):
        import torch.distributed as dist
        src, dst, _ = self._buffers()
        dist.all_gather_into_tensor(dst, src)
        return dst

    def all_gather_async(self, compute_stream):
        import torch
        import torch.distributed as dist
        if self._side is None:
            self._side = torch.cuda.Stream(device=torch.device('cuda', self.eng.device))
        src, dst, done = self._buffers()
        self._side.wait_stream(compute_stream)          # the job's finalize kernels
        with torch.cuda.stream(self._side):
            dist.all_gather_into_tensor(dst, src)
            done.record(self._side)
        return dst

    def before_next_job(self, compute_stream):
        """Call before starting a job: its result region was last read by the collective issued two
        jobs ago; make the compute stream wait for that one (normally long finished)."""
        for ptr, (_, _, done) in self._views.items():
            if ptr != self._last:                       # the next job writes the OTHER region
                compute_stream.wait_event(done)

    def join(self, compute_stream):
        """Order everything issued on the side stream before whatever comes next on compute_stream."""
        if self._side is not None:
            compute_stream.wait_stream(self._side)

    def wait(self):
        import torch
        if self._side is not None:
            torch.cuda.current_stream().wait_stream(self._side)
            self._side.synchronize()


class PeerTableGather(TableGather):
    """The same exchange without a collective kernel: every rank PUSHES its packed tables straight
    into every other rank's gathered buffer over NVLink with the copy engines (peer-mapped symmetric
    memory from ``torch.distributed._symmetric_memory``; one ``cudaMemcpyAsync`` per peer), followed
    by one signal-pad barrier.  No SM is taken from the statistics kernels of the next job, which is
    what an NCCL all_gather running beside them costs (bench.py, N = 2..8).

    Gathered buffers are double-buffered like the result regions: the push of job k+2 into a peer's
    slot happens after the barrier of job k+1, which that peer enters only after everything it had
    queued before (including readers of job k's slot)."""

    def __init__(self, engine, kinds, world, group=None):
        super().__init__(engine, kinds, world)
        import torch.distributed as dist
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self._sym = None          # (nbytes, local symmetric tensor, handle, per-parity per-rank views)
        self._parity = 0

    def _symmetric(self, nbytes):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        if self._sym is not None and self._sym[0] == nbytes:
            return self._sym
        dev = torch.device('cuda', self.eng.device)
        biggest = torch.tensor([nbytes], dtype=torch.int64, device=dev)
        dist.all_reduce(biggest, op=dist.ReduceOp.MAX, group=self.group)
        if int(biggest) != nbytes:
            raise ValueError('PeerTableGather needs equally sized exchange regions on every rank '
                             '(%d bytes here, %d elsewhere)' % (nbytes, int(biggest)))
        slot = (nbytes + 255) // 256 * 256
        local = symm.empty(2 * self.world * slot, dtype=torch.uint8, device=dev)
        hdl = symm.rendezvous(local, self.group)
        views = [[hdl.get_buffer(r, (nbytes,), torch.uint8, (par * self.world + self.rank) * slot)
                  for r in range(self.world)] for par in range(2)]
        mine = [local[par * self.world * slot:(par + 1) * self.world * slot].view(self.world, slot)[:, :nbytes] for par in range(2)]
        self._sym = (nbytes, local, hdl, views, mine)
        return self._sym

    def _push(self):
        import torch
        ptr, nbytes, _ = self.eng.exchange_region()
        dev = torch.device('cuda', self.eng.device)
        src = self._views.get(ptr)
        if src is None or src[0].numel() != nbytes:
            src = (torch.as_tensor(_DeviceBytes(ptr, nbytes), device=dev), None, torch.cuda.Event())
            self._views[ptr] = src
        self._last = ptr
        _, _, hdl, views, mine = self._symmetric(nbytes)
        par = self._parity
        self._parity ^= 1
        # one stream: spreading the copies over several streams was measured slower at 8 GPUs (0.35 vs
        # 0.264 ms per step) - the extra stream/event calls cost more host time than the copies overlap
        for k in range(self.world):                      # start with the neighbour: spread the links
            r = (self.rank + 1 + k) % self.world
            views[par][r].copy_(src[0], non_blocking=True)
        hdl.barrier(channel=par, timeout_ms=20000)
        return mine[par], src[2]

    def all_gather(self):
        dst, _ = self._push()
        return dst

    def all_gather_async(self, compute_stream):
        import torch
        if self._side is None:
            self._side = torch.cuda.Stream(device=torch.device('cuda', self.eng.device))
        self._side.wait_stream(compute_stream)
        with torch.cuda.stream(self._side):
            dst, done = self._push()
            done.record(self._side)
        return dst


class PeerExchange:
    """The table exchange through the C ABI (kgs_exchange_setup / kgs_exchange_push, include/kgs.h):
    torch supplies only the peer-mapped allocation and its rendezvous (symmetric memory); the
    pushes, the barrier kernel, the double buffering and the ordering against the next job are the
    library's.  One call per step from the host.  Every rank's tables must fit `slot_bytes`."""

    FLAG_BYTES = 4096

    def __init__(self, engine, world, rank, slot_bytes, group=None, multicast=False, self_copy=True):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.eng, self.world, self.rank = engine, world, rank
        self.multicast = False
        self.group = group if group is not None else dist.group.WORLD
        self.slot = (int(slot_bytes) + 255) // 256 * 256
        dev = torch.device('cuda', engine.device)
        agreed = torch.tensor([self.slot], dtype=torch.int64, device=dev)
        dist.all_reduce(agreed, op=dist.ReduceOp.MAX, group=self.group)
        self.slot = int(agreed)
        self.local = symm.empty(2 * world * self.slot + self.FLAG_BYTES, dtype=torch.uint8, device=dev)
        self.local.zero_()
        torch.cuda.synchronize(dev)
        self.hdl = symm.rendezvous(self.local, self.group)
        bases = [int(p) for p in self.hdl.buffer_ptrs]
        flags = [p + 2 * world * self.slot for p in bases]
        engine.exchange_setup(world, rank, bases, flags, self.slot)
        self._base = bases[rank]
        engine.exchange_set_self_copy(self_copy)   # False: this rank reads its own tables from its result region
        if multicast:                           # one push through the switch's multicast address, if every rank has one
            mc = int(getattr(self.hdl, 'multicast_ptr', 0) or 0)
            have = torch.tensor([1 if mc else 0], dtype=torch.int64, device=dev)
            dist.all_reduce(have, op=dist.ReduceOp.MIN, group=self.group)
            if int(have):
                engine.exchange_set_multicast(mc)
                self.multicast = True
        dist.barrier(group=self.group)          # nobody pushes before every rank's flag words are zero

    def push(self, want_view=True):
        """Queue the exchange of the finalized job; returns the [world, bytes] view of the gathered
        tables (filled when the exchange stream has passed the push: exchange_join / exchange_sync).
        want_view=False skips building that view: one C call per step and nothing else on the host."""
        ptr, nbytes = self.eng.exchange_push()
        if not want_view:
            return None
        half = (ptr - self._base) // (self.world * self.slot)
        a = half * self.world * self.slot
        return self.local[a:a + self.world * self.slot].view(self.world, self.slot)[:, :nbytes]


class NcclExchange:
    """The same exchange as ONE ncclAllGather issued by the library (kgs_exchange_nccl) on torch's
    communicator (or on the library's own: Engine.nccl_init).  Receive buffers alternate."""

    def __init__(self, engine, world, slot_bytes, comm=None):
        import torch
        self.eng, self.world, self.slot, self.comm = engine, world, int(slot_bytes), comm
        dev = torch.device('cuda', engine.device)
        self.recv = [torch.empty(world * self.slot, dtype=torch.uint8, device=dev) for _ in range(2)]
        self._i = 0

    @staticmethod
    def torch_comm(device):
        """ncclComm_t of torch's default process group for `device`, as an integer."""
        import torch
        import torch.distributed as dist
        backend = dist.distributed_c10d._get_default_group()._get_backend(torch.device('cuda', device))
        return int(backend._comm_ptr())

    def push(self, want_view=True):
        buf = self.recv[self._i]
        self._i ^= 1
        n = self.eng.exchange_nccl(buf.data_ptr(), buf.numel(), self.comm)
        return buf[:self.world * n].view(self.world, n) if want_view else None


def decode_packed(buf, layout):
    """One shard's packed exchange region (uint8 array) -> {kind: {count,sum,max,mean,pct}}.
    mean and the percentiles are recomputed with the sender's own operation order (one fp64 divide;
    rule R5: rank = q*(n-1), v[lo]*(1-w) + v[hi]*w), so they equal the sender's tables bit for bit."""
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    out = {}
    for kind, t in layout['kinds'].items():
        n = t['n_cells']

        def arr(name, dt):
            off, nb = t[name]
            return buf[off:off + nb].view(dt)
        if layout.get('compact'):
            count = arr('count', np.uint16).astype(np.uint64)
            total = arr('sum', np.uint16).astype(np.uint64) | (arr('sum_hi', np.uint8).astype(np.uint64) << np.uint64(16))
            vmax = arr('max', np.uint8).astype(np.uint32)
            off, nb = layout['escape']                      # cells above 65 534 samples: full numbers from the list
            n_esc = int(buf[off:off + 4].view(np.uint32)[0])
            cap = (nb - 16) // 16
            if n_esc > cap:
                raise ValueError('compact exchange tables: %d cells on an escape list of %d (sender should use the '
                                 'packed format for windows this long)' % (n_esc, cap))
            if n_esc:
                ent = buf[off + 16:off + 16 + 16 * n_esc].view(np.uint32).reshape(n_esc, 4)
                cell = ent[:, 0].astype(np.int64) - layout['cell_begin'][kind]
                mine = (cell >= 0) & (cell < n)
                count[cell[mine]] = ent[mine, 1].astype(np.uint64)
                total[cell[mine]] = ent[mine, 2].astype(np.uint64) | (ent[mine, 3].astype(np.uint64) << np.uint64(32))
        else:
            count = arr('count', np.uint32).astype(np.uint64)
            total = arr('sum', np.uint64 if layout['sum64'] else np.uint32).astype(np.uint64)
            vmax = arr('max', np.uint16).astype(np.uint32)
        cf = count.astype(np.float64)
        mean = np.divide(total.astype(np.float64), cf, out=np.zeros(n), where=count > 0)
        qs = layout['qs']
        pct = np.full((len(qs), n), np.nan)
        if qs:
            ranks = arr('rank', np.uint32 if layout['u16'] else np.uint16).reshape(len(qs), n).astype(np.uint32)
            shift, mask = (16, 0xFFFF) if layout['u16'] else (8, 0xFF)
            nz = count > 0
            for iq, q in enumerate(qs):
                vlo = (ranks[iq] & mask).astype(np.float64)
                vhi = (ranks[iq] >> shift).astype(np.float64)
                rank = np.float64(q) * (cf - 1.0)
                w = rank - np.floor(rank)
                pct[iq, nz] = (vlo * (1.0 - w) + vhi * w)[nz]
        out[kind] = {'count': count, 'sum': total, 'max': vmax, 'mean': mean, 'pct': pct}
    return out


# ---- host-memory variant of the exchange (any torch.distributed backend, e.g. gloo) -------------

_FIELDS = (('count', np.uint64), ('sum', np.uint64), ('max', np.uint32), ('mean', np.float64), ('pct', np.float64))


def pack_tables(table):
    """One kind's {count,sum,max,mean,pct} -> a single uint8 array (header: n_cells, n_q)."""
    n_cells = int(np.asarray(table['count']).size)
    pct = np.asarray(table['pct'], dtype=np.float64).reshape(-1, n_cells) if n_cells else np.zeros((0, 0))
    head = np.array([n_cells, pct.shape[0]], dtype=np.uint64)
    parts = [head.view(np.uint8)]
    for name, dt in _FIELDS:
        arr = pct if name == 'pct' else np.asarray(table[name], dtype=dt)
        parts.append(np.ascontiguousarray(arr, dtype=dt).reshape(-1).view(np.uint8))
    return np.concatenate(parts)


def unpack_tables(buf):
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    n_cells, n_q = (int(x) for x in buf[:16].view(np.uint64))
    out, at = {}, 16
    for name, dt in _FIELDS:
        n = n_cells * (n_q if name == 'pct' else 1)
        nbytes = n * np.dtype(dt).itemsize
        arr = buf[at:at + nbytes].view(dt).copy()
        out[name] = arr.reshape(n_q, n_cells) if name == 'pct' else arr
        at += nbytes
    return out


def all_gather_tables(table):
    """all_gather of one kind's table over the default process group; shards may differ in size
    (buffers are padded to the largest).  Returns the list of per-rank tables, rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size()
    mine = torch.from_numpy(pack_tables(table))
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([mine.numel()], dtype=torch.int64))
    longest = max(int(s) for s in sizes)
    padded = torch.zeros(longest, dtype=torch.uint8)
    padded[:mine.numel()] = mine
    got = [torch.zeros(longest, dtype=torch.uint8) for _ in range(world)]
    dist.all_gather(got, padded)
    return [unpack_tables(g[:int(s)].numpy()) for g, s in zip(got, sizes)]


def cluster_totals(per_type, group=None, device=None):
    """The one cross-shard REDUCTION of the reference: who_use_gpu.py:18-20 sums GPU counts over all
    pods, in total and per GPU type.  Each rank passes the totals of its own shard of pods
    ({gpu type: count}); every rank gets back (grand total, {gpu type: count}) of the whole cluster,
    types in first-seen order across ranks (rank 0's order first).  One all_gather of the type names
    (strings) and one all_reduce(SUM) of a small int64 vector - NCCL when `device` is a CUDA device,
    any backend otherwise."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    names = [None] * world
    dist.all_gather_object(names, list(per_type.keys()), group=group)
    order = []
    for rank_names in names:
        for t in rank_names:
            if t not in order:
                order.append(t)
    vec = torch.tensor([int(per_type.get(t, 0)) for t in order], dtype=torch.int64, device=device)
    if len(order):
        dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
    totals = dict(zip(order, (int(v) for v in vec.tolist())))
    return sum(totals.values()), totals

