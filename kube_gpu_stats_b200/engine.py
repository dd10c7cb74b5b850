"""Engine: a thin object over one libkgs handle (one GPU, two CUDA streams).

This is host plumbing only - every computation happens in libkgs.so.  The drop-in
modules (``gpu_util_stats.py``, ``who_use_gpu.py``), ``bench.py`` and the tests
all go through this class, i.e. through the C ABI of include/kgs.h.
"""
import ctypes as C

import numpy as np

from . import _ffi

KIND_ID = {'pod': _ffi.KEY_POD, 'gpu': _ffi.KEY_GPU, 'node': _ffi.KEY_NODE}
METRIC_ID = {'sm': _ffi.METRIC_SM, 'mem': _ffi.METRIC_MEM, 'fb': _ffi.METRIC_FB}


class KgsError(RuntimeError):
    def __init__(self, code, text):
        super().__init__('libkgs error %d: %s' % (code, text))
        self.code = code


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


class Engine:
    def __init__(self, device=0):
        self._lib = _ffi.lib()
        h = C.c_void_p()
        rc = self._lib.kgs_create(int(device), C.byref(h))
        if rc:
            raise KgsError(rc, self._lib.kgs_last_error(None).decode())
        self._h = h
        self.device = device
        self._keep = []          # host arrays an asynchronous copy may still be reading
        self._pinned = []
        self._spec = None
        self._n_keys = {}
        self.n_resident = 0
        self._matrix = (0, 0)
        self.matrix_generation = 0   # bumped by every matrix_parse: a holder of a decoded answer can tell it is still resident

    # -- plumbing ---------------------------------------------------------------------
    def _ck(self, rc):
        if rc:
            raise KgsError(rc, self._lib.kgs_last_error(self._h).decode())

    def close(self):
        if self._h:
            for p in self._pinned:
                self._lib.kgs_free_pinned(self._h, p)
            self._pinned = []
            self._lib.kgs_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        self._ck(self._lib.kgs_sync(self._h))
        self._keep = []

    def pinned(self, n, dtype=np.uint32):
        """A numpy array over n elements of page-locked host memory (freed on close)."""
        nbytes = int(n) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self._ck(self._lib.kgs_alloc_pinned(self._h, nbytes, C.byref(p)))
        self._pinned.append(p)
        buf = (C.c_char * max(nbytes, 1)).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype, count=int(n))

    # -- ingestion --------------------------------------------------------------------
    def ingest(self, ts, gpu_gid, pid, util):
        """Queue the copy of one batch (four u32 columns) to the device; asynchronous."""
        cols = [_u32(ts), _u32(gpu_gid), _u32(pid), _u32(util)]
        n = len(cols[0])
        if any(len(c) != n for c in cols):
            raise ValueError('columns differ in length')
        self._keep.append(cols)
        self._ck(self._lib.kgs_ingest(self._h, *[c.ctypes.data for c in cols], n))
        self.n_resident = n

    def batch_reserve(self, capacity):
        """Make an empty batch of the given capacity resident; fill it with batch_append."""
        self._ck(self._lib.kgs_batch_reserve(self._h, int(capacity)))
        self.n_resident = 0

    def batch_append(self, ts, gpu_gid, pid, util):
        cols = [_u32(ts), _u32(gpu_gid), _u32(pid), _u32(util)]
        n = len(cols[0])
        self._keep.append(cols)
        self._ck(self._lib.kgs_batch_append(self._h, *[c.ctypes.data for c in cols], n))
        self.n_resident += n

    def reorder(self):
        """Sort the resident batch into (gpu, pid, input order) on the GPU (see kgs_reorder_batch)."""
        self._ck(self._lib.kgs_reorder_batch(self._h))
        self._keep = []

    def set_auto_reorder(self, on):
        """Accumulate batches that interleave their series from a series-major copy (default on)."""
        self._ck(self._lib.kgs_set_auto_reorder(self._h, 1 if on else 0))

    def fetch_batch(self):
        """The resident batch's four columns back on the host (testing / inspection)."""
        cols = [np.empty(self.n_resident, dtype=np.uint32) for _ in range(4)]
        self._ck(self._lib.kgs_fetch_batch(self._h, *[c.ctypes.data for c in cols]))
        return cols

    def adopt_device_batch(self, d_ts, d_gid, d_pid, d_util, n):
        self._ck(self._lib.kgs_adopt_device_batch(self._h, int(d_ts), int(d_gid), int(d_pid), int(d_util), int(n)))
        self.n_resident = int(n)

    # -- join -------------------------------------------------------------------------
    def build_join(self, keys, pod_ids, n_pods):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        pod_ids = _u32(pod_ids)
        if len(keys) != len(pod_ids):
            raise ValueError('keys and pod_ids differ in length')
        self._ck(self._lib.kgs_build_join(self._h, keys.ctypes.data, pod_ids.ctypes.data, len(keys), int(n_pods)))
        self._n_keys['pod'] = int(n_pods)

    def probe(self, gpus_per_node=8):
        out = np.empty(self.n_resident, dtype=np.uint32)
        self._ck(self._lib.kgs_probe(self._h, int(gpus_per_node), out.ctypes.data))
        self._keep = []
        return out

    # -- windowed statistics ------------------------------------------------------------
    def make_spec(self, kinds=('pod',), metric='sm', gpus_per_node=8, t0=0, step=60, width=None,
                  n_windows=1, n_gpus=0, n_nodes=0, qs=(), total_samples_hint=0):
        sp = _ffi.WindowSpec()
        sp.key_mask = 0
        for k in kinds:
            sp.key_mask |= 1 << KIND_ID[k]
        sp.metric = METRIC_ID[metric] if isinstance(metric, str) else int(metric)
        sp.gpus_per_node = gpus_per_node
        sp.t0, sp.step, sp.width = int(t0), int(step), int(step if width is None else width)
        sp.n_windows, sp.n_gpus, sp.n_nodes = int(n_windows), int(n_gpus), int(n_nodes)
        if len(qs) > _ffi.MAX_Q:
            raise ValueError('at most %d percentiles' % _ffi.MAX_Q)
        sp.n_q = len(qs)
        for i, q in enumerate(qs):
            sp.q[i] = float(q)
        sp.total_samples_hint = int(total_samples_hint)
        return sp

    def set_run_log_capacity(self, entries):
        self._ck(self._lib.kgs_set_run_log_capacity(self._h, int(entries)))

    def stats_begin(self, spec):
        self._spec = spec
        self._n_keys['gpu'] = spec.n_gpus
        self._n_keys['node'] = spec.n_nodes
        self._ck(self._lib.kgs_stats_begin(self._h, C.byref(spec)))

    def stats_accumulate(self):
        self._ck(self._lib.kgs_stats_accumulate(self._h))

    def stats_finalize(self):
        self._ck(self._lib.kgs_stats_finalize(self._h))

    def window_stats_device(self, spec):
        """begin + accumulate + finalize on the resident batch; results stay on the device."""
        self._spec = spec
        self._n_keys['gpu'] = spec.n_gpus
        self._n_keys['node'] = spec.n_nodes
        self._ck(self._lib.kgs_window_stats(self._h, C.byref(spec)))

    def job_counter(self, which):
        """One of the job's kernel-side counters (kgs_job_counter): 'pool_entries', 'cache_ctas',
        'big_cells', 'wide_cells'."""
        idx = {'pool_entries': _ffi.JC_POOL_ENTRIES, 'cache_ctas': _ffi.JC_CACHE_CTAS,
               'big_cells': _ffi.JC_BIG_CELLS, 'wide_cells': _ffi.JC_WIDE_CELLS}[which]
        out = C.c_uint64()
        self._ck(self._lib.kgs_job_counter(self._h, idx, C.byref(out)))
        return out.value

    def fetch(self, kind, want=('count', 'sum', 'max', 'mean', 'pct')):
        sp = self._spec
        ncell = self._n_keys[kind] * sp.n_windows
        out = {}
        if 'count' in want:
            out['count'] = np.empty(ncell, dtype=np.uint32)
        if 'sum' in want:
            out['sum'] = np.empty(ncell, dtype=np.uint64)
        if 'max' in want:
            out['max'] = np.empty(ncell, dtype=np.uint32)
        if 'mean' in want:
            out['mean'] = np.empty(ncell, dtype=np.float64)
        if 'pct' in want:
            out['pct'] = np.empty((sp.n_q, ncell), dtype=np.float64)

        def ptr(name):
            return out[name].ctypes.data if name in out and out[name].size else None

        self._ck(self._lib.kgs_fetch_stats(self._h, KIND_ID[kind], ptr('count'), ptr('sum'), ptr('max'),
                                           ptr('mean'), ptr('pct')))
        self._keep = []
        if 'count' in out:
            out['count'] = out['count'].astype(np.uint64)
        return out

    def fetch_all(self, kinds, with_mean=True):
        """Every table of the given kinds with ONE device-to-host copy of the result region into a
        pinned staging buffer.  Returns {kind: {count,sum,max,mean,pct}} as views of that buffer
        (valid until the next fetch_all)."""
        _, exchange, total = self.result_region()
        nbytes = total if with_mean else exchange
        if getattr(self, '_stage', None) is None or self._stage.size < nbytes:
            self._stage = self.pinned(nbytes + 4096, dtype=np.uint8)
        self._ck(self._lib.kgs_fetch_region(self._h, self._stage.ctypes.data, nbytes))
        self._keep = []
        sp = self._spec
        out = {}
        fields = (('count', 0, np.uint32), ('sum', 1, np.uint64), ('max', 2, np.uint32), ('pct', 4, np.float64))
        for k in kinds:
            ncell = self._n_keys[k] * sp.n_windows
            t = {}
            for name, which, dt in fields + ((('mean', 3, np.float64),) if with_mean else ()):
                off, nb = C.c_size_t(), C.c_size_t()
                self._ck(self._lib.kgs_result_offset(self._h, KIND_ID[k], which, C.byref(off), C.byref(nb)))
                arr = self._stage[off.value:off.value + nb.value].view(dt)
                t[name] = arr.reshape(sp.n_q, ncell) if name == 'pct' else arr
            if not with_mean:
                cnt = t['count'].astype(np.float64)
                t['mean'] = np.divide(t['sum'].astype(np.float64), cnt, out=np.zeros(ncell), where=cnt > 0)
            out[k] = t
        return out

    def window_stats(self, kinds=('pod',), **kw):
        """One-shot job on the resident batch.  Returns {kind: {count,sum,max,mean,pct}}.
        If the run log overflows (input far from collector order) the job is repeated once
        with a run log sized for the worst case."""
        spec = self.make_spec(kinds=kinds, **kw)
        for attempt in (0, 1):
            self.window_stats_device(spec)
            try:
                return {k: self.fetch(k) for k in kinds}
            except KgsError as e:
                if e.code != _ffi.KGS_ERR_CAPACITY or attempt:
                    raise
                self.set_run_log_capacity(2 * (self.n_resident + 4096) * len(kinds))
        raise AssertionError('unreachable')

    def result_device_ptr(self, kind, which):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._lib.kgs_result_device_ptr(self._h, KIND_ID[kind], int(which), C.byref(p), C.byref(n)))
        return p.value, n.value

    def result_region(self):
        """(device address, exchange bytes, total bytes) of the finalized job's result region."""
        p, e, t = C.c_void_p(), C.c_size_t(), C.c_size_t()
        self._ck(self._lib.kgs_result_region(self._h, C.byref(p), C.byref(e), C.byref(t)))
        return p.value, e.value, t.value

    def set_packed_exchange(self, on):
        """From the next job on, finalize also writes the packed exchange tables (include/kgs.h):
        False/0 off, True/1 packed (10 + 2*n_q bytes per cell), 2 or 'compact' (6 + 2*n_q)."""
        mode = 2 if on == 'compact' else int(on)
        self._ck(self._lib.kgs_set_packed_exchange(self._h, mode))

    def exchange_region(self):
        """(device address, bytes, packed flag) of what one collective has to move for this job."""
        p, n, k = C.c_void_p(), C.c_size_t(), C.c_int()
        self._ck(self._lib.kgs_exchange_region(self._h, C.byref(p), C.byref(n), C.byref(k)))
        return p.value, n.value, k.value

    def packed_layout(self, kinds):
        """{kind: {table: (offset, nbytes)}} of the packed exchange tables, plus what decoding needs."""
        sp = self._spec
        mode = self.exchange_region()[2]
        out = {'n_windows': sp.n_windows, 'qs': [sp.q[i] for i in range(sp.n_q)], 'u16': sp.metric == _ffi.METRIC_FB,
               'sum64': mode == 2, 'compact': mode == 3, 'kinds': {}, 'cell_begin': {}}
        begin = 0
        for name, kid in (('pod', 0), ('gpu', 1), ('node', 2)):      # global cell numbering: kinds of the job, in this order
            if sp.key_mask & (1 << kid):
                out['cell_begin'][name] = begin
                begin += self._n_keys[name] * sp.n_windows
        if mode == 3:
            off, nb = C.c_size_t(), C.c_size_t()
            self._ck(self._lib.kgs_packed_offset(self._h, KIND_ID[kinds[0]], 6, C.byref(off), C.byref(nb)))
            out['escape'] = (off.value, nb.value)
        for k in kinds:
            t = {}
            for name, which in (('count', 0), ('sum', 1), ('max', 2), ('rank', 4)) + ((('sum_hi', 5),) if mode == 3 else ()):
                off, nb = C.c_size_t(), C.c_size_t()
                self._ck(self._lib.kgs_packed_offset(self._h, KIND_ID[k], which, C.byref(off), C.byref(nb)))
                t[name] = (off.value, nb.value)
            t['n_cells'] = self._n_keys[k] * sp.n_windows
            out['kinds'][k] = t
        return out

    # -- table exchange between the GPUs of one box (kgs_exchange_*) ---------------------
    def exchange_setup(self, n_ranks, my_rank, gathered_ptrs, flag_ptrs, slot_bytes):
        """Register the peer-mapped gathered buffers and flag words (device addresses valid on THIS
        device, one per rank; see include/kgs.h)."""
        arr = C.c_void_p * int(n_ranks)
        self._ck(self._lib.kgs_exchange_setup(self._h, int(n_ranks), int(my_rank), arr(*[int(p) for p in gathered_ptrs]),
                                              arr(*[int(p) for p in flag_ptrs]), int(slot_bytes)))

    def exchange_set_multicast(self, mc_ptr):
        """One multicast address for all ranks' gathered buffers (kgs_exchange_set_multicast); 0/None = off."""
        self._ck(self._lib.kgs_exchange_set_multicast(self._h, int(mc_ptr) if mc_ptr else None))

    def exchange_set_self_copy(self, on):
        self._ck(self._lib.kgs_exchange_set_self_copy(self._h, 1 if on else 0))

    def exchange_push(self):
        """Push the finalized job's tables into every rank's gathered buffer (asynchronous, exchange
        stream).  Returns (device address of this job's gathered half on this device, bytes per rank)."""
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._lib.kgs_exchange_push(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def exchange_join(self):
        self._ck(self._lib.kgs_exchange_join(self._h))

    def exchange_sync(self):
        self._ck(self._lib.kgs_exchange_sync(self._h))

    def nccl_unique_id(self):
        buf = (C.c_char * 128)()
        rc = self._lib.kgs_nccl_unique_id(buf)
        if rc:
            raise KgsError(rc, self._lib.kgs_last_error(None).decode())
        return bytes(buf)

    def nccl_init(self, unique_id, n_ranks, rank):
        buf = (C.c_char * 128).from_buffer_copy(unique_id)
        self._ck(self._lib.kgs_nccl_init(self._h, buf, int(n_ranks), int(rank)))

    def exchange_nccl(self, recv_ptr, recv_bytes, comm=None):
        """One ncclAllGather of the finalized job's tables into recv_ptr (asynchronous, exchange
        stream); comm = a caller's ncclComm_t as an integer, None = the handle's own (nccl_init)."""
        n = C.c_size_t()
        self._ck(self._lib.kgs_exchange_nccl(self._h, comm, int(recv_ptr), int(recv_bytes), C.byref(n)))
        return n.value

    # -- the reference's statistic ------------------------------------------------------
    def segmented_mean(self, values, offsets, mul=0.0, div=0.0):
        values = np.ascontiguousarray(values, dtype=np.float64)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        ns = len(offsets) - 1
        out = np.empty(max(ns, 0), dtype=np.float64)
        if ns > 0:
            self._ck(self._lib.kgs_segmented_mean_f64(self._h, values.ctypes.data, offsets.ctypes.data, ns,
                                                      float(mul), float(div), out.ctypes.data))
        return out

    # -- value decode of a Prometheus answer (gpu_util_stats.py:25,:47-48,:73-74) ---------
    def matrix_parse(self, text, result_type='matrix'):
        """Decode the body of a /query_range (``matrix``) or /query (``vector``) answer on the GPU.
        ``text``: bytes-like, or (address, nbytes) of a pinned buffer.  Returns (n_series, n_samples);
        the CSR stays in device memory for matrix_mean / matrix_fetch.  Raises ValueError where
        the reference's ``float()`` / ``.json()`` would."""
        if isinstance(text, tuple):
            addr, nbytes = text
            keep = None
        else:
            keep = np.frombuffer(text, dtype=np.uint8)
            addr, nbytes = keep.ctypes.data if len(keep) else 0, len(keep)
        kind = {'matrix': _ffi.RESULT_MATRIX, 'vector': _ffi.RESULT_VECTOR}[result_type]
        ns, nv = C.c_uint64(), C.c_uint64()
        self._matrix = (0, 0)
        self.matrix_generation += 1
        rc = self._lib.kgs_matrix_parse(self._h, addr, nbytes, kind, C.byref(ns), C.byref(nv))
        if rc == _ffi.KGS_ERR_PARSE:
            raise ValueError(self._lib.kgs_last_error(self._h).decode())
        self._ck(rc)
        self._matrix = (ns.value, nv.value)
        return self._matrix

    def matrix_fetch(self, ts=True, values=True, spans=True):
        """(ts, values, offsets, spans) of the decoded answer as numpy arrays (None where not asked)."""
        n_series, n_samples = self._matrix
        a_ts = np.empty(n_samples, dtype=np.float64) if ts else None
        a_val = np.empty(n_samples, dtype=np.float64) if values else None
        a_off = np.empty(n_series + 1, dtype=np.uint64)
        a_sp = np.empty((n_series, 4), dtype=np.uint64) if spans else None
        ptr = lambda a: a.ctypes.data if a is not None and a.size else None
        self._ck(self._lib.kgs_matrix_fetch(self._h, ptr(a_ts), ptr(a_val), a_off.ctypes.data, ptr(a_sp)))
        return a_ts, a_val, a_off, a_sp

    def matrix_mean(self, mul=0.0, div=0.0):
        """Per-series statistic of the reference over the decoded answer (see segmented_mean)."""
        out = np.empty(self._matrix[0], dtype=np.float64)
        self._ck(self._lib.kgs_matrix_mean(self._h, float(mul), float(div), out.ctypes.data if len(out) else None))
        return out

    # -- measurement --------------------------------------------------------------------
    def launch_count(self):
        n = C.c_uint64()
        self._ck(self._lib.kgs_launch_count(self._h, C.byref(n)))
        return n.value

    def set_profiling(self, on):
        self._ck(self._lib.kgs_set_profiling(self._h, 1 if on else 0))

    def reset_kernel_times(self):
        self._ck(self._lib.kgs_reset_kernel_times(self._h))

    def kernel_times(self):
        """{kernel name: (total ms, launches)} accumulated while profiling was on."""
        out = {}
        for kid, name in enumerate(_ffi.KERNEL_NAMES):
            ms, n = C.c_double(), C.c_uint64()
            self._ck(self._lib.kgs_kernel_time(self._h, kid, C.byref(ms), C.byref(n)))
            out[name] = (ms.value, n.value)
        return out

    def stream(self, which='compute'):
        """cudaStream_t of the handle's copy / compute / exchange stream as an integer."""
        p = C.c_void_p()
        self._ck(self._lib.kgs_stream(self._h, {'copy': 0, 'compute': 1, 'exchange': 2}[which], C.byref(p)))
        return p.value or 0

    def mark(self, which):
        self._ck(self._lib.kgs_mark(self._h, which))

    def elapsed_ms(self):
        ms = C.c_double()
        self._ck(self._lib.kgs_elapsed_ms(self._h, C.byref(ms)))
        return ms.value
