"""CPU: host-side logic of the library that needs no GPU - the look at a batch's order at ingest
(kgs_input_order_class, exported for exactly this) and the decoding of the compact exchange tables
(sharding.decode_packed) from a region written by hand, escape list included."""
import ctypes as C

import numpy as np
import pytest

from conftest import assert_stats_equal
from kube_gpu_stats_b200 import sharding, synth


def order_class(built_lib, gid, pid):
    from kube_gpu_stats_b200 import _ffi
    lib = _ffi.lib()
    gid = np.ascontiguousarray(gid, dtype=np.uint32)
    pid = np.ascontiguousarray(pid, dtype=np.uint32)
    return lib.kgs_input_order_class(gid.ctypes.data, pid.ctypes.data, len(gid))


def test_order_classes(built_lib):
    tr = synth.cluster_trace(4, 8, 1, 600, n_tenancies=40, seed=1)
    assert order_class(built_lib, tr.gpu_gid, tr.pid) == 0                      # collector order: long runs
    tr8 = synth.cluster_trace(2, 8, 8, 300, n_tenancies=150, pids_per_pod=2, seed=1)
    assert order_class(built_lib, tr8.gpu_gid, tr8.pid) == 2                    # 8 processes per GPU, second by second
    o = np.argsort(tr.ts, kind='stable')
    assert order_class(built_lib, tr.gpu_gid[o], tr.pid[o]) == 2                # 32 GPUs second by second: period 32
    big = synth.cluster_trace(100, 8, 1, 60, seed=2)
    o = np.argsort(big.ts, kind='stable')
    assert order_class(built_lib, big.gpu_gid[o], big.pid[o]) == 1              # 800 GPUs second by second: period beyond the look
    fl = synth.config(3, scale=0.001)
    assert order_class(built_lib, fl.gpu_gid, fl.pid) == 1                      # a random stream
    assert order_class(built_lib, tr.gpu_gid[:100], tr.pid[:100]) == 0          # too short to judge
    from kube_gpu_stats_b200 import _ffi
    assert _ffi.lib().kgs_input_order_class(None, None, 5) == -1


def test_decode_compact_tables_with_escape_list():
    """A compact region laid out by hand (one kind, 6 cells, two percentiles): count u16 with 0xFFFF for the
    cell on the escape list, sum as u16 + u8, max u8, selected values u8|u8<<8."""
    n, qs = 6, (0.5, 0.95)
    count = np.array([0, 1, 60, 65534, 70000, 3], dtype=np.uint64)
    total = np.array([0, 17, 3000, 6_000_000, 7_000_123, 299], dtype=np.uint64)
    vmax = np.array([0, 17, 99, 100, 255, 100], dtype=np.uint8)
    vlo = np.array([[0, 17, 50, 91, 100, 99], [0, 17, 97, 99, 200, 100]], dtype=np.uint16)
    vhi = np.array([[0, 17, 51, 92, 101, 100], [0, 17, 98, 100, 201, 100]], dtype=np.uint16)
    cell_begin = 1000                                          # cells of an earlier kind precede these in the global numbering
    esc_cap = 4
    buf = np.zeros(4096, dtype=np.uint8)
    buf[0:4] = np.array([2], dtype=np.uint32).view(np.uint8)   # two entries: one of this kind, one of another
    ent = np.array([[cell_begin + 4, 70000, 7_000_123 & 0xFFFFFFFF, 7_000_123 >> 32], [5, 99999, 1, 0]], dtype=np.uint32)
    buf[16:16 + 32] = ent.view(np.uint8).reshape(-1)
    off = {'sum': 256, 'sum_hi': 512, 'count': 768, 'max': 1024, 'rank': 1280}
    c16 = np.where(count > 65534, 0xFFFF, count).astype(np.uint16)
    buf[off['count']:off['count'] + 2 * n] = c16.view(np.uint8)
    buf[off['sum']:off['sum'] + 2 * n] = (total & 0xFFFF).astype(np.uint16).view(np.uint8)
    buf[off['sum_hi']:off['sum_hi'] + n] = ((total >> 16) & 0xFF).astype(np.uint8)
    buf[off['max']:off['max'] + n] = vmax
    buf[off['rank']:off['rank'] + 4 * n] = (vlo | (vhi << 8)).astype(np.uint16).reshape(-1).view(np.uint8)
    layout = {'n_windows': 3, 'qs': list(qs), 'u16': False, 'sum64': False, 'compact': True, 'escape': (0, 16 + 16 * esc_cap),
              'cell_begin': {'gpu': cell_begin},
              'kinds': {'gpu': {'count': (off['count'], 2 * n), 'sum': (off['sum'], 2 * n), 'sum_hi': (off['sum_hi'], n),
                                'max': (off['max'], n), 'rank': (off['rank'], 4 * n), 'n_cells': n}}}
    got = sharding.decode_packed(buf, layout)['gpu']
    cf = count.astype(np.float64)
    mean = np.divide(total.astype(np.float64), cf, out=np.zeros(n), where=count > 0)
    pct = np.full((2, n), np.nan)
    for iq, q in enumerate(qs):
        rank = q * (cf - 1.0)
        w = rank - np.floor(rank)
        v = vlo[iq].astype(np.float64) * (1.0 - w) + vhi[iq].astype(np.float64) * w
        pct[iq, count > 0] = v[count > 0]
    assert_stats_equal(got, {'count': count, 'sum': total, 'max': vmax.astype(np.uint32), 'mean': mean, 'pct': pct})
    buf[0:4] = np.array([esc_cap + 1], dtype=np.uint32).view(np.uint8)          # more entries than the list holds: loud
    with pytest.raises(ValueError, match='escape list'):
        sharding.decode_packed(buf, layout)
