"""(node,pid)->pod join through the C ABI vs the oracle (rule R2).  Bit-exact."""
import numpy as np
import pytest

from oracle import window_stats as ws

pytestmark = pytest.mark.gpu


def _random_join(rng, m, n, n_nodes=50, gpn=8, dup_frac=0.2, miss_frac=0.3):
    node = rng.integers(0, n_nodes, size=m).astype(np.uint64)
    pid = rng.integers(1, 5000, size=m).astype(np.uint64)
    keys = (node << np.uint64(32)) | pid
    ndup = int(m * dup_frac)
    if ndup:                                   # repeated keys: the LAST pair must win
        keys[rng.integers(0, m, size=ndup)] = keys[rng.integers(0, m, size=ndup)]
    pods = rng.integers(0, 1000, size=m).astype(np.uint32)
    pick = rng.integers(0, m, size=n)
    g_node = (keys[pick] >> np.uint64(32)).astype(np.uint32)
    g_pid = (keys[pick] & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    miss = rng.random(n) < miss_frac
    g_pid[miss] += 7000                        # pids that are in no pair
    gid = g_node * gpn + rng.integers(0, gpn, size=n).astype(np.uint32)
    return keys, pods, gid.astype(np.uint32), g_pid


@pytest.mark.parametrize('m,n', [(1, 1), (10, 3), (1000, 4096), (5000, 100003), (200000, 1000001)])
def test_probe_matches_oracle(engine, m, n):
    rng = np.random.default_rng(m * 31 + n)
    keys, pods, gid, pid = _random_join(rng, m, n)
    engine.build_join(keys, pods, 1000)
    z = np.zeros(n, dtype=np.uint32)
    engine.ingest(z, gid, pid, z)
    got = engine.probe(8)
    want = ws.probe_np(keys, pods, gid, pid, 8)
    np.testing.assert_array_equal(got, want)
    if m <= 1000:                              # the dict statement of R2 on small cases
        table = ws.build_table(keys, pods)
        np.testing.assert_array_equal(got, np.array(ws.probe_py(table, gid, pid, 8), dtype=np.uint32))


def test_probe_empty_table_and_empty_batch(engine):
    engine.build_join(np.zeros(0, np.uint64), np.zeros(0, np.uint32), 0)
    gid = np.arange(10, dtype=np.uint32)
    engine.ingest(gid, gid, gid, gid)
    np.testing.assert_array_equal(engine.probe(8), np.full(10, ws.POD_NONE, np.uint32))
    e = np.zeros(0, np.uint32)
    engine.ingest(e, e, e, e)
    assert engine.probe(8).shape == (0,)


def test_rebuild_replaces_table(engine):
    gid = np.array([0, 8, 16], np.uint32)
    pid = np.array([5, 5, 5], np.uint32)
    engine.ingest(gid, gid, pid, gid)
    engine.build_join(np.array([5, (1 << 32) | 5], np.uint64), np.array([1, 2], np.uint32), 3)
    np.testing.assert_array_equal(engine.probe(8), [1, 2, ws.POD_NONE])
    engine.build_join(np.array([(2 << 32) | 5], np.uint64), np.array([9], np.uint32), 10)
    np.testing.assert_array_equal(engine.probe(8), [ws.POD_NONE, ws.POD_NONE, 9])


def test_errors(engine):
    from kube_gpu_stats_b200.engine import KgsError
    from kube_gpu_stats_b200 import _ffi
    with pytest.raises(KgsError) as e:
        engine.probe(8)
    assert e.value.code == _ffi.KGS_ERR_STATE
    with pytest.raises(KgsError) as e:
        engine.build_join(np.array([0xFFFFFFFFFFFFFFFF], np.uint64), np.array([1], np.uint32), 2)
    assert e.value.code == _ffi.KGS_ERR_ARG
