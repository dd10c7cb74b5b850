"""GPU: property tests (hypothesis) - random small traces in arbitrary order, random window specs,
random join tables with repeated keys and misses.  Everything must equal the oracle bit for bit."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from conftest import assert_stats_equal
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu

KINDS = {'pod': ws.KEY_POD, 'gpu': ws.KEY_GPU, 'node': ws.KEY_NODE}


@st.composite
def traces(draw):
    seed = draw(st.integers(0, 2**31 - 1))
    n = draw(st.sampled_from([0, 1, 7, 8, 9, 255, 256, 257, 700, 3000]))
    order = draw(st.sampled_from(['collector', 'shuffled', 'by_time']))
    procs = draw(st.integers(1, 4))
    step = draw(st.sampled_from([1, 5, 7, 60]))
    wmul = draw(st.sampled_from([0, 1, 1, 3]))               # 0: sampled windows (width < step)
    n_windows = draw(st.integers(1, 40))
    t0_off = draw(st.integers(-20, 50))
    metric = draw(st.sampled_from(['sm', 'mem', 'fb']))
    qs = draw(st.sampled_from([(), (0.5,), (0.0, 0.95, 1.0), (0.99, 0.5)]))
    return seed, n, order, procs, step, wmul, n_windows, t0_off, metric, qs


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])
@given(traces())
def test_random_traces_match_oracle(engine, case):
    seed, n, order, procs, step, wmul, n_windows, t0_off, metric, qs = case
    rng = np.random.default_rng(seed)
    gpn, n_gpus = 4, 8
    # n samples: gpu streams with `procs` interleaved processes, 1 Hz from second 1000
    gid = np.sort(rng.integers(0, n_gpus + 1, size=n)).astype(np.uint32)      # gpu 8 is out of the key domain (R6)
    slot = rng.integers(0, procs, size=n)
    ts = (1000 + np.arange(n) // max(1, procs * 2)).astype(np.uint32)
    pid = (100 + slot + 10 * (gid % 3)).astype(np.uint32)
    util = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    if metric != 'fb':
        util &= np.uint32(0x7F7F7F7F) if rng.random() < 0.5 else np.uint32(0xFFFFFFFF)
    if order == 'shuffled':
        p = rng.permutation(n)
        ts, gid, pid, util = ts[p], gid[p], pid[p], util[p]
    elif order == 'by_time':
        p = np.argsort(ts, kind='stable')
        ts, gid, pid, util = ts[p], gid[p], pid[p], util[p]
    m = int(rng.integers(0, 12))
    keys = (rng.integers(0, 3, size=m).astype(np.uint64) << np.uint64(32)) | rng.integers(100, 125, size=m).astype(np.uint64)
    pods = rng.integers(0, 6, size=m).astype(np.uint32)
    width = step * wmul if wmul else max(1, step // 2)
    t0 = 1000 + t0_off
    engine.build_join(keys, pods, 5)                                             # pod 5 is out of the key domain
    engine.ingest(ts, gid, pid, util)
    np.testing.assert_array_equal(engine.probe(gpn), ws.probe_np(keys, pods, gid, pid, gpn))
    got = engine.window_stats(kinds=('pod', 'gpu', 'node'), metric=metric, gpus_per_node=gpn, t0=t0, step=step,
                              width=width, n_windows=n_windows, n_gpus=n_gpus, n_nodes=2, qs=qs)
    for kind, nk in (('pod', 5), ('gpu', n_gpus), ('node', 2)):
        want = ws.window_stats_np(ts, gid, pid, util, keys, pods, key_kind=KINDS[kind], metric={'sm': 0, 'mem': 1, 'fb': 2}[metric],
                                  gpus_per_node=gpn, t0=t0, step=step, width=width, n_windows=n_windows, n_keys=nk, qs=qs)
        assert_stats_equal(got[kind], want, what='%s %s' % (kind, case))
