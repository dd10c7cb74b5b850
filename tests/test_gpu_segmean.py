"""The reference's per-series mean (gpu_util_stats.py:77-80, :51) on the GPU vs the oracle."""
import numpy as np
import pytest

from conftest import bits
from oracle import ref_gpu_util_stats as ref
from oracle import window_stats as ws

pytestmark = pytest.mark.gpu


def test_short_series_bit_exact(engine):
    """<= 64 points per series (the reference asks for 25): naive left-to-right sum, bit for bit."""
    rng = np.random.default_rng(0)
    lens = rng.integers(0, 65, size=3000)
    lens[:3] = [0, 1, 64]
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    vals = rng.uniform(0, 100, size=int(off[-1]))
    got = engine.segmented_mean(vals, off)
    want = np.array(ws.segmented_mean_py(vals, off))
    np.testing.assert_array_equal(bits(got), bits(want))
    assert got[0] == 0.0                                     # empty series -> 0.0 (GUS:79-80)


def test_long_series_within_1e12(engine):
    rng = np.random.default_rng(1)
    lens = np.array([65, 100, 3600, 86400, 7, 0, 1000003])
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    vals = rng.uniform(0, 100, size=int(off[-1]))
    got = engine.segmented_mean(vals, off)
    want = np.array(ws.segmented_mean_py(vals, off))
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=0)


def test_server_weighting(engine):
    """GUS:51: (sum * step_sec) / time_range_sec, in that order."""
    rng = np.random.default_rng(2)
    lens = rng.integers(1, 26, size=200)
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    vals = rng.uniform(0, 100, size=int(off[-1]))
    got = engine.segmented_mean(vals, off, mul=3600, div=86400)
    want = np.array([ref.naive_sum(vals[int(a):int(b)]) * 3600 / 86400 for a, b in zip(off[:-1], off[1:])])
    np.testing.assert_array_equal(bits(got), bits(want))
