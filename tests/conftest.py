import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def built_lib():
    """libkgs.so, compiled if stale (nvcc cross-compiles without a GPU)."""
    from kube_gpu_stats_b200 import build
    return build.build_lib()


@pytest.fixture()
def engine(built_lib):
    from kube_gpu_stats_b200.engine import Engine
    eng = Engine(0)
    yield eng
    eng.close()


def bits(a):
    """fp64 arrays compared bit for bit (NaN included)."""
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


def assert_stats_equal(got, want, what=''):
    for name in ('count', 'sum', 'max'):
        np.testing.assert_array_equal(np.asarray(got[name]).astype(np.uint64),
                                      np.asarray(want[name]).astype(np.uint64), err_msg='%s %s' % (what, name))
    np.testing.assert_array_equal(bits(got['mean']), bits(want['mean']), err_msg='%s mean bits' % what)
    gp, wp = np.asarray(got['pct'], dtype=np.float64), np.asarray(want['pct'], dtype=np.float64)
    assert gp.shape == wp.shape, (what, gp.shape, wp.shape)
    np.testing.assert_array_equal(bits(gp), bits(wp), err_msg='%s percentile bits' % what)
