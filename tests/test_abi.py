"""CPU: libkgs.so builds, loads, exports every symbol include/kgs.h declares, and refuses to run
without a device instead of falling back (no compute calls here)."""
import ctypes as C
import os
import re

from kube_gpu_stats_b200 import _ffi


def test_library_builds_and_exports_header(built_lib):
    assert os.path.exists(built_lib)
    declared = _ffi.declared_symbols()
    assert len(declared) >= 28 and 'kgs_window_stats' in declared and 'kgs_join_probe_noalloc' in declared
    lib = C.CDLL(built_lib)
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    bound = _ffi.load(built_lib)
    assert bound.kgs_version().startswith(b'libkgs')


def test_binding_covers_every_declared_function(built_lib):
    src = open(_ffi.__file__).read()
    for name in _ffi.declared_symbols():
        assert re.search(r"\b%s\b" % name, src), 'no ctypes signature for ' + name


def test_window_spec_layout_matches_header():
    """ctypes mirror of kgs_window_spec: 10 u32, 8 doubles, one u64 - 40 + 64 + 8 bytes."""
    assert C.sizeof(_ffi.WindowSpec) == 112
    assert _ffi.WindowSpec.q.offset == 40 and _ffi.WindowSpec.total_samples_hint.offset == 104


def test_no_silent_cpu_fallback(built_lib):
    """Without a CUDA device kgs_create must fail loudly; with one it must succeed."""
    lib = _ffi.load(built_lib)
    h = C.c_void_p()
    rc = lib.kgs_create(0, C.byref(h))
    if rc == _ffi.KGS_OK:
        assert lib.kgs_destroy(h) == _ffi.KGS_OK
    else:
        assert rc == _ffi.KGS_ERR_CUDA and not h.value
        assert b'no CUDA device' in lib.kgs_last_error(None)
        from kube_gpu_stats_b200.engine import Engine, KgsError
        try:
            Engine(0)
            raise AssertionError('Engine() must raise without a GPU')
        except KgsError as e:
            assert e.code == _ffi.KGS_ERR_CUDA


def test_null_and_bad_arguments_are_rejected(built_lib):
    lib = _ffi.load(built_lib)
    assert lib.kgs_create(0, None) == _ffi.KGS_ERR_ARG
    assert lib.kgs_sync(None) == _ffi.KGS_ERR_ARG
    assert lib.kgs_stats_accumulate(None) == _ffi.KGS_ERR_ARG
    assert lib.kgs_destroy(None) == _ffi.KGS_OK
    rc = lib.kgs_join_probe_noalloc(0, 16, 16, _ffi.NOALLOC_MAX_PAIRS + 1, 16, 16, 4, 8, 16, None)
    assert rc == _ffi.KGS_ERR_CAPACITY


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under kube_gpu_stats_b200/ may import it."""
    pkg = os.path.dirname(_ffi.__file__)
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f
