"""ctypes binding of libkgs.so (the C ABI in include/kgs.h).

There is no fallback: if the library has not been built, or exports less than the
header declares, importing this module raises.
"""
import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('KGS_LIB_PATH') or os.path.join(HERE, 'libkgs.so')   # KGS_LIB_PATH: experiments with another build
HEADER_PATH = os.path.join(HERE, '..', 'include', 'kgs.h')

KGS_OK, KGS_ERR_ARG, KGS_ERR_CUDA, KGS_ERR_STATE, KGS_ERR_CAPACITY, KGS_ERR_NOMEM, KGS_ERR_PARSE = range(7)
RESULT_MATRIX, RESULT_VECTOR = 0, 1
KEY_POD, KEY_GPU, KEY_NODE = 0, 1, 2
METRIC_SM, METRIC_MEM, METRIC_FB = 0, 1, 2
POD_NONE = 0xFFFFFFFF
MAX_Q = 8
NOALLOC_MAX_PAIRS = 65536
JC_POOL_ENTRIES, JC_CACHE_CTAS, JC_BIG_CELLS, JC_WIDE_CELLS = range(4)
(K_BUILD, K_BUILD_FIN, K_PROBE, K_ACCUM, K_FINALIZE, K_FINALIZE_BIG, K_SEGMEAN, K_CLEAR, K_FINALIZE_WIDE, K_FINALIZE_BASIC, K_SORT, K_PARSE, K_FINALIZE_GROUPS, K_EXCHANGE) = range(14)
KERNEL_NAMES = ['join_build', 'join_resolve', 'join_probe', 'accum', 'finalize_cells', 'finalize_big',
                'segmean', 'join_clear', 'finalize_wide', 'finalize_basic', 'reorder', 'parse', 'finalize_groups', 'exchange']


class WindowSpec(C.Structure):
    _fields_ = [('key_mask', C.c_uint32), ('metric', C.c_uint32), ('gpus_per_node', C.c_uint32),
                ('t0', C.c_uint32), ('step', C.c_uint32), ('width', C.c_uint32),
                ('n_windows', C.c_uint32), ('n_gpus', C.c_uint32), ('n_nodes', C.c_uint32),
                ('n_q', C.c_uint32), ('q', C.c_double * MAX_Q), ('total_samples_hint', C.c_uint64)]


def declared_symbols(header_path=HEADER_PATH):
    """Names of every function the header declares (used by the tests and by load())."""
    text = open(header_path).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(kgs_[a-z0-9_]+)\s*\(', text)))


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise RuntimeError(
            'libkgs.so is missing (%s). Build it: python -c "import __graft_entry__ as g; g.build()" '
            '- kube_gpu_stats_b200 has no CPU or pure-Python path.' % path)
    lib = C.CDLL(path)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    if missing:
        raise RuntimeError('libkgs.so does not export: ' + ', '.join(missing))
    H, u32p, u64p, f64p, vp = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_double), C.c_void_p
    sig = {
        'kgs_create': [C.c_int, C.POINTER(H)],
        'kgs_destroy': [H],
        'kgs_sync': [H],
        'kgs_alloc_pinned': [H, C.c_size_t, C.POINTER(vp)],
        'kgs_free_pinned': [H, vp],
        'kgs_ingest': [H, vp, vp, vp, vp, C.c_uint64],
        'kgs_batch_reserve': [H, C.c_uint64],
        'kgs_batch_append': [H, vp, vp, vp, vp, C.c_uint64],
        'kgs_reorder_batch': [H],
        'kgs_set_auto_reorder': [H, C.c_int],
        'kgs_input_order_class': [vp, vp, C.c_uint64],
        'kgs_fetch_batch': [H, vp, vp, vp, vp],
        'kgs_adopt_device_batch': [H, vp, vp, vp, vp, C.c_uint64],
        'kgs_build_join': [H, vp, vp, C.c_uint64, C.c_uint32],
        'kgs_probe': [H, C.c_uint32, vp],
        'kgs_join_probe_noalloc': [C.c_int, vp, vp, C.c_uint64, vp, vp, C.c_uint64, C.c_uint32, vp, vp],
        'kgs_stats_begin': [H, C.POINTER(WindowSpec)],
        'kgs_stats_accumulate': [H],
        'kgs_stats_finalize': [H],
        'kgs_window_stats': [H, C.POINTER(WindowSpec)],
        'kgs_set_run_log_capacity': [H, C.c_uint64],
        'kgs_job_counter': [H, C.c_int, u64p],
        'kgs_fetch_stats': [H, C.c_int, vp, vp, vp, vp, vp],
        'kgs_result_device_ptr': [H, C.c_int, C.c_int, C.POINTER(vp), C.POINTER(C.c_size_t)],
        'kgs_result_region': [H, C.POINTER(vp), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)],
        'kgs_set_packed_exchange': [H, C.c_int],
        'kgs_exchange_region': [H, C.POINTER(vp), C.POINTER(C.c_size_t), C.POINTER(C.c_int)],
        'kgs_packed_offset': [H, C.c_int, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)],
        'kgs_result_offset': [H, C.c_int, C.c_int, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)],
        'kgs_fetch_region': [H, vp, C.c_size_t],
        'kgs_exchange_setup': [H, C.c_int, C.c_int, vp, vp, C.c_size_t],
        'kgs_exchange_set_multicast': [H, vp],
        'kgs_exchange_set_self_copy': [H, C.c_int],
        'kgs_exchange_push': [H, C.POINTER(vp), C.POINTER(C.c_size_t)],
        'kgs_exchange_join': [H],
        'kgs_exchange_sync': [H],
        'kgs_nccl_unique_id': [vp],
        'kgs_nccl_init': [H, vp, C.c_int, C.c_int],
        'kgs_exchange_nccl': [H, vp, vp, C.c_size_t, C.POINTER(C.c_size_t)],
        'kgs_segmented_mean_f64': [H, vp, vp, C.c_uint64, C.c_double, C.c_double, vp],
        'kgs_matrix_parse': [H, vp, C.c_uint64, C.c_int, u64p, u64p],
        'kgs_matrix_fetch': [H, vp, vp, vp, vp],
        'kgs_matrix_mean': [H, C.c_double, C.c_double, vp],
        'kgs_launch_count': [H, u64p],
        'kgs_set_profiling': [H, C.c_int],
        'kgs_kernel_time': [H, C.c_int, f64p, u64p],
        'kgs_reset_kernel_times': [H],
        'kgs_stream': [H, C.c_int, C.POINTER(vp)],
        'kgs_mark': [H, C.c_int],
        'kgs_elapsed_ms': [H, f64p],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    lib.kgs_last_error.argtypes = [H]
    lib.kgs_last_error.restype = C.c_char_p
    lib.kgs_version.argtypes = []
    lib.kgs_version.restype = C.c_char_p
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = load()
    return _lib
