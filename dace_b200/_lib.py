"""ctypes binding of ``libdace_b200_reduce.so`` (the C ABI in include/dace_b200_reduce.h).

The library is built in-tree (``dace_b200/lib/``) by ``__graft_entry__.build()`` /
``make -C dace_b200/csrc``.  There is deliberately no fallback: if the shared
library is missing, every entry point raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'lib', 'libdace_b200_reduce.so')
CSRC_DIR = os.path.join(_HERE, 'csrc')
INCLUDE_DIR = os.path.join(os.path.dirname(_HERE), 'include')

_lib = None

_c = ctypes
_PROTOTYPES = {
    'dace_b200_version': (_c.c_int, []),
    'dace_b200_status_string': (_c.c_char_p, [_c.c_int]),
    'dace_b200_workspace_bytes': (_c.c_size_t, []),
    'dace_b200_workspace_init': (_c.c_int, [_c.c_void_p, _c.c_size_t, _c.c_void_p]),
    'dace_b200_reduce_nd_workspace_bytes': (_c.c_size_t, [
        _c.c_int, _c.c_int, _c.c_int, _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64), _c.c_uint32
    ]),
    'dace_b200_reduce_nd_plan': (_c.c_size_t, [
        _c.c_int, _c.c_int, _c.c_int, _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64), _c.c_uint32, _c.c_char_p, _c.c_size_t
    ]),
    'dace_b200_reduce_nd': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int,
        _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64), _c.POINTER(_c.c_int64), _c.c_uint32, _c.c_void_p,
        _c.c_void_p, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_reduce_full': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p, _c.c_void_p,
        _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_reduce_rows': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_size_t, _c.c_size_t,
        _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_reduce_rows_noalloc': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_size_t, _c.c_size_t,
        _c.c_void_p, _c.c_void_p
    ]),
    'dace_b200_reduce_cols': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_size_t, _c.c_size_t,
        _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_reduce_rows_max_sum': (_c.c_int, [
        _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_size_t, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_reduce_arg': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_size_t, _c.c_size_t,
        _c.c_int64, _c.c_int64, _c.c_int64, _c.c_size_t, _c.c_size_t, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_wcr_scatter_workspace_bytes': (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_size_t,
                                                            _c.c_size_t]),
    'dace_b200_wcr_scatter': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p,
        _c.c_size_t, _c.c_void_p, _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_peer_buffer_bytes': (_c.c_size_t, [_c.c_int, _c.c_size_t]),
    'dace_b200_peer_alloc': (_c.c_int, [_c.c_size_t, _c.POINTER(_c.c_void_p)]),
    'dace_b200_peer_free': (_c.c_int, [_c.c_void_p]),
    'dace_b200_peer_export': (_c.c_int, [_c.c_void_p, _c.c_char_p]),
    'dace_b200_peer_open': (_c.c_int, [_c.c_char_p, _c.POINTER(_c.c_void_p)]),
    'dace_b200_peer_close': (_c.c_int, [_c.c_void_p]),
    'dace_b200_peer_allreduce': (_c.c_int, [
        _c.c_int, _c.c_int, _c.c_void_p, _c.c_size_t, _c.POINTER(_c.c_void_p), _c.c_int, _c.c_int, _c.c_uint64,
        _c.c_size_t, _c.c_void_p
    ]),
    'dace_b200_comm_unique_id': (_c.c_int, [_c.c_char_p]),
    'dace_b200_comm_create': (_c.c_int, [_c.c_int, _c.c_int, _c.c_char_p, _c.POINTER(_c.c_void_p)]),
    'dace_b200_comm_create_from_env': (_c.c_int, [_c.POINTER(_c.c_void_p)]),
    'dace_b200_comm_destroy': (_c.c_int, [_c.c_void_p]),
    'dace_b200_comm_rank': (_c.c_int, [_c.c_void_p]),
    'dace_b200_comm_world': (_c.c_int, [_c.c_void_p]),
    'dace_b200_comm_uses_peer_memory': (_c.c_int, [_c.c_void_p]),
    'dace_b200_comm_set_default': (_c.c_int, [_c.c_void_p]),
    'dace_b200_comm_default': (_c.c_int, [_c.POINTER(_c.c_void_p)]),
    'dace_b200_comm_allreduce': (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_PROTOTYPES)


class B200LibraryError(RuntimeError):
    """Raised when the CUDA library is missing or a call returns a CUDA error."""


def build(verbose: bool = False, jobs: int = 8) -> str:
    """Compile the kernels for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ['make', '-C', CSRC_DIR, '-j%d' % jobs]
    res = subprocess.run(cmd, capture_output=not verbose, text=True)
    if res.returncode != 0:
        raise B200LibraryError('building libdace_b200_reduce.so failed:\n%s\n%s' %
                               (res.stdout or '', res.stderr or ''))
    return LIB_PATH


def load():
    """Load the shared library (once) and attach prototypes.  No fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200LibraryError('%s not found: run `python -c "import __graft_entry__ as g; g.build()"` or '
                               '`make -C dace_b200/csrc`; there is no CPU fallback' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    for name, (res, args) in _PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status: int, what: str = 'dace_b200 call'):
    """Mirror of DACE_GPU_CHECK + __dace_gpu_last_error -> RuntimeError
    (reference dace/codegen/compiled_sdfg.py:583-592)."""
    if status != 0:
        msg = load().dace_b200_status_string(status)
        raise B200LibraryError('%s failed: %s (cudaError %d)' % (what, msg.decode() if msg else '?', status))
