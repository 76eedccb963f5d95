"""CPU checks of the drop-in boundary: the shared library builds/loads and
exports exactly what include/dace_b200_reduce.h declares (no compute calls)."""
import ctypes
import os
import re

import dace_b200
from dace_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'dace_b200_reduce.h')


def _declared():
    src = open(HEADER).read()
    return sorted(set(re.findall(r'DACE_B200_API\s+[\w\s\*]+?\b(dace_b200_\w+)\s*\(', src)))


def test_header_declares_what_python_binds():
    assert _declared() == sorted(_lib.EXPORTED_SYMBOLS)


def test_library_loads_and_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), 'run __graft_entry__.build() first'
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), name
    lib = dace_b200.load()
    assert lib.dace_b200_version() == 2001
    assert lib.dace_b200_workspace_bytes() == 8 << 20
    assert lib.dace_b200_status_string(801).decode() == 'operation not supported'


def test_enum_values_match_reference_reductiontype():
    """dace::ReductionType numeric values (reference dace/runtime/include/dace/types.h:301-317)."""
    src = open(HEADER).read()
    want = {'CUSTOM': 0, 'MIN': 1, 'MAX': 2, 'SUM': 3, 'PRODUCT': 4, 'LOGICAL_AND': 5, 'BITWISE_AND': 6,
            'LOGICAL_OR': 7, 'BITWISE_OR': 8, 'LOGICAL_XOR': 9, 'BITWISE_XOR': 10}
    for k, v in want.items():
        assert re.search(r'DACE_B200_%s = %d\b' % (k, v), src), k
    from dace_b200 import ReductionType
    assert [ReductionType[n].value for n in ('Min', 'Max', 'Sum', 'Product', 'Logical_And', 'Bitwise_And',
                                             'Logical_Or', 'Bitwise_Or', 'Logical_Xor', 'Bitwise_Xor')] == list(range(1, 11))


def test_no_product_code_touches_the_oracle():
    """The oracle is test infrastructure: nothing under dace_b200/ may reference it."""
    for dirpath, _, files in os.walk(os.path.join(ROOT, 'dace_b200')):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f), errors='ignore').read()
                assert 'oracle' not in text.lower(), os.path.join(dirpath, f)


def test_compute_paths_neither_allocate_nor_synchronise_nor_read_the_environment():
    """Contract of include/dace_b200_reduce.h: the kernel files and the planner contain no cudaMalloc / cudaFree /
    synchronise / getenv.  Allowed exceptions: the peer-buffer setup entry points (dace_b200_peer_alloc / _free,
    not compute calls) and the one-time read of the tuning knobs when the library is loaded (read_tuning)."""
    csrc = os.path.join(ROOT, 'dace_b200', 'csrc')
    banned = re.compile(r'\b(cudaMalloc\w*|cudaFree\w*|cudaEventSynchronize|cudaStreamSynchronize|cudaDeviceSynchronize|'
                        r'cudaMemcpy|cudaMemset|getenv)\s*\(')
    for f in sorted(os.listdir(csrc)):
        if not f.endswith(('.cu', '.cuh', '.h')) or f == 'comm.cu':     # comm.cu = communicator set-up (NCCL, IPC)
            continue
        text = open(os.path.join(csrc, f)).read()
        text = re.sub(r'//[^\n]*', '', text)
        if f == 'api.cu':
            a, b = text.index('static Tuning read_tuning()'), text.index('static const Tuning g_tuning')
            text = text[:a] + text[b:]
            a, b = text.index('int dace_b200_peer_alloc('), text.index('int dace_b200_peer_export(')
            text = text[:a] + text[b:]
        hits = banned.findall(text)
        assert not hits, (f, hits)
    # the shared library's compute entry points carry their scratch in the signature
    src = open(HEADER).read()
    for fn in ('dace_b200_reduce_nd', 'dace_b200_reduce_full', 'dace_b200_reduce_rows', 'dace_b200_reduce_cols',
               'dace_b200_reduce_arg', 'dace_b200_wcr_scatter'):
        proto = re.search(r'DACE_B200_API int %s\((.*?)\);' % fn, src, re.S).group(1)
        assert 'void* workspace' in proto and 'size_t workspace_bytes' in proto, fn


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, '_lib', None)
    monkeypatch.setattr(_lib, 'LIB_PATH', str(tmp_path / 'nope.so'))
    try:
        _lib.load()
    except _lib.B200LibraryError as e:
        assert 'no CPU fallback' in str(e)
    else:
        raise AssertionError('load() must raise when the CUDA library is missing')


def test_single_process_communicator_needs_neither_nccl_nor_a_gpu():
    """world == 1: dace_b200_comm_* is pure host logic (what a generated program gets when it is not launched
    under a multi-rank launcher): create / query / no-op all-reduce / default / destroy."""
    import ctypes
    import os
    from dace_b200 import _lib
    lib = _lib.load()
    c = ctypes.c_void_p()
    assert lib.dace_b200_comm_create(0, 1, None, ctypes.byref(c)) == 0 and c.value
    assert lib.dace_b200_comm_rank(c) == 0 and lib.dace_b200_comm_world(c) == 1
    assert lib.dace_b200_comm_uses_peer_memory(c) == 0
    buf = (ctypes.c_float * 4)(1, 2, 3, 4)
    assert lib.dace_b200_comm_allreduce(c, 3, 0, buf, 4, None) == 0 and list(buf) == [1, 2, 3, 4]
    assert lib.dace_b200_comm_allreduce(c, 3, 99, buf, 4, None) == 801          # unknown dtype: cudaErrorNotSupported
    assert lib.dace_b200_comm_create(2, 2, None, ctypes.byref(ctypes.c_void_p())) == 1   # rank out of range
    assert lib.dace_b200_comm_destroy(c) == 0
    saved = {k: os.environ.pop(k, None) for k in ('WORLD_SIZE', 'RANK')}
    try:
        d = ctypes.c_void_p()
        assert lib.dace_b200_comm_default(ctypes.byref(d)) == 0 and lib.dace_b200_comm_world(d) == 1
        d2 = ctypes.c_void_p()
        assert lib.dace_b200_comm_default(ctypes.byref(d2)) == 0 and d2.value == d.value     # process-wide
        assert lib.dace_b200_comm_destroy(d) == 0
    finally:
        for k, v in saved.items():
            if v is not None:
                os.environ[k] = v


def test_planner_has_multi_pass_plans_for_views_without_a_gpu():
    """The dry run behind dace_b200_reduce_nd_workspace_bytes walks the same planner as a real call and needs no GPU:
    layouts that once fell to one warp of the strided kernel (padded windows folded completely or over interleaved
    axes) must report scratch beyond the base workspace, i.e. a multi-pass streaming plan exists for them; layouts
    with a direct kernel must not."""
    import ctypes
    lib = _lib.load()
    base = lib.dace_b200_workspace_bytes()

    def need(shape, istr, mask, dt=0):
        n = len(shape)
        return lib.dace_b200_reduce_nd_workspace_bytes(dt, dt, n, (ctypes.c_int64 * n)(*shape), (ctypes.c_int64 * n)(*istr), mask)

    pitch2, pitch3 = [16384, 1], [512 * 512, 512, 1]
    assert need([16384, 16000], pitch2, 0b11) == base + 16384 * 4                       # rows -> mid[R1] -> full
    assert need([1024, 500, 500], pitch3, 0b111) > base + 1024 * 500 * 4 - 1            # rows on two levels -> mid[A, B]
    assert need([1024, 500, 500], pitch3, 0b101) > base + 1024 * 500 * 4 - 1            # axes (0, 2)
    assert need([1024, 500, 500], pitch3, 0b110) > base                                 # axes (1, 2)
    assert need([32, 32, 510, 510], [32 * 512 * 512, 512 * 512, 512, 1], 0b1101) > base   # [N, C, H, W] interior, (0, 2, 3)
    assert need([4096, 512, 8], [4096, 8, 1], 0b101) == base + 256 * ((512 * 8 * 4 + 255) // 256)   # outer axis first
    # direct kernels: nothing beyond the base
    assert need([4096, 4096], [4096, 1], 0b11) == base
    assert need([1024, 500, 500], pitch3, 0b100) == base                                # rows on two levels, no scratch
    assert need([1024, 500, 500], pitch3, 0b001) == base                                # column fold with an outer axis


def test_planner_trace_keeps_views_off_the_strided_kernel():
    """dace_b200_reduce_nd_plan names the kernels a call would launch (no GPU needed).  Over contiguous arrays,
    windows padded on one to three levels and transposed views, every axis mask: the strided thread-per-output
    kernel ('generic') may only serve the layouts listed here -- folds over outer axes of a window (thread per
    output is a streaming plan there: measured 5.5-5.8 TB/s, profiles/r2_window_sweep.txt) and three transposed
    views whose output order differs from their input order."""
    import ctypes
    lib = _lib.load()

    def plan(shape, istr, mask, dt=0):
        n = len(shape)
        buf = ctypes.create_string_buffer(256)
        need = lib.dace_b200_reduce_nd_plan(dt, dt, n, (ctypes.c_int64 * n)(*shape), (ctypes.c_int64 * n)(*istr), mask, buf, 256)
        assert need == lib.dace_b200_reduce_nd_workspace_bytes(dt, dt, n, (ctypes.c_int64 * n)(*shape), (ctypes.c_int64 * n)(*istr), mask)
        return buf.value.decode()

    def dense(full):
        st = [1] * len(full)
        for i in range(len(full) - 2, -1, -1):
            st[i] = st[i + 1] * full[i + 1]
        return st

    cases = []
    for shape in ([4096, 4096], [512, 512, 1024], [64, 64, 64, 1024], [1 << 20, 8], [8, 1 << 20], [100000, 33], [64, 256, 16, 512]):
        cases += [('contig', shape, dense(shape), m) for m in range(1, 1 << len(shape))]
    for full, win in (([1024, 512, 512], [1024, 500, 500]), ([32, 32, 512, 512], [32, 32, 510, 510]),
                      ([4096, 4096], [4000, 4000]), ([16, 64, 64, 4096], [16, 60, 60, 4000])):
        cases += [('window', win, dense(full), m) for m in range(1, 1 << len(win))]
    for shape, perm in (([512, 512, 1024], (2, 0, 1)), ([512, 512, 1024], (1, 0, 2)), ([4096, 4096], (1, 0)), ([64, 64, 64, 1024], (0, 2, 1, 3))):
        st = dense(shape)
        cases += [('permuted', [shape[q] for q in perm], [st[q] for q in perm], m) for m in range(1, 1 << len(shape))]

    allowed = {
        ('window', (1024, 500, 500), 0b1), ('window', (32, 32, 510, 510), 0b1), ('window', (32, 32, 510, 510), 0b10),
        ('window', (32, 32, 510, 510), 0b11), ('window', (16, 60, 60, 4000), 0b1), ('window', (16, 60, 60, 4000), 0b10),
        ('permuted', (1024, 512, 512), 0b10), ('permuted', (64, 64, 64, 1024), 0b1), ('permuted', (64, 64, 64, 1024), 0b1000),
        ('permuted', (64, 64, 64, 1024), 0b1001),
    }
    generic = {(k, tuple(sh), m) for k, sh, st, m in cases if 'generic' in plan(sh, st, m)}
    assert generic <= allowed, sorted(generic - allowed)

    # spot checks of the trace itself
    assert plan([4096, 4096], [4096, 1], 0b10) == 'rows'
    assert plan([4096, 4096], [4096, 1], 0b01) == 'cols'
    assert plan([4096, 4096], [4096, 1], 0b11) == 'split'
    assert plan([1024, 500, 500], [262144, 512, 1], 0b100) == 'rows2'
    assert plan([16, 60, 60, 4000], [16777216, 262144, 4096, 1], 0b1111).startswith('peel: inner-first:')
    assert plan([16, 60, 60, 4000], [16777216, 262144, 4096, 1], 0b1000) == 'peel: rows2'
    assert plan([16, 60, 60, 4000], [16777216, 262144, 4096, 1], 0b0011) == 'peel: cols'       # slices [R, K1, K2] of a folded axis
    assert plan([16, 60, 60, 1000], [16777216, 262144, 4096, 1], 0b1110).startswith('peel:')   # 16 outputs: never a warp each
    # 16-bit in and out (dtype codes of include/dace_b200_reduce.h): fp32 between the passes, one small launch rounds
    import torch
    from dace_b200 import api
    BF16, F16 = api.dtype_code(torch.bfloat16), api.dtype_code(torch.float16)
    for dt in (BF16, F16):
        assert plan([1024, 500, 500], [262144, 512, 1], 0b111, dt) == 'inner-first: rows2 split finish'
        assert plan([1024, 500, 500], [262144, 512, 1], 0b101, dt) == 'inner-first: rows2 cols finish'
        assert plan([16384, 16000], [16384, 1], 0b11, dt) == 'inner-first: rows split finish'
        assert plan([512, 512, 1024], [524288, 1024, 1], 0b101, dt) == 'two-pass(rows,cols)'
