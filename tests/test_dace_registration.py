"""CPU: the 'B200' expansion plugs into the REAL dace.library registry and emits
the C-ABI call (codegen-text test in the style of reference
tests/library/reduce_cub_workspace_test.py:27-84).  Needs /root/reference; skipped
on the GPU box where it does not exist."""
import os

import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def dace(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache')
    cwd = os.getcwd()
    os.chdir(d)
    mod = import_reference(str(d / '.dacecache'))
    import dace_b200.expansion  # noqa: F401  (registers the expansion)
    yield mod
    os.chdir(cwd)


def _sdfg(dace, name, axes, wcr='lambda a, b: a + b', identity=0, gpu=True):
    from oracle.build_ref import _reduce_sdfg
    sdfg = _reduce_sdfg(dace, name, dace.float32, dace.float32, 3, axes, wcr, identity, 'B200')
    if gpu:
        sdfg.apply_gpu_transformations()
        from dace_b200.expansion import use_b200_reductions
        use_b200_reductions(sdfg)
    return sdfg


def test_registered_in_reduce_implementations(dace):
    from dace.libraries.standard.nodes.reduce import Reduce
    from dace_b200.expansion import ExpandReduceB200, B200Reduce
    assert Reduce.implementations['B200'] is ExpandReduceB200
    from dace_b200.expansion import B200FusedReduce
    assert ExpandReduceB200.environments == [B200Reduce, B200FusedReduce]   # C ABI + the header for custom-lambda functors
    assert set(Reduce.implementations) >= {'pure', 'OpenMP', 'CUDA (device)', 'GPUAuto', 'B200'}
    assert Reduce.default_implementation == 'pure'   # importing the plug-in does not hijack the default


def test_generated_code_calls_the_c_abi(dace):
    sdfg = _sdfg(dace, 'b200_codegen_mid', [1])
    code = {c.title: c.clean_code for c in sdfg.generate_code()}
    frame = next(v for k, v in code.items() if 'Frame' in k or k == 'b200_codegen_mid')
    assert 'dace_b200_reduce.h' in frame
    assert frame.count('dace_b200_reduce_nd(') == 1
    assert 'DACE_GPU_CHECK((cudaError_t)dace_b200_reduce_nd(3, 0, 0, _in, _out, 3,' in frame
    assert '__b200_shape[] = {(int64_t)(D0), (int64_t)(D1), (int64_t)(D2)}' in frame
    import re
    istr = re.search(r'__b200_istr\[\] = \{(.*?)\};', frame).group(1).replace(' ', '')
    assert istr in ('(int64_t)(D1*D2),(int64_t)(D2),(int64_t)(1)', '(int64_t)((D1*D2)),(int64_t)(D2),(int64_t)(1)')
    import re
    m = re.search(r', 2u, &__b200_identity, (__b200_ws_\w+), \1_bytes,', frame)
    assert m, 'the call passes the workspace owned by this node'
    ws = m.group(1)
    assert '__dace_current_stream' in frame
    assert 'void *%s = nullptr; size_t %s_bytes = 0;' % (ws, ws) in frame        # per-node storage, like CUB's
    assert '__b200_need = dace_b200_reduce_nd_workspace_bytes(0, 0, 3, __b200_shape, __b200_istr, 2u);' in frame
    assert 'cudaMalloc(&%s, %s_bytes)' % (ws, ws) in frame                       # allocated once in __dace_init
    assert 'dace_b200_workspace_init(%s, %s_bytes, nullptr)' % (ws, ws) in frame
    assert 'cudaFree(%s)' % ws in frame                                          # freed in __dace_exit
    assert '__state->dace_b200_workspace' not in frame
    assert 'cub::' not in ''.join(code.values())


def test_identity_none_passes_null(dace):
    sdfg = _sdfg(dace, 'b200_codegen_noident', [0], 'lambda a, b: max(a, b)', None)
    frame = ''.join(c.clean_code for c in sdfg.generate_code())
    assert 'dace_b200_reduce_nd(2, 0, 0,' in frame and ', 1u, nullptr, __b200_ws_b200_codegen_noident_' in frame


def test_environment_links_the_prebuilt_library(dace):
    from dace.codegen import compiler
    from dace_b200 import _lib
    sdfg = _sdfg(dace, 'b200_codegen_env', [2])
    sdfg.expand_library_nodes()
    envs = set()
    for n, _ in sdfg.all_nodes_recursive():
        envs |= set(getattr(n, 'environments', []) or [])
    assert any('B200Reduce' in e for e in envs)
    flags, link_flags = compiler.get_environment_flags([dace.library.get_environment(e) for e in envs])
    joined = ' '.join(list(flags) + list(link_flags))
    assert _lib.LIB_PATH in joined and _lib.INCLUDE_DIR in joined and '-Wl,-rpath,' in joined


def test_no_cpu_fallback_and_custom_rejected(dace):
    cpu = _sdfg(dace, 'b200_codegen_cpu', [1], gpu=False)
    with pytest.raises(ValueError, match='no CPU fallback|GPU_Global'):
        cpu.expand_library_nodes()
    custom = _sdfg(dace, 'b200_codegen_custom', [1], 'lambda a, b: a * a + b')
    with pytest.raises(NotImplementedError):
        custom.expand_library_nodes()
    ok = _sdfg(dace, 'b200_codegen_custommin', [1], 'lambda a, b: a if a < b else b', 1e30)
    assert 'dace_b200_reduce_nd(1, 0, 0,' in ''.join(c.clean_code for c in ok.generate_code())


# ---- in-kernel expansions (dace_b200/block_expansion.py) ----

def _load_build_programs():
    import importlib.util
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'dace_e2e', 'build_programs.py')
    spec = importlib.util.spec_from_file_location('dace_e2e_build_programs_for_tests', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_block_expansions_registered(dace):
    from dace.libraries.standard.nodes.reduce import Reduce
    from dace_b200 import block_expansion as BE
    assert Reduce.implementations['B200 (block)'] is BE.ExpandReduceB200Block
    assert Reduce.implementations['B200 (block allreduce)'] is BE.ExpandReduceB200BlockAll
    assert BE.B200BlockReduce.headers['cuda'] == ['dace_b200_block_reduce.cuh']
    assert BE.B200BlockReduce.cmake_libraries == []      # header-only: nothing to link


def test_block_expansion_generated_kernel_calls_the_header(dace):
    bp = _load_build_programs()
    from dace_b200 import block_expansion  # noqa: F401
    sdfg = bp._blockreduce_sdfg(dace, 'b200_cg_block_i32', 'lambda a, b: a + b', dace.int32, nblocks=3, nthreads=96, items=4)
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    assert '#include "dace_b200_block_reduce.cuh"' in code
    assert 'dace_b200::block_reduce<int, 96, ' in code
    assert 'reinterpret_cast<const int (&)[4]>(*_in), dace_b200::op<3, int>()' in code
    assert 'cub::' not in code
    custom = bp._blockreduce_sdfg(dace, 'b200_cg_block_custom', 'lambda a, b: a + b + a * b', dace.int64, 2, 64, 1)
    code = ''.join(c.clean_code for c in custom.generate_code())
    assert 'struct __b200_reduce_b200_cg_block_custom_' in code and 'dace_b200::block_reduce<int64_t, 64, ' in code \
        or 'dace_b200::block_reduce<long long, 64, ' in code


def test_block_expansion_errors_match_the_reference(dace):
    """Outside a kernel / without identity: same ValueErrors as reference reduce.py:665-716."""
    from oracle.build_ref import _reduce_sdfg
    from dace_b200 import block_expansion  # noqa: F401
    host = _reduce_sdfg(dace, 'b200_cg_block_host', dace.float32, dace.float32, 1, None, 'lambda a, b: a + b', 0,
                        'B200 (block)')
    with pytest.raises(ValueError, match='within a GPU kernel|register'):
        host.expand_library_nodes()
    bp = _load_build_programs()
    noident = bp._blockreduce_sdfg(dace, 'b200_cg_block_noident', 'lambda a, b: a + b', dace.float32, 2, 32, 2)
    for n, _ in noident.all_nodes_recursive():
        if isinstance(n, block_expansion.Reduce):
            n.identity = None
    with pytest.raises(ValueError, match='initial value'):
        noident.expand_library_nodes()
