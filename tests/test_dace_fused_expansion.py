"""CPU: FuseMapIntoReduce (dace_b200/fused_expansion.py) on SDFGs from the real frontend: what matches,
what the generated functor looks like, what is left alone.  Needs /root/reference; skipped on the GPU box."""
import os

import numpy as np
import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def dace(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache_fused')
    cwd = os.getcwd()
    os.chdir(d)
    mod = import_reference(str(d / '.dacecache'))
    yield mod
    os.chdir(cwd)


def _den(dace, name, axis=1, gpu=True, reuse_tmp=False):
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def prog(A: dace.float32[N, M], mx: dace.float32[N], s: dace.float32[N if axis == 1 else M], t2: dace.float32[N, M]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = np.exp(A[i, j] - mx[i])
        s[:] = np.sum(tmp, axis=axis)
        if reuse_tmp:
            t2[:] = tmp

    sdfg = prog.to_sdfg(simplify=True)
    sdfg.name = name
    if gpu:
        sdfg.apply_gpu_transformations()
    return sdfg


def test_map_body_becomes_a_load_functor(dace):
    from dace_b200 import fused_expansion as FE
    from dace.libraries.standard.nodes.reduce import Reduce
    sdfg = _den(dace, 'fused_cg')
    assert FE.fuse_maps_into_reductions(sdfg) == 1
    assert not [n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, Reduce)]
    assert 'tmp' not in sdfg.arrays                       # the transient is gone
    code = {c.title + c.language: c.clean_code for c in sdfg.generate_code()}
    cu = ''.join(v for k, v in code.items() if k.endswith('cu'))
    frame = ''.join(v for k, v in code.items() if k.endswith('cpp'))
    assert '#include "dace_b200_fused_reduce.cuh"' in cu
    assert 'struct __b200_load_fused_cg_' in cu and 'const float __in1 = gpu_A[__flat];' in cu
    assert 'const float __in2 = gpu_mx[i];' in cu and '__out = exp(__in1);' in cu
    assert 'dace_b200::fused_rows<3, float, float>(__load, __out, (ptrdiff_t)1, (size_t)(N), (size_t)(M), &__identity' in cu
    # vector interface: the operand read at the map's flat index comes through 32-byte accesses, the row-indexed one
    # stays a scalar read; the call site works out the widest legal access from n and the pointer
    assert 'static constexpr int kVec = 8;' in cu and 'using value_type = float;' in cu
    assert 'float __b200_v_gpu_A[__V]; dace_b200::stream_load<float, __V>(gpu_A + __flat0, __b200_v_gpu_A);' in cu
    assert 'const float __in1 = __b200_v_gpu_A[__k];' in cu and cu.count('const float __in2 = gpu_mx[i];') == 2
    assert '((size_t)gpu_A % (__v * sizeof(float))) == 0' in cu and '__workspace_bytes, __stream, __b200_max_vec);' in cu
    assert 'DACE_GPU_CHECK(__b200_fused_fused_cg_' in frame and ', __b200_ws_fused_cg_' in frame     # per-kernel workspace
    assert '__global__' not in cu.split('struct __b200_load_')[0] or 'reduce_atomic' not in cu   # no WCR map kernel
    assert FE.fuse_maps_into_reductions(sdfg) == 0


def test_things_that_must_not_be_fused(dace):
    from dace_b200 import fused_expansion as FE
    col = _den(dace, 'fused_axis0', axis=0)                                         # column fold: fused_cols
    assert FE.fuse_maps_into_reductions(col) == 1
    assert 'dace_b200::fused_cols<3, float, float>' in ''.join(c.clean_code for c in col.generate_code())
    N, M, K = (dace.symbol(x, dace.int64) for x in 'NMK')

    @dace.program
    def mid(A: dace.float32[N, M, K], s: dace.float32[N, K]):
        tmp = np.ndarray([N, M, K], dace.float32)
        for i, j, k in dace.map[0:N, 0:M, 0:K]:
            tmp[i, j, k] = A[i, j, k] * 2
        s[:] = np.sum(tmp, axis=1)

    sd = mid.to_sdfg(simplify=True)
    sd.apply_gpu_transformations()
    assert FE.fuse_maps_into_reductions(sd) == 0                                    # a middle axis: left to the 'B200' Reduce
    assert FE.fuse_maps_into_reductions(_den(dace, 'fused_host', gpu=False)) == 0   # host storage: untouched
    assert FE.fuse_maps_into_reductions(_den(dace, 'fused_reuse', reuse_tmp=True)) == 0   # the transient has another reader


def test_tasklet_code_may_use_map_parameters_and_symbols(dace):
    from dace_b200 import fused_expansion as FE
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def prog(A: dace.float64[N, M], s: dace.float64[N]):
        tmp = np.ndarray([N, M], dace.float64)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = A[i, j] * i / N + j
        s[:] = np.sum(tmp, axis=1)

    sdfg = prog.to_sdfg(simplify=True)
    sdfg.name = 'fused_params'
    sdfg.apply_gpu_transformations()
    if FE.fuse_maps_into_reductions(sdfg) == 0:
        pytest.skip('the frontend did not produce a plain tasklet body for this program')
    cu = ''.join(c.clean_code for c in sdfg.generate_code() if c.language == 'cu')
    assert 'const long long i = (long long)__row;' in cu and 'const long long j = (long long)__col;' in cu
    assert 'int64_t N;' in cu or 'long long N;' in cu


def test_wcr_maps_lower_to_the_fused_kernels(dace):
    """`s[i] += f(A[i, j])` (AugAssignToWCR) and explicit WCR outputs: rows / cols / all forms; indirect WCR maps
    (histograms) and maps that read their own output are not touched."""
    from dace_b200 import fused_expansion as FE
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def rows(A: dace.float32[N, M], s: dace.float32[N]):
        for i, j in dace.map[0:N, 0:M]:
            s[i] += A[i, j] * A[i, j]

    @dace.program
    def cols(A: dace.float64[N, M], s: dace.float64[M]):
        for i, j in dace.map[0:N, 0:M]:
            with dace.tasklet:
                xin << A[i, j]
                res >> s(1, lambda x, y: min(x, y))[j]
                res = xin + 1

    @dace.program
    def hist(keys: dace.int32[N], bins: dace.int32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    @dace.program
    def middle(A: dace.float32[N, M, N], s: dace.float32[N, N]):
        for i, j, k in dace.map[0:N, 0:M, 0:N]:
            s[i, k] += A[i, j, k]

    def lowered(p, name):
        sdfg = p.to_sdfg(simplify=True)
        sdfg.name = name
        sdfg.apply_gpu_transformations()
        return FE.lower_wcr_maps(sdfg), sdfg

    n, sd = lowered(rows, 'wcr_cg_rows')
    assert n == 1
    code = ''.join(c.clean_code for c in sd.generate_code())
    assert 'dace_b200::fused_rows<3, float, float>(__load, __out, (ptrdiff_t)1, (size_t)(N), (size_t)(M), (const float*)nullptr' in code
    assert 'reduce_atomic' not in code and 'atomicAdd' not in code
    n, sd = lowered(cols, 'wcr_cg_cols')
    assert n == 1 and 'dace_b200::fused_cols<1, double, double>' in ''.join(c.clean_code for c in sd.generate_code())
    assert lowered(hist, 'wcr_cg_hist')[0] == 0          # indirection: ScatterReduce's job
    assert lowered(middle, 'wcr_cg_middle')[0] == 0      # a middle axis folded: left to the code generator
