"""CPU: FuseRowMaxSum (dace_b200/fuse_node.py) on SDFGs produced by the real frontend.
Needs /root/reference; skipped on the GPU box."""
import os

import numpy as np
import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def dace(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache_fuse')
    cwd = os.getcwd()
    os.chdir(d)
    mod = import_reference(str(d / '.dacecache'))
    yield mod
    os.chdir(cwd)


def _stats(dace, name, dt=None, sum_axis=1, gpu=True):
    dt = dt or dace.float32
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def prog(A: dt[N, M], mx_out: dt[N], sm_out: dt[N if sum_axis == 1 else M]):
        mx_out[:] = np.max(A, axis=1)
        sm_out[:] = np.sum(A, axis=sum_axis)

    sdfg = prog.to_sdfg(simplify=True)
    sdfg.name = name
    if gpu:
        sdfg.apply_gpu_transformations()
    return sdfg


def test_pair_is_fused_into_one_call(dace):
    from dace_b200 import auto as AUTO, fuse_node as FN
    from dace.libraries.standard.nodes.reduce import Reduce
    auto = _stats(dace, 'fuse_auto')
    assert AUTO.apply_b200(auto)['max_sum'] == 1          # whole-program selection includes the rewrite
    sdfg = _stats(dace, 'fuse_cg')
    AUTO.apply_b200(sdfg, fuse=False)
    assert FN.fuse_row_max_sum(sdfg) == 1
    assert not [n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, Reduce)]
    assert len([n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, FN.RowsMaxSum)]) == 1
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    assert code.count('dace_b200_reduce_rows_max_sum(0, _in, _max, _sum, (size_t)(N), (size_t)(M),') == 1
    assert 'dace_b200_reduce_nd(' not in code
    assert FN.fuse_row_max_sum(sdfg) == 0


def test_pairs_that_must_not_be_fused(dace):
    from dace_b200 import fuse_node as FN
    assert FN.fuse_row_max_sum(_stats(dace, 'fuse_axis0', sum_axis=0)) == 0      # different axes
    assert FN.fuse_row_max_sum(_stats(dace, 'fuse_f64', dt=dace.float64)) == 0   # kernel is f32 / bf16 / f16 only
    assert FN.fuse_row_max_sum(_stats(dace, 'fuse_host', gpu=False)) == 0        # host storage: no CPU fallback, no rewrite
    # a max identity that is not "minus infinity": semantics would change
    sdfg = _stats(dace, 'fuse_ident')
    from dace.libraries.standard.nodes.reduce import Reduce
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, Reduce) and 'max' in n.wcr:
            n.identity = 0
    assert FN.fuse_row_max_sum(sdfg) == 0
