"""CPU: the np.argmax / np.argmin rewrite (dace_b200/arg_node.py) against the real frontend output.
Needs /root/reference; skipped on the GPU box."""
import os

import numpy as np
import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def dace(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache_arg')
    cwd = os.getcwd()
    os.chdir(d)
    mod = import_reference(str(d / '.dacecache'))
    yield mod
    os.chdir(cwd)


def _sdfg(dace, func, dt, axis, ndim=2):
    syms = [dace.symbol(s, dace.int64) for s in 'NMK'][:ndim]
    kept = [s for i, s in enumerate(syms) if i != axis]
    fn = np.argmax if func == 'max' else np.argmin

    @dace.program
    def prog(A: dt[tuple(syms)], out: dace.int32[tuple(kept)]):
        out[:] = fn(A, axis=axis)

    sdfg = prog.to_sdfg(simplify=True)
    sdfg.name = 'arg%s_%s_axis%d' % (func, dt.to_string(), axis)
    return sdfg


@pytest.mark.parametrize('func,axis,ndim', [('max', 1, 2), ('min', 0, 2), ('min', 1, 3)])
def test_frontend_map_is_rewritten_and_calls_the_c_abi(dace, func, axis, ndim):
    from dace_b200 import arg_node as AN
    sdfg = _sdfg(dace, func, dace.float64, axis, ndim)
    sdfg.apply_gpu_transformations()
    assert AN.use_b200_argreduce(sdfg) == 1
    nodes = [n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, AN.ArgReduce)]
    assert len(nodes) == 1 and nodes[0].func == func and nodes[0].axis == axis
    assert not any('_reduce_' in getattr(n, 'label', '') and 'arg' in n.label for n, _ in sdfg.all_nodes_recursive()
                   if not isinstance(n, AN.ArgReduce))
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    red = 12 if func == 'max' else 11
    # struct {int idx; double val}: idx at 0, val at 8, 16 bytes per pair; init == nullptr (fold into the init state's pairs)
    assert 'dace_b200_reduce_arg(%d, 1, 2, _in, (char *)_out + 0, (char *)_out + 8,' % red in code
    assert '(size_t)16, (size_t)16, nullptr' in code
    assert 'wcr_custom' not in code            # the CAS-loop kernel is gone
    assert '_val_and_idx' in code and '.idx' in code   # init state and .idx extraction stay as the frontend generated them


def test_struct_layouts(dace):
    from dace_b200 import arg_node as AN
    f32 = dace.struct('_val_and_idx', idx=dace.int32, val=dace.float32)
    f64 = dace.struct('_val_and_idx', idx=dace.int32, val=dace.float64)
    i64 = dace.struct('_val_and_idx', idx=dace.int32, val=dace.int64)
    assert AN._struct_layout(f32) == (0, 4, 8) and AN._struct_layout(f64) == (0, 8, 16) and AN._struct_layout(i64) == (0, 8, 16)


def test_other_maps_are_left_alone_and_host_arrays_rejected(dace):
    from dace_b200 import arg_node as AN, auto as AUTO
    N = dace.symbol('N', dace.int64)

    @dace.program
    def plain(A: dace.float32[N], B: dace.float32[N]):
        for i in dace.map[0:N]:
            B[i] = A[i] * 2

    assert AN.use_b200_argreduce(plain.to_sdfg(simplify=True)) == 0
    host = _sdfg(dace, 'max', dace.float32, 1)
    assert AUTO.apply_b200(host)['arg'] == 0           # host storage: auto-selection does not touch it
    assert AN.use_b200_argreduce(host) == 1            # explicit request ...
    with pytest.raises(ValueError, match='no CPU fallback'):
        host.expand_library_nodes()                    # ... fails loudly instead of falling back


def test_flat_argmax_is_kept_on_the_data_flow_side(dace):
    """np.argmax(A) over all elements: the frontend reads `pair[0].idx` on an interstate edge, which the reference's
    own GPU transformation rejects; lower_flat_arg_results replaces the symbol by a memlet, after which the program
    moves to the GPU and the 1-D arg-reduce map becomes an ArgReduce node."""
    from dace.sdfg.validation import InvalidSDFGInterstateEdgeError
    from dace_b200 import arg_node as AN
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def flat(A: dace.float32[N, M], out: dace.int64[1]):
        out[0] = np.argmax(A)

    rejected = flat.to_sdfg(simplify=True)
    with pytest.raises(InvalidSDFGInterstateEdgeError):
        rejected.apply_gpu_transformations()               # the reference as it is
    sdfg = flat.to_sdfg(simplify=True)
    assert AN.lower_flat_arg_results(sdfg) == 1
    assert AN.lower_flat_arg_results(sdfg) == 0
    assert not any(e.data.assignments for e in sdfg.all_interstate_edges())
    sdfg.validate()
    sdfg.apply_gpu_transformations()
    assert AN.use_b200_argreduce(sdfg) == 1
    node = next(n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, AN.ArgReduce))
    assert node.func == 'max' and node.axis == 0           # axis 0 of the flat view
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    assert 'dace_b200_reduce_arg(12, 0, 2, _in,' in code and '__pair.idx' in code and 'wcr_custom' not in code


def test_flat_arg_lowering_leaves_other_interstate_reads_alone(dace):
    from dace_b200 import arg_node as AN
    N = dace.symbol('N', dace.int64)

    @dace.program
    def branchy(A: dace.float32[N], out: dace.int64[1]):
        k = np.argmax(A)
        if k > 3:                                          # the symbol feeds a condition: not the frontend's store shape
            out[0] = k
        else:
            out[0] = 0

    sdfg = branchy.to_sdfg(simplify=True)
    before = sum(len(e.data.assignments) for e in sdfg.all_interstate_edges())
    assert AN.lower_flat_arg_results(sdfg) == 0
    assert sum(len(e.data.assignments) for e in sdfg.all_interstate_edges()) == before
