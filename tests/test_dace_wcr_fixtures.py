"""which of the reference's own WCR fixtures the backend rewrites and which it deliberately leaves on the
reference's per-element atomics - stated program by program, so that the boundary is a decision and not an accident.

Each fixture is built from the reference's test / sample source (imported from /root/reference where it is a
@dace.program, rebuilt with explicit connectors where the original uses connector-less edges that the reference's own
GPU transformation rejects), moved to the GPU with ``apply_gpu_transformations`` and passed through
``dace_b200.auto.apply_b200``.  The generated code of every result must still come out of the code generator."""
import importlib.util
import os

import pytest

from oracle.refenv import REFERENCE_ROOT, import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='needs the reference checkout')


def _load(rel, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_ROOT, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _census(dace, sdfg):
    from dace.libraries.standard.nodes.reduce import Reduce
    from dace_b200.scatter_node import ScatterReduce
    out = {'scatter': 0, 'b200_reduce': 0, 'other_reduce': 0, 'fused_kernels': 0, 'wcr_edges': 0}
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, ScatterReduce):
            out['scatter'] += 1
        elif isinstance(n, Reduce):
            out['b200_reduce' if n.implementation == 'B200' else 'other_reduce'] += 1
        elif isinstance(n, dace.nodes.Tasklet) and n.label.startswith('fused_map_reduce_b200'):
            out['fused_kernels'] += 1
    out['wcr_edges'] = sum(1 for e, _ in sdfg.all_edges_recursive() if getattr(e.data, 'wcr', None) is not None)
    return out


def _through(dace, sdfg):
    from dace_b200 import auto
    sdfg.apply_gpu_transformations()
    counts = auto.apply_b200(sdfg)
    sdfg.validate()
    sdfg.generate_code()
    return counts, _census(dace, sdfg)


def _multiple_cr(dace):
    N = dace.symbol('N')
    sdfg = dace.SDFG('multiple_cr')
    sdfg.add_array('A', [N], dace.int64)
    sdfg.add_array('s', [1], dace.int64)
    sdfg.add_array('p', [1], dace.int64)
    state = sdfg.add_state()
    me, mx = state.add_map('mymap', dict(i='0:N'))
    a = state.add_read('A')
    tasklets = []
    for name, arr, wcr in (('task1', 's', 'lambda a, b: a + b'), ('task2', 'p', 'lambda a, b: a * b')):
        t = state.add_tasklet(name, {'a'}, {'b'}, 'b = a')
        state.add_memlet_path(a, me, t, dst_conn='a', memlet=dace.Memlet('A[i]'))
        state.add_memlet_path(t, mx, state.add_write(arr), src_conn='b', memlet=dace.Memlet('%s[0]' % arr, wcr=wcr))
        tasklets.append(t)
    return sdfg, state, tasklets


def test_multiple_cr_two_wcr_outputs_of_one_map_are_split_and_fused():
    """reference tests/multiple_cr_test.py: one map, two tasklets, `s += a` and `p *= a` into scalars.  The outputs come
    from disjoint tasklets, so the map is split into one map per output (dace_b200/split_wcr.py) and each becomes a
    fused tree reduction: two streaming passes instead of two same-address atomics per element."""
    dace = import_reference()
    sdfg, _, _ = _multiple_cr(dace)
    counts, census = _through(dace, sdfg)
    assert counts['split'] == 1 and counts['wcr_fused'] == 2 and counts['scatter'] == 0
    assert census['wcr_edges'] == 0 and census['fused_kernels'] == 2


def test_wcr_outputs_that_share_a_tasklet_stay_on_reference_atomics():
    """The same two outputs fed by ONE tasklet (or through a shared intermediate): splitting would duplicate work the
    program wrote once, and the fused rewrite takes maps with one WCR output - left alone, stated here."""
    dace = import_reference()
    N = dace.symbol('N')
    sdfg = dace.SDFG('multiple_cr_shared')
    sdfg.add_array('A', [N], dace.int64)
    sdfg.add_array('s', [1], dace.int64)
    sdfg.add_array('p', [1], dace.int64)
    state = sdfg.add_state()
    me, mx = state.add_map('mymap', dict(i='0:N'))
    t = state.add_tasklet('both', {'a'}, {'b', 'c'}, 'b = a\nc = a')
    state.add_memlet_path(state.add_read('A'), me, t, dst_conn='a', memlet=dace.Memlet('A[i]'))
    state.add_memlet_path(t, mx, state.add_write('s'), src_conn='b', memlet=dace.Memlet('s[0]', wcr='lambda a, b: a + b'))
    state.add_memlet_path(t, mx, state.add_write('p'), src_conn='c', memlet=dace.Memlet('p[0]', wcr='lambda a, b: a * b'))
    counts, census = _through(dace, sdfg)
    assert counts['split'] == 0 and counts['wcr_fused'] == 0 and census['wcr_edges'] >= 2


def test_split_is_undone_when_a_piece_would_not_be_lowered():
    """One of the two outputs uses an operator the sm_100a kernels do not serve (`a - b`): the map is left whole."""
    dace = import_reference()
    sdfg, state, _ = _multiple_cr(dace)
    for e in state.edges():
        if e.data.wcr is not None and e.data.data == 'p':
            e.data.wcr = 'lambda a, b: a - b'
    before = len(state.nodes())
    counts, census = _through(dace, sdfg)
    assert counts['split'] == 0 and counts['wcr_fused'] == 0 and census['wcr_edges'] >= 2


def test_confres_reduce_nodes_move_the_mixed_wcr_map_stays():
    """reference tests/confres_test.py: a map with plain outputs AND two WCR outputs into fixed elements (left alone:
    not a pure reduction map), three associative Reduce nodes (-> 'B200'), and `lambda a, b: a - b` (left on the
    reference's `pure`: not associative, a tree over it is a different number)."""
    dace = import_reference()
    mod = _load('tests/confres_test.py', 'ref_confres_test')
    sdfg = mod.confres_test.to_sdfg(simplify=True)
    counts, census = _through(dace, sdfg)
    assert census['b200_reduce'] == 3 and census['other_reduce'] == 1
    assert counts['scatter'] == 0 and counts['wcr_fused'] == 0 and census['wcr_edges'] >= 2


def test_mapreduce_histogram_the_reduce_node_moves():
    """reference tests/mapreduce_test.py:16-31,274-291 (uint8 keys -> uint32[256] through a [BINS, H, W] one-hot
    transient): the Reduce over axes (1, 2) is taken by 'B200' (rows form after merging); the one-hot map writes each
    element once (no WCR) and stays a generated kernel."""
    dace = import_reference()
    from dace.transformation.dataflow import MapReduceFusion
    mod = _load('tests/mapreduce_test.py', 'ref_mapreduce_test')
    sdfg = mod.histogram.to_sdfg()
    sdfg.simplify()
    sdfg.apply_transformations(MapReduceFusion)
    counts, census = _through(dace, sdfg)
    assert census['b200_reduce'] + census['fused_kernels'] + census['scatter'] >= 1 and census['other_reduce'] == 0


def test_explicit_histogram_sample_computed_index_stays():
    """reference samples/explicit/histogram.py: `out[min(int(a * BINS), BINS - 1)] = 1` - the bin index is COMPUTED in
    the tasklet from a float.  Left alone: ScatterReduce takes index ARRAYS (`bins[keys[i]] += w[i]`, explicit
    `out[k] = v`); materialising the computed indices first would add a pass over the data."""
    dace = import_reference()
    mod = _load('samples/explicit/histogram.py', 'ref_explicit_histogram')
    sdfg = mod.histogram.to_sdfg(simplify=True)
    counts, census = _through(dace, sdfg)
    assert counts['scatter'] == 0 and census['wcr_edges'] >= 1


def test_cr_complex_values_stay():
    """reference tests/cr_complex_test.py: complex128 sum through a WCR map.  Left alone: complex types have no
    sm_100a kernel here (the reference resolves them with a non-atomic read-modify-write / critical section)."""
    dace = import_reference()

    @dace.program
    def cr_complex(inp: dace.complex128[12], output: dace.complex128[1]):
        @dace.map(_[0:12])
        def tasklet(i):
            a << inp[i]
            b >> output(1, lambda a, b: a + b)
            b = a

    counts, census = _through(dace, cr_complex.to_sdfg(simplify=True))
    assert sum(counts.values()) == 0 and census['wcr_edges'] >= 1


def test_what_is_rewritten_for_contrast():
    """The shapes that ARE taken (details in tests/test_dace_scatter_node.py / test_dace_fused_expansion.py): indirect
    `bins[keys[i]] += w[i]` -> ScatterReduce, and `s[i] += f(A[i, j])` -> fused tree reduction."""
    dace = import_reference()
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def weighted(keys: dace.int32[N], w: dace.float32[N], bins: dace.float32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += w[i]

    @dace.program
    def rowsq(A: dace.float32[N, M], s: dace.float32[N]):
        for i, j in dace.map[0:N, 0:M]:
            s[i] += A[i, j] * A[i, j]

    c1, s1 = _through(dace, weighted.to_sdfg(simplify=True))
    assert c1['scatter'] == 1 and s1['scatter'] == 1        # (the node's own output memlet keeps the WCR until expansion)
    c2, s2 = _through(dace, rowsq.to_sdfg(simplify=True))
    assert c2['wcr_fused'] == 1 and s2['fused_kernels'] == 1
