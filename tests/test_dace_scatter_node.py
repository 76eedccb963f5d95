"""CPU: the ScatterReduce library node and the WCR-map rewrite, against the REAL
reference (needs /root/reference; skipped on the GPU box).

* the rewrite matches the map shapes the Python frontend produces
  (reference samples/explicit/histogram.py:13-26, tests/npbench/misc/azimint_hist_test.py:39-59)
  and leaves other maps alone;
* with the 'pure' expansion the rewritten program is compiled by the reference
  for the CPU and must give the results of the untouched program (NumPy);
* with the 'B200' expansion the generated host code calls ``dace_b200_wcr_scatter``
  and refuses non-GPU storage (no CPU fallback).
"""
import os

import numpy as np
import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def env(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache_scatter')
    cwd = os.getcwd()
    os.chdir(d)
    dace = import_reference(str(d / '.dacecache'))
    from dace_b200 import scatter_node as SN
    N, M = dace.symbol('N'), dace.symbol('M')

    @dace.program
    def hist_aug(keys: dace.int32[N], bins: dace.int32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    @dace.program
    def hist_explicit(keys: dace.int32[N], bins: dace.int32[M]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = 1

    @dace.program
    def hist_weighted(keys: dace.int32[N], w: dace.float32[N], bins: dace.float32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += w[i]

    @dace.program
    def scat_max_shifted(keys: dace.int64[N], w: dace.float64[N], bins: dace.float64[M]):
        for i in dace.map[2:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i - 2]
                out >> bins(1, lambda x, y: max(x, y))[:]
                out[k] = v

    @dace.program
    def not_a_scatter(keys: dace.int32[N], w: dace.float32[N], bins: dace.float32[M]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = v * 2

    @dace.program
    def elementwise(a: dace.float32[N], b: dace.float32[N]):
        for i in dace.map[0:N]:
            b[i] = a[i] + 1

    progs = dict(hist_aug=hist_aug, hist_explicit=hist_explicit, hist_weighted=hist_weighted,
                 scat_max_shifted=scat_max_shifted, not_a_scatter=not_a_scatter, elementwise=elementwise)
    yield dace, SN, progs
    os.chdir(cwd)


def _rewritten(env, name, impl):
    dace, SN, progs = env
    sdfg = progs[name].to_sdfg(simplify=True)
    applied = sdfg.apply_transformations_repeated(SN.WCRMapToScatterReduce, options={'implementation': impl})
    nodes = [n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, SN.ScatterReduce)]
    return sdfg, applied, nodes


def test_library_registration(env):
    dace, SN, _ = env
    import dace.library
    assert set(SN.ScatterReduce.implementations) == {'pure', 'B200'}
    assert SN.ScatterReduce.default_implementation == 'pure'
    assert SN.ExpandScatterB200.environments[0].__name__ == 'B200Reduce'
    assert 'dace_b200' in dace.library._DACE_REGISTERED_LIBRARIES
    assert SN.WCRMapToScatterReduce in SN._dace_library_transformations


@pytest.mark.parametrize('name,const,has_val,red', [('hist_aug', 1, False, 'Sum'), ('hist_explicit', 1, False, 'Sum'),
                                                     ('hist_weighted', None, True, 'Sum'),
                                                     ('scat_max_shifted', None, True, 'Max')])
def test_rewrite_matches_frontend_shapes(env, name, const, has_val, red):
    dace, SN, _ = env
    sdfg, applied, nodes = _rewritten(env, name, 'pure')
    assert applied == 1 and len(nodes) == 1
    node = nodes[0]
    assert node.const_value == const and ('_val' in node.in_connectors) == has_val
    assert str(node) == 'ScatterReduce (%s)' % red
    assert not any(isinstance(n, dace.nodes.MapEntry) for n, _ in sdfg.all_nodes_recursive())
    sdfg.validate()
    state = next(s for s in sdfg.states() if node in s.nodes())
    out_edge = state.out_edges(node)[0]
    assert out_edge.data.wcr is not None          # still an update of bins, not an overwrite, for dataflow analyses
    if name == 'scat_max_shifted':                # map range 2:N, w[i - 2]  ->  keys[2:N], w[0:N-2]
        ins = {e.dst_conn: str(e.data.subset) for e in state.in_edges(node)}
        assert ins['_idx'] == '2:N' and ins['_val'] == '0:N - 2'


@pytest.mark.parametrize('name', ['not_a_scatter', 'elementwise'])
def test_rewrite_leaves_other_maps_alone(env, name):
    _, applied, nodes = _rewritten(env, name, 'pure')
    assert applied == 0 and not nodes


def test_pure_expansion_runs_like_the_original_program(env):
    rng = np.random.default_rng(0)
    n, m = 5000, 97
    keys = rng.integers(0, m, n).astype(np.int32)
    sdfg, _, _ = _rewritten(env, 'hist_aug', 'pure')
    bins = rng.integers(0, 5, m).astype(np.int32)
    exp = bins + np.bincount(keys, minlength=m).astype(np.int32)
    sdfg(keys=keys, bins=bins, N=n, M=m)
    np.testing.assert_array_equal(bins, exp)

    sdfg, _, _ = _rewritten(env, 'scat_max_shifted', 'pure')
    k64, w = keys.astype(np.int64), rng.standard_normal(n)
    bins = np.full(m, -0.5)
    exp = bins.copy()
    np.maximum.at(exp, k64[2:], w[:n - 2])
    sdfg(keys=k64, w=w, bins=bins, N=n, M=m)
    np.testing.assert_array_equal(bins, exp)


def test_b200_expansion_emits_the_c_abi_call(env):
    dace, SN, progs = env
    sdfg = progs['hist_weighted'].to_sdfg(simplify=True)
    sdfg.name = 'b200_scatter_codegen'
    sdfg.apply_gpu_transformations()
    assert SN.use_b200_scatter(sdfg) == 1
    code = {c.title: c.clean_code for c in sdfg.generate_code()}
    frame = next(v for k, v in code.items() if 'Frame' in k or k == 'b200_scatter_codegen')
    assert 'dace_b200_reduce.h' in frame
    assert 'DACE_GPU_CHECK((cudaError_t)dace_b200_wcr_scatter(3, 2, 0, _idx, _val, nullptr, (size_t)(N),' in frame
    import re
    m = re.search(r'_bins, \(size_t\)\(M\), (__b200_ws_\w+), \1_bytes, \(dace_b200_stream_t\)__dace_current_stream\)\);', frame)
    assert m, 'the call passes the scratch owned by this node'
    # sized in __dace_init from the program's symbols (per-element values: the query answers 0, a stub is kept)
    assert '__b200_need = dace_b200_wcr_scatter_workspace_bytes(3, 2, 0, 1, (size_t)(N), (size_t)(M));' in frame
    assert 'cudaMalloc(&%s, %s_bytes)' % (m.group(1), m.group(1)) in frame and 'cudaFree(%s)' % m.group(1) in frame
    assert 'reduce_atomic' not in ''.join(code.values())      # the per-element atomic kernel is gone
    assert not any(k.endswith('cuda') and 'hist_weighted' in v and '__global__' in v for k, v in code.items())


def test_b200_expansion_refuses_host_storage(env):
    dace, SN, progs = env
    sdfg, applied, _ = _rewritten(env, 'hist_aug', 'B200')     # no apply_gpu_transformations
    assert applied == 1
    with pytest.raises(ValueError, match='GPU_Global or CPU_Pinned'):
        sdfg.expand_library_nodes()


def test_auto_optimize_selects_the_backend(env):
    """the reference's auto_optimize(device=GPU) would pick GPUAuto for the
    Reduce nodes (auto_optimize.py:439-458) and leave the WCR map as an atomic kernel."""
    dace, SN, _ = env
    from dace_b200 import auto as AUTO
    N, M = dace.symbol('N'), dace.symbol('M')

    @dace.program
    def mixed(A: dace.float32[N, M], keys: dace.int32[N], bins: dace.int32[M], rs: dace.float32[N], mx: dace.float32[1]):
        rs[:] = np.sum(A, axis=1)
        mx[0] = np.max(A)
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    import dace.transformation.auto.auto_optimize as ao
    before = ao.set_fast_implementations
    sdfg = AUTO.auto_optimize(mixed.to_sdfg(simplify=True), dace.DeviceType.GPU)
    assert ao.set_fast_implementations is before          # the hook is removed again
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    assert code.count('dace_b200_reduce_nd(') == 2 and code.count('dace_b200_wcr_scatter(') == 1
    assert 'reduce_atomic' not in code
    assert AUTO.find_fast_library(dace.DeviceType.GPU) == ['B200'] and AUTO.find_fast_library(dace.DeviceType.CPU) == []
