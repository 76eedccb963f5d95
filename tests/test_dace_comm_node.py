"""CPU: the multi-GPU sharding transformation (dace_b200/comm_node.py) against the REAL dace IR.

* semantics with one process: after shard_reductions() the program - run through the reference's own
  'pure' expansions on the host - still computes what NumPy computes; identities count exactly once
  (neutral seed + one fold), scatter goes through a neutral private copy;
* generated code of the GPU variant: one dace_b200_comm_allreduce per reduction over the sharded
  axis, none for the reduction that keeps it, communicator taken in __dace_init.
Needs /root/reference; skipped on the GPU box."""
import os

import numpy as np
import pytest

from oracle.refenv import import_reference, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def dace(tmp_path_factory):
    d = tmp_path_factory.mktemp('dacecache_comm')
    cwd = os.getcwd()
    os.chdir(d)
    mod = import_reference(str(d / '.dacecache'))
    yield mod
    os.chdir(cwd)


def _program(dace, name, f, i):
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def prog(A: f[N, M], keys: dace.int32[N], cs: f[M], tot: f[1], rs: f[N], bins: i[M]):
        cs[:] = np.sum(A, axis=0)
        tot[0] = dace.reduce(lambda a, b: max(a, b), A, identity=0.25)
        rs[:] = np.sum(A, axis=1)
        for k in dace.map[0:N]:
            bins[keys[k]] += 1

    sdfg = prog.to_sdfg(simplify=True)
    sdfg.name = name
    return sdfg


def test_neutral_elements(dace):
    from dace_b200 import comm_node as C
    from dace_b200.redtypes import ReductionType as R
    assert C.neutral_element(R.Sum, dace.float32) == 0 and C.neutral_element(R.Product, dace.int64) == 1
    assert C.neutral_element(R.Max, dace.int32) == -2**31 and C.neutral_element(R.Min, dace.int32) == 2**31 - 1
    assert C.neutral_element(R.Max, dace.float32) == np.finfo(np.float32).min
    assert C.neutral_element(R.Bitwise_And, dace.int32) == -1
    assert C.neutral_element(R.Bitwise_And, dace.uint32) == 2**32 - 1
    assert C.neutral_element(R.Logical_And, dace.int32) == 1 and C.neutral_element(R.Logical_Or, dace.int32) == 0


def test_sharded_program_single_process_semantics(dace):
    from dace_b200 import comm_node as C, scatter_node as SN
    sdfg = _program(dace, 'comm_cpu_semantics', dace.float64, dace.int64)
    assert SN.use_b200_scatter(sdfg) == 1
    counts = C.shard_reductions(sdfg, {'A': 0, 'keys': 0})
    assert counts == {'reduce': 2, 'scatter': 1, 'local': 1}
    assert C.shard_reductions(sdfg, {'A': 0, 'keys': 0}) == {'reduce': 0, 'scatter': 0, 'local': 1}   # idempotent
    combines = [n for n, _ in sdfg.all_nodes_recursive() if isinstance(n, C.AllReduceCombine)]
    assert len(combines) == 3
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, (C.AllReduceCombine, SN.ScatterReduce)):
            n.implementation = 'pure'
    sdfg.validate()
    rng = np.random.default_rng(0)
    a = rng.standard_normal((50, 40))
    keys = rng.integers(0, 40, 50).astype(np.int32)
    cs, tot, rs, bins = np.zeros(40), np.zeros(1), np.zeros(50), np.full(40, 7, np.int64)
    sdfg(A=a, keys=keys, cs=cs, tot=tot, rs=rs, bins=bins, N=50, M=40)
    assert np.allclose(cs, a.sum(0)) and np.allclose(rs, a.sum(1)) and tot[0] == max(a.max(), 0.25)
    assert np.array_equal(bins, 7 + np.bincount(keys, minlength=40))
    sdfg(A=-np.abs(a), keys=keys, cs=cs, tot=tot, rs=rs, bins=bins, N=50, M=40)
    assert tot[0] == 0.25


def test_gpu_variant_emits_one_collective_per_sharded_reduction(dace):
    from dace_b200 import auto as AUTO, comm_node as C
    sdfg = _program(dace, 'comm_gpu_codegen', dace.float32, dace.int32)
    sdfg.apply_gpu_transformations()
    AUTO.apply_b200(sdfg)
    assert C.shard_reductions(sdfg, {'A': 0, 'keys': 0}) == {'reduce': 2, 'scatter': 1, 'local': 1}
    sdfg.validate()
    code = ''.join(c.clean_code for c in sdfg.generate_code())
    assert code.count('dace_b200_comm_allreduce(__state->dace_b200_comm, ') == 3
    assert 'dace_b200_comm_allreduce(__state->dace_b200_comm, 3, 0, _outbuf, (size_t)(M)' in code     # column sums, f32 Sum
    assert 'dace_b200_comm_allreduce(__state->dace_b200_comm, 2, 0, _outbuf, (size_t)(1)' in code     # max
    assert 'dace_b200_comm_allreduce(__state->dace_b200_comm, 3, 2, _outbuf, (size_t)(M)' in code     # bins, i32 Sum
    assert 'void *dace_b200_comm = nullptr;' in code and 'dace_b200_comm_default(&__state->dace_b200_comm)' in code
    assert code.count('dace_b200_reduce_nd(') == 3 and code.count('dace_b200_wcr_scatter(') == 1
    assert 'MPI_' not in code and 'nccl' not in code    # NCCL is behind the C ABI, bound at run time


def test_rejections(dace):
    from dace_b200 import comm_node as C
    from oracle.build_ref import _reduce_sdfg
    noident = _reduce_sdfg(dace, 'comm_noident', dace.float32, dace.float32, 2, [0], 'lambda a, b: a + b', None, 'pure')
    with pytest.raises(NotImplementedError, match='identity None|cannot combine'):
        C.shard_reductions(noident, {'A': 0})
    kept = _reduce_sdfg(dace, 'comm_kept', dace.float32, dace.float32, 2, [1], 'lambda a, b: a + b', 0, 'pure')
    assert C.shard_reductions(kept, {'A': 0}) == {'reduce': 0, 'scatter': 0, 'local': 1}
    host = _reduce_sdfg(dace, 'comm_host', dace.float32, dace.float32, 2, [0], 'lambda a, b: a + b', 0, 'pure')
    assert C.shard_reductions(host, {'A': 0})['reduce'] == 1
    with pytest.raises(ValueError, match='GPU_Global'):     # the B200 combine has no CPU fallback
        host.expand_library_nodes()


def test_fusing_before_sharding_is_refused(dace):
    """A map -> reduce chain fused into one kernel hides the reduction from shard_reductions(): it must say so."""
    from dace_b200 import auto as AUTO, comm_node as C
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)

    @dace.program
    def prog(A: dace.float32[N, M], s: dace.float32[M]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = A[i, j] * A[i, j]
        s[:] = np.sum(tmp, axis=0)

    fused_first = prog.to_sdfg(simplify=True)
    fused_first.apply_gpu_transformations()
    assert AUTO.apply_b200(fused_first)['fused'] == 1
    with pytest.raises(NotImplementedError, match='BEFORE the fusing'):
        C.shard_reductions(fused_first, {'A': 0})
    right_order = prog.to_sdfg(simplify=True)
    right_order.name = 'comm_then_fuse'
    right_order.apply_gpu_transformations()
    AUTO.apply_b200(right_order, fuse=False)
    assert C.shard_reductions(right_order, {'A': 0})['reduce'] == 1
    assert AUTO.apply_b200(right_order)['fused'] == 1            # the local partial is still computed by the fused kernel
    code = ''.join(c.clean_code for c in right_order.generate_code())
    assert 'dace_b200::fused_cols<3, float, float>' in code and code.count('dace_b200_comm_allreduce(') == 1
