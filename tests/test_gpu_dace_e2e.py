"""GPU: DaCe-generated programs whose Reduce nodes were expanded with 'B200'.

tests/dace_e2e/_built/*.so are ordinary DaCe build products (unmodified reference
frontend + codegen + CMake, built in the container where /root/reference exists by
tests/dace_e2e/build_programs.py).  They are loaded here the way
``CompiledSDFG`` loads them (reference dace/codegen/compiled_sdfg.py:233-240):
``__dace_init_<n>`` -> ``__program_<n>(state, arrays..., symbols...)`` ->
``__dace_exit_<n>``; host NumPy arrays in, host arrays out - the generated code
does the H2D/D2H copies and calls ``dace_b200_reduce_nd`` on its own stream.
"""
import os

import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref_loader

pytestmark = pytest.mark.gpu

BUILT = os.path.join(os.path.dirname(__file__), 'dace_e2e', '_built')
if not ref_loader.available(BUILT):
    pytest.skip('tests/dace_e2e/_built not present (built where /root/reference exists)', allow_module_level=True)


def prog(name):
    return ref_loader.RefProgram(name, BUILT)


def close_f32(got, ref, a, axes):
    scale = np.abs(a.astype(np.float64)).sum(axis=axes).reshape(np.shape(ref))
    return np.all(np.abs(got.astype(np.float64) - ref.astype(np.float64)) <= 1e-5 * np.maximum(scale, 1e-30))


def test_manifest_every_program_calls_the_c_abi(cuda):
    m = ref_loader.manifest(BUILT)
    assert len(m['programs']) >= 10
    assert all(e['b200_calls'] >= 1 for e in m['programs'].values())


def test_c1_np_sum_axis1_dace_program(cuda):
    """BASELINE C1 shape: np.sum(A, axis=1) on 4096x4096 float32 in a @dace.program."""
    a = np.random.default_rng(0).random((4096, 4096), dtype=np.float32)
    b = np.zeros(4096, np.float32)
    p = prog('b200_c1_npsum_axis1')
    p(A=a, B=b, N=4096, M=4096)
    ref = O.reduce(O.SUM, a, (1, ), identity=0, wide=True)
    assert close_f32(b, ref, a, (1, ))
    b2 = np.zeros(100, np.float32)
    p(A=np.ascontiguousarray(a[:100, :77]), B=b2, N=100, M=77)   # symbolic sizes: same .so, new shape
    assert close_f32(b2, a[:100, :77].astype(np.float64).sum(1), a[:100, :77], (1, ))


def test_np_max_axis0_int64(cuda):
    a = np.random.default_rng(1).integers(-10**12, 10**12, (777, 333))
    b = np.zeros(333, np.int64)
    prog('b200_np_max_axis0_i64')(A=a, B=b, N=777, M=333)
    np.testing.assert_array_equal(b, a.max(0))
    np.testing.assert_array_equal(b, O.reduce(O.MAX, a, (0, ), identity=np.iinfo(np.int64).min))


def test_two_reduce_nodes_in_one_program(cuda):
    a = np.random.default_rng(2).standard_normal((513, 2050)).astype(np.float32)
    mx, sm = np.zeros(513, np.float32), np.zeros(513, np.float32)
    prog('b200_softmax_stats')(A=a, mx_out=mx, sm_out=sm, N=513, M=2050)
    np.testing.assert_array_equal(mx, a.max(1))
    assert close_f32(sm, a.astype(np.float64).sum(1), a, (1, ))


@pytest.mark.parametrize('name,shape,axes,dt,red', [
    ('b200_full_sum_f32', (3000017, ), None, np.float32, O.SUM),
    ('b200_col_sum_f32', (5000, 520), (0, ), np.float32, O.SUM),
    ('b200_row_sum_i32', (999, 1001), (1, ), np.int32, O.SUM),
    ('b200_mid2_sum_f32', (123, 21, 26, 8), (1, 2), np.float32, O.SUM),
    ('b200_outer_inner_sum_f64', (64, 55, 33), (0, 2), np.float64, O.SUM),
])
def test_hand_built_reduce_nodes(name, shape, axes, dt, red, cuda):
    rng = np.random.default_rng(3)
    a = (rng.uniform(-10, 10, shape) if np.issubdtype(dt, np.floating) else rng.integers(-99, 99, shape)).astype(dt)
    kept = [s for i, s in enumerate(shape) if axes is not None and i not in axes] or [1]
    b = np.zeros(kept, dt)
    prog(name)(A=a, B=b, **{'D%d' % i: s for i, s in enumerate(shape)})
    ref = O.reduce(red, a, axes, identity=0, wide=True).reshape(b.shape)
    if dt == np.float32:
        assert close_f32(b, ref, a, axes)
    elif dt == np.float64:
        np.testing.assert_allclose(b, ref, rtol=1e-12, atol=1e-9)
    else:
        np.testing.assert_array_equal(b, ref)


def test_bf16_rows_fp32_out(cuda):
    rng = np.random.default_rng(4)
    bits = O.f32_to_bf16_bits(rng.standard_normal((300, 4096)).astype(np.float32))
    b = np.zeros(300, np.float32)
    prog('b200_row_max_bf16_f32')(A=bits, B=b, D0=300, D1=4096)
    np.testing.assert_array_equal(b, O.reduce(O.MAX, bits, (1, ), identity=-3.4028234663852886e38, in_bf16=True,
                                              out_dtype=np.float32))


def test_identity_none_folds_into_output(cuda):
    """identity=None: the fold starts from what the output holds (reference operations.py:251).  Device-resident
    arrays, so the starting values are really there (the reference's GPU transformation does not copy a
    write-only host array in)."""
    import torch
    rng = np.random.default_rng(5)
    a = rng.uniform(-1, 1, (400, 60))
    b0 = rng.uniform(0.9, 1.1, 60)               # some above, some below the column maxima
    ta, tb = torch.from_numpy(a).to(cuda), torch.from_numpy(b0).to(cuda)
    prog('b200_noident_col_max_f64')(A=ta.data_ptr(), B=tb.data_ptr(), D0=400, D1=60)
    torch.cuda.synchronize()
    exp = np.maximum(b0, a.max(0))
    assert (exp != a.max(0)).any() and (exp != b0).any()
    np.testing.assert_array_equal(tb.cpu().numpy(), exp)


def test_two_independent_reductions_in_one_state_do_not_share_a_workspace(cuda):
    """Two large reductions that both combine through tickets + partials, in one state: DaCe's default stream
    assignment runs them on different streams.  Each expanded node owns its workspace, so results stay exact
    (integers-as-floats, sums < 2^24) call after call."""
    import torch
    n, rows = (1 << 24) + 4096, 1 << 13
    g = torch.Generator(device=cuda)
    g.manual_seed(11)
    a = (torch.rand(n, device=cuda, generator=g) < 0.25).float()              # exact sum < 2^24
    c = torch.randint(0, 3, (rows, 4096), device=cuda, generator=g).float()    # exact column sums
    s = torch.zeros(1, device=cuda)
    cs = torch.zeros(4096, device=cuda)
    want_s, want_cs = a.double().sum().item(), c.double().sum(0)
    p = prog('b200_two_reductions_one_state')
    for _ in range(20):
        s.fill_(-1)
        cs.fill_(-1)
        p(A=a.data_ptr(), C=c.data_ptr(), s=s.data_ptr(), cs=cs.data_ptr(), D0=n, D1=rows)
        torch.cuda.synchronize()
        assert s.item() == want_s
        assert torch.equal(cs.double(), want_cs)


def test_persistent_transients_program(cuda):
    """make_transients_persistent: device copies of the host arrays are allocated in __dace_init."""
    a = np.random.default_rng(8).random(3000017, dtype=np.float32)
    b = np.zeros(1, np.float32)
    p = prog('b200_full_sum_f32_persistent')
    for _ in range(3):
        b[0] = -1
        p(A=a, B=b, D0=a.size)
        assert abs(float(b[0]) - float(a.sum(dtype=np.float64))) <= 1e-5 * float(a.sum(dtype=np.float64))


def test_offset_subrange_memlets(cuda):
    """reference tests/reduce_offsets_test.py:26-66."""
    rng = np.random.default_rng(6)
    a, b = rng.random((50, 50)), rng.random(25)
    exp = b.copy()
    exp[4:11] = a[25:50, 13:20].sum(0)
    prog('b200_offsets')(A=a, B=b)
    np.testing.assert_allclose(b, exp, rtol=1e-12)


def test_device_resident_arrays(cuda):
    """GPU_Global arrays: the program receives device pointers, nothing is copied."""
    import ctypes
    import torch
    x = torch.rand(1 << 22, device=cuda)
    out = torch.zeros(1, device=cuda)
    p = prog('b200_full_sum_f32_resident')
    state = p._init(x.numel())
    p._prog(state, ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(out.data_ptr()), x.numel())
    torch.cuda.synchronize()
    p._exit(state)
    assert abs(out.item() - x.double().sum().item()) <= 1e-5 * x.double().abs().sum().item()


# ---- WCR maps rewritten into the ScatterReduce library node (dace_b200/scatter_node.py) ----

def test_histogram_augassign_program(cuda):
    """``bins[keys[i]] += 1`` in a @dace.program (the npbench azimint_hist pattern)."""
    rng = np.random.default_rng(7)
    for n, m in ((100000, 256), (1 << 22, 1 << 20), (5, 3)):
        keys = rng.integers(0, m, n).astype(np.int32)
        bins = rng.integers(0, 9, m).astype(np.int32)
        exp = O.wcr_scatter(O.SUM, bins.copy(), keys)
        prog('b200_hist_augassign')(keys=keys, bins=bins, N=n, M=m)
        np.testing.assert_array_equal(bins, exp)
        np.testing.assert_array_equal(bins, exp)


def test_histogram_explicit_wcr_tasklet(cuda):
    """reference samples/explicit/histogram.py:13-26 (explicit WCR memlet), uint32 bins."""
    rng = np.random.default_rng(8)
    keys = rng.integers(0, 1000, 300000).astype(np.int32)
    bins = np.zeros(1000, np.uint32)
    prog('b200_hist_explicit_wcr')(keys=keys, bins=bins, N=keys.size, M=bins.size)
    np.testing.assert_array_equal(bins, np.bincount(keys, minlength=1000).astype(np.uint32))


def test_weighted_histogram_f32(cuda):
    rng = np.random.default_rng(9)
    keys = rng.integers(0, 4096, 200000).astype(np.int32)
    w = rng.random(200000, dtype=np.float32)
    bins = np.zeros(4096, np.float32)
    prog('b200_hist_weighted_f32')(keys=keys, w=w, bins=bins, N=keys.size, M=bins.size)
    ref = np.bincount(keys, weights=w.astype(np.float64), minlength=4096)
    np.testing.assert_allclose(bins, ref, rtol=1e-5)


def test_scatter_max_f64(cuda):
    rng = np.random.default_rng(10)
    keys = rng.integers(0, 777, 50000)
    w = rng.standard_normal(50000)
    bins = np.full(777, -0.25)
    exp = O.wcr_scatter(O.MAX, bins.copy(), keys, w)
    prog('b200_scatter_max_f64')(keys=keys, w=w, bins=bins, N=keys.size, M=bins.size)
    np.testing.assert_array_equal(bins, exp)


def test_auto_optimized_program_uses_the_backend_everywhere(cuda):
    """dace_b200.auto.auto_optimize: reference auto_optimize(device=GPU) + automatic selection; the
    program has two Reduce nodes and one WCR map, all three served by the library."""
    assert ref_loader.manifest(BUILT)['programs']['b200_auto_optimized']['b200_calls'] == 3
    rng = np.random.default_rng(11)
    n, m = 3000, 700
    a = rng.standard_normal((n, m)).astype(np.float32)
    keys = rng.integers(0, m, n).astype(np.int32)
    bins = np.zeros(m, np.int32)
    rs, mx = np.zeros(n, np.float32), np.zeros(1, np.float32)
    prog('b200_auto_optimized')(A=a, keys=keys, bins=bins, rs=rs, mx=mx, N=n, M=m)
    assert close_f32(rs, a.astype(np.float64).sum(1), a, (1, ))
    assert mx[0] == a.max()
    np.testing.assert_array_equal(bins, np.bincount(keys, minlength=m).astype(np.int32))


# ---- in-kernel reductions: 'B200 (block)' / 'B200 (block allreduce)' (dace_b200/block_expansion.py) ----

@pytest.mark.parametrize('name,dt,nblocks,nthreads,items,red', [
    ('b200_blockreduce_f32', np.float32, 2, 32, 2, 'sum'),          # reference tests/blockreduce_cudatest.py:11-56
    ('b200_blockreduce_i32_max', np.int32, 5, 160, 3, 'max'),
    ('b200_blockreduce_f64_min', np.float64, 3, 1024, 1, 'min'),
    ('b200_blockreduce_custom_i64', np.int64, 4, 72, 2, 'affine'),  # custom lambda a + b + a*b -> generated functor
])
def test_block_reduce_inside_a_generated_kernel(cuda, name, dt, nblocks, nthreads, items, red):
    rng = np.random.default_rng(11)
    n = nblocks * nthreads * items
    if red == 'affine':
        a = rng.integers(-3, 4, n).astype(dt)
    elif np.issubdtype(dt, np.floating):
        a = rng.random(n).astype(dt)
    else:
        a = rng.integers(-10**9, 10**9, n).astype(dt)
    b = np.full(nblocks, 77).astype(dt)
    prog(name)(A=a, B=b)
    blocks = a.reshape(nblocks, -1)
    if red == 'sum':
        assert close_f32(b, blocks.astype(np.float64).sum(1), blocks, (1, ))
        assert close_f32(b, O.reduce(O.SUM, blocks, (1, ), identity=0, wide=True), blocks, (1, ))
    elif red == 'max':
        np.testing.assert_array_equal(b, O.reduce(O.MAX, blocks, (1, ), identity=np.iinfo(dt).min))
    elif red == 'min':
        np.testing.assert_array_equal(b, O.reduce(O.MIN, blocks, (1, ), identity=np.finfo(dt).max))
    else:
        np.testing.assert_array_equal(b, np.prod(1 + blocks, axis=1, dtype=np.int64) - 1)


def test_block_allreduce_through_reduce_expansion(cuda):
    """dace.reduce(max, A, axis=1) -> ReduceExpansion -> inner node 'B200 (block allreduce)': one block per row,
    96 threads, result stored by thread 0 (reference block_allreduce_cudatest.py:17-45)."""
    a = np.random.default_rng(12).standard_normal((301, 96)).astype(np.float32)
    b = np.zeros(301, np.float32)
    prog('b200_blockallreduce')(A=a, B=b, N=301)
    np.testing.assert_array_equal(b, O.reduce(O.MAX, a, (1, ), identity=0))   # identity 0 like the reference test
    np.testing.assert_array_equal(b, np.maximum(a.max(1), 0))


# ---- multi-GPU sharding transformation (dace_b200/comm_node.py), single rank: the combine is a no-op ----

def test_sharded_program_single_rank(cuda):
    """b200_sharded was rewritten by shard_reductions({'A': 0, 'keys': 0}): three AllReduceCombine nodes.
    With one process (no WORLD_SIZE) the communicator has world 1 and the program must behave like the
    unsharded one - in particular identities count exactly once (non-neutral identity 0.25 of the max)."""
    m = ref_loader.manifest(BUILT)['programs']['b200_sharded']
    assert m['comm_calls'] == 3
    rng = np.random.default_rng(21)
    a = rng.standard_normal((777, 300)).astype(np.float32)
    keys = rng.integers(0, 300, 777).astype(np.int32)
    cs, tot, rs = np.zeros(300, np.float32), np.zeros(1, np.float32), np.zeros(777, np.float32)
    bins = np.full(300, 7, np.int32)
    p = prog('b200_sharded')
    p(A=a, keys=keys, cs=cs, tot=tot, rs=rs, bins=bins, N=777, M=300)
    assert close_f32(cs, a.astype(np.float64).sum(0), a, (0, ))
    assert close_f32(rs, a.astype(np.float64).sum(1), a, (1, ))
    assert tot[0] == max(a.max(), np.float32(0.25))
    np.testing.assert_array_equal(bins, 7 + np.bincount(keys, minlength=300).astype(np.int32))
    neg = -np.abs(a) - 1
    p(A=neg, keys=keys, cs=cs, tot=tot, rs=rs, bins=bins, N=777, M=300)
    assert tot[0] == np.float32(0.25)                 # the identity, folded once
    np.testing.assert_array_equal(bins, 7 + 2 * np.bincount(keys, minlength=300).astype(np.int32))


# ---- np.argmax / np.argmin in @dace.programs: struct-WCR map -> ArgReduce -> dace_b200_reduce_arg ----

def test_np_argmax_argmin_dace_programs(cuda):
    rng = np.random.default_rng(31)
    a = rng.standard_normal((513, 2050)).astype(np.float32)
    out = np.full(513, -5, np.int32)
    prog('b200_argmax_f32_axis1')(A=a, out=out, N=513, M=2050)
    np.testing.assert_array_equal(out, np.argmax(a, axis=1))
    ridx, _ = O.reduce_arg(O.MAX_LOCATION, a, 1)
    np.testing.assert_array_equal(out, ridx)
    t = rng.integers(0, 4, (300, 700)).astype(np.float64)          # many ties: the first occurrence wins
    out0 = np.zeros(700, np.int32)
    prog('b200_argmin_f64_axis0')(A=t, out=out0, N=300, M=700)
    np.testing.assert_array_equal(out0, np.argmin(t, axis=0))
    m = rng.integers(-9, 9, (33, 1000, 17)).astype(np.int32)
    outm = np.zeros((33, 17), np.int32)
    prog('b200_argmin_i32_mid')(A=m, out=outm, N=33, M=1000, K=17)
    np.testing.assert_array_equal(outm, np.argmin(m, axis=1))
    # one long row: the split-row kernel (cross-CTA combine of (val, idx) pairs)
    big = rng.standard_normal((3, 1 << 21)).astype(np.float32)
    outb = np.zeros(3, np.int32)
    prog('b200_argmax_f32_axis1')(A=big, out=outb, N=3, M=1 << 21)
    np.testing.assert_array_equal(outb, np.argmax(big, axis=1))


def test_interiors_of_larger_arrays(cuda):
    """np.sum over A[1:-1, 1:-1] and per-plane sum / row max over V[:, 1:-1, 1:-1]: Reduce nodes whose input memlets
    are windows (a pitch on one and on two levels); large enough that the one-warp strided kernel these shapes used
    to end in would take seconds."""
    rng = np.random.default_rng(5)
    a = rng.standard_normal((3000, 4100))
    s = np.zeros(1)
    prog('b200_interior_sum')(A=a, s=s, N=3000, M=4100)
    np.testing.assert_allclose(s[0], a[1:-1, 1:-1].sum(), rtol=1e-11)
    v = rng.standard_normal((40, 300, 500)).astype(np.float32)
    ps = np.zeros(40, np.float32)
    rm = np.zeros((40, 298), np.float32)
    prog('b200_interior_plane_sums')(V=v, s=ps, t=rm, K=40, N=300, M=500)
    w = v[:, 1:-1, 1:-1].astype(np.float64)
    assert np.all(np.abs(ps - w.sum((1, 2))) <= 1e-5 * np.abs(w).sum((1, 2)))
    np.testing.assert_array_equal(rm, v[:, 1:-1, 1:-1].max(2))


def test_np_argmax_over_all_elements(cuda):
    """np.argmax(A) / np.argmin(A) without an axis: flat index of the first extreme element (NumPy's rule and the
    reference frontend's sequential order); the reference itself cannot put these programs on a GPU."""
    rng = np.random.default_rng(33)
    a = rng.standard_normal((777, 3001)).astype(np.float32)
    out = np.array([-7], np.int64)
    prog('b200_flat_argmax_f32')(A=a, out=out, N=777, M=3001)
    assert out[0] == np.argmax(a)
    a[5, 17] = a[600, 2] = a.max() + 1                             # a tie: the first occurrence wins
    prog('b200_flat_argmax_f32')(A=a, out=out, N=777, M=3001)
    assert out[0] == 5 * 3001 + 17
    big = rng.standard_normal((4, 1 << 21)).astype(np.float32)     # one long flat row: split over CTAs
    prog('b200_flat_argmax_f32')(A=big, out=out, N=4, M=1 << 21)
    assert out[0] == np.argmax(big)
    m = rng.integers(-9, 9, (33, 100, 17)).astype(np.int32)        # many ties
    prog('b200_flat_argmin_i32')(A=m, out=out, N=33, M=100, K=17)
    assert out[0] == np.argmin(m)
    ridx, _ = O.reduce_arg(O.MIN_LOCATION, m.reshape(1, -1), 1)
    assert out[0] == ridx[0]


# ---- row max + row sum fused into one pass (dace_b200/fuse_node.py -> dace_b200_reduce_rows_max_sum) ----

def test_fused_row_max_sum_programs(cuda):
    import torch
    m = ref_loader.manifest(BUILT)['programs']
    assert m['b200_softmax_stats_fused']['b200_calls'] == 1       # one library call where the unfused program has two
    assert m['b200_softmax_stats']['b200_calls'] == 2
    a = np.random.default_rng(41).standard_normal((513, 2050)).astype(np.float32)
    mx, sm = np.zeros(513, np.float32), np.zeros(513, np.float32)
    prog('b200_softmax_stats_fused')(A=a, mx_out=mx, sm_out=sm, N=513, M=2050)
    np.testing.assert_array_equal(mx, a.max(1))
    np.testing.assert_array_equal(mx, O.reduce(O.MAX, a, (1, ), identity=-3.4e38))
    assert close_f32(sm, a.astype(np.float64).sum(1), a, (1, ))
    # BASELINE C3 layout, device-resident bf16 in, fp32 out (scaled down): hand-built Reduce pair, fused
    x = torch.randn(4096, 8192, device=cuda).to(torch.bfloat16)
    tmx = torch.zeros(4096, dtype=torch.float32, device=cuda)
    tsm = torch.zeros(4096, dtype=torch.float32, device=cuda)
    torch.cuda.synchronize()
    prog('b200_c3_fused_bf16')(A=x.data_ptr(), mx=tmx.data_ptr(), sm=tsm.data_ptr(), N=4096, M=8192)
    torch.cuda.synchronize()
    bits = x.view(torch.int16).cpu().numpy().view(np.uint16)
    np.testing.assert_array_equal(tmx.cpu().numpy(), O.reduce(O.MAX, bits, (1, ), identity=-3e38, in_bf16=True,
                                                              out_dtype=np.float32))
    ref = x.double().sum(1).cpu().numpy()
    scale = x.double().abs().sum(1).cpu().numpy()
    assert np.all(np.abs(tsm.cpu().numpy() - ref) <= 1e-5 * scale)


# ---- the other frontend producers: np.any / np.all / np.mean / np.min(initial=) ----

def test_np_any_all_mean_min_programs(cuda):
    rng = np.random.default_rng(51)
    a = (rng.random((400, 333)) < 0.004).astype(np.int32) * rng.integers(1, 9, (400, 333)).astype(np.int32)
    a[7, :] = 3                                         # a row that is all-true
    any_out, all_out = np.full(400, 9, np.int32), np.full(400, 9, np.int32)
    prog('b200_np_any_all')(A=a, any_out=any_out, all_out=all_out, N=400, M=333)
    np.testing.assert_array_equal(any_out, np.any(a, axis=1).astype(np.int32))
    np.testing.assert_array_equal(any_out, O.reduce(O.LOR, a, (1, ), identity=0))
    # Reference quirk (replacements/reduction.py:107-109): np.all folds `x and y` from identity 0, so it is false
    # everywhere - checked against the reference's CPU run when this test was written; the oracle restates it.
    np.testing.assert_array_equal(all_out, O.reduce(O.LAND, a, (1, ), identity=0))
    np.testing.assert_array_equal(all_out, np.zeros(400, np.int32))
    f = rng.standard_normal((1234, 77))
    b = np.zeros(77)
    prog('b200_np_mean_axis0')(A=f, B=b, N=1234, M=77)
    np.testing.assert_allclose(b, f.mean(0), rtol=1e-12, atol=1e-15)
    m = rng.integers(-3, 50, (90, 41)).astype(np.int32)
    m[:, 5] = 20                                        # never below the initial value
    bm = np.zeros(41, np.int32)
    prog('b200_np_min_initial')(A=m, B=bm, N=90, M=41)
    np.testing.assert_array_equal(bm, np.minimum(m.min(0), 7))


# ---- Map -> transient -> Reduce fused into one generated kernel (dace_b200/fused_expansion.py) ----

def test_fused_map_reduce_programs(cuda):
    import torch
    rng = np.random.default_rng(61)
    # softmax denominators, device-resident: s[i] = sum_j exp(A[i, j] - mx[i]); fused and unfused twins agree
    a = torch.randn(777, 3001, device=cuda)
    mx = a.max(dim=1).values.contiguous()
    ref = torch.exp(a.double() - mx.double()[:, None]).sum(1)
    for name in ('b200_fused_softmax_den', 'b200_unfused_softmax_den'):
        s = torch.zeros(777, device=cuda)
        torch.cuda.synchronize()
        prog(name)(A=a.data_ptr(), mx=mx.data_ptr(), s=s.data_ptr(), N=777, M=3001)
        torch.cuda.synchronize()
        assert torch.allclose(s.double(), ref, rtol=2e-5), name
    # all axes, one split row through the workspace: sum(A * A) in fp64
    f = rng.standard_normal((1500, 2111))
    out = np.zeros(1)
    prog('b200_fused_sumsq')(A=f, s=out, N=1500, M=2111)
    np.testing.assert_allclose(out[0], (f * f).sum(), rtol=1e-12)
    # explicit tasklet, integer max: exact
    x = rng.integers(-10**6, 10**6, (333, 4099)).astype(np.int32)
    y = rng.integers(-10**6, 10**6, (333, 4099)).astype(np.int32)
    o = np.zeros(333, np.int32)
    prog('b200_fused_absdiff_max')(A=x, B=y, out=o, N=333, M=4099)
    np.testing.assert_array_equal(o, np.abs(x - y).max(1))
    # 3-D, reduce the last axis, a broadcast operand and a map parameter in the expression
    t = rng.standard_normal((19, 23, 500))
    w = rng.standard_normal(500)
    o3 = np.zeros((19, 23))
    prog('b200_fused_weighted_3d')(A=t, w=w, out=o3, N=19, M=23, K=500)
    np.testing.assert_allclose(o3, (t * w[None, None, :] + np.arange(19)[:, None, None]).sum(2), rtol=1e-12, atol=1e-9)


def test_fused_programs_take_the_vector_loads_when_the_call_allows(cuda):
    """The generated load functors carry vec() for operands at the map's flat index; the call site picks the widest
    legal access from n and the pointers at run time.  Same programs as above at shapes / pointers that do (n a
    multiple of 8, aligned) and do not (base pointer 4 bytes off) qualify - same results either way."""
    import torch
    g = torch.Generator(device=cuda)
    g.manual_seed(5)
    for rows, n in ((512, 4096), (3000, 256), (40000, 64), (3, 1 << 21)):
        buf = torch.randn(rows * n + 8, device=cuda, generator=g)
        for off in (0, 1):
            a = buf[off:off + rows * n].view(rows, n)
            mx = a.max(dim=1).values.contiguous()
            ref = torch.exp(a.double() - mx.double()[:, None]).sum(1)
            s = torch.zeros(rows, device=cuda)
            prog('b200_fused_softmax_den')(A=a.data_ptr(), mx=mx.data_ptr(), s=s.data_ptr(), N=rows, M=n)
            torch.cuda.synchronize()
            assert torch.allclose(s.double(), ref, rtol=2e-5), (rows, n, off)
    rng = np.random.default_rng(66)
    f = rng.standard_normal((1024, 2048))
    out = np.zeros(1)
    prog('b200_fused_sumsq')(A=f, s=out, N=1024, M=2048)
    np.testing.assert_allclose(out[0], (f * f).sum(), rtol=1e-12)
    x = rng.integers(-10**6, 10**6, (333, 4096)).astype(np.int32)
    y = rng.integers(-10**6, 10**6, (333, 4096)).astype(np.int32)
    o = np.zeros(333, np.int32)
    prog('b200_fused_absdiff_max')(A=x, B=y, out=o, N=333, M=4096)
    np.testing.assert_array_equal(o, np.abs(x - y).max(1))
    xc = rng.integers(-10**6, 10**6, (4099, 8192)).astype(np.int32)
    yc = rng.integers(-10**6, 10**6, (4099, 8192)).astype(np.int32)
    oc = np.full(8192, -5, np.int32)
    prog('b200_wcr_cols_max_i32')(A=xc, B=yc, out=oc, N=4099, M=8192)
    np.testing.assert_array_equal(oc, np.maximum((xc - yc).max(0), -5))
    # custom lambdas (a + b + a * b on int64): rows, columns and everything, vector-eligible sizes
    for name, shape, dim in (('b200_custom_app_i64_rows', (7, 3000000), 1), ('b200_custom_app_i64_cols', (30001, 1024), 0),
                             ('b200_custom_app_i64_full', (5000000, ), 0)):
        xi = torch.randint(-3, 4, shape, device=cuda, dtype=torch.int64, generator=g) * 2
        kept = [sz for i, sz in enumerate(shape) if i != dim] if len(shape) > 1 else [1]
        oi = torch.zeros(kept, dtype=torch.int64, device=cuda)
        prog(name)(A=xi.data_ptr(), B=oi.data_ptr(), **{'D%d' % i: sz for i, sz in enumerate(shape)})
        torch.cuda.synchronize()
        fu = (xi.cpu().numpy() + 1).astype(np.uint64)
        with np.errstate(over='ignore'):
            refi = (np.multiply.reduce(fu, axis=dim, dtype=np.uint64) - np.uint64(1)).astype(np.int64)
        np.testing.assert_array_equal(oi.cpu().numpy().reshape(refi.shape), refi)


def test_fused_map_reduce_column_fold(cuda):
    """Per-channel sum of squared deviations over [N, H*W, C] (channels last): leading axes folded, last kept."""
    rng = np.random.default_rng(62)
    a = rng.standard_normal((37, 301, 96)).astype(np.float32)
    mean = a.mean((0, 1)).astype(np.float32)
    out = np.zeros(96, np.float32)
    prog('b200_fused_colstats')(A=a, mean=mean, out=out, N=37, M=301, K=96)
    ref = ((a.astype(np.float64) - mean.astype(np.float64)) ** 2).sum((0, 1))
    np.testing.assert_allclose(out, ref, rtol=2e-5)


def test_sharded_fused_program_single_rank(cuda):
    a = np.random.default_rng(71).standard_normal((2001, 333)).astype(np.float32)
    sq = np.zeros(333, np.float32)
    prog('b200_sharded_fused')(A=a, sq=sq, N=2001, M=333)
    ref = (a.astype(np.float64) ** 2).sum(0)
    assert np.all(np.abs(sq - ref) <= 1e-5 * ref)


# ---- reductions written as WCR maps (no indirection) -> fused tree-reduction kernels ----

def test_wcr_map_reductions(cuda):
    rng = np.random.default_rng(81)
    a = rng.standard_normal((513, 3001)).astype(np.float32)
    mx = a.max(1)
    s = np.full(513, 2.5, np.float32)                        # WCR accumulates into what is there
    prog('b200_wcr_rows')(A=a, mx=mx, s=s, N=513, M=3001)
    ref = 2.5 + np.exp(a.astype(np.float64) - mx.astype(np.float64)[:, None]).sum(1)
    np.testing.assert_allclose(s, ref, rtol=2e-5)
    f = rng.standard_normal((1111, 777))
    tot = np.array([10.0])
    p = prog('b200_wcr_all_f64')
    p(A=f, s=tot, N=1111, M=777)
    np.testing.assert_allclose(tot[0], 10.0 + (f * f).sum(), rtol=1e-12)
    p(A=f, s=tot, N=1111, M=777)                             # accumulates again
    np.testing.assert_allclose(tot[0], 10.0 + 2 * (f * f).sum(), rtol=1e-12)
    x = rng.integers(-10**6, 10**6, (4099, 333)).astype(np.int32)
    y = rng.integers(-10**6, 10**6, (4099, 333)).astype(np.int32)
    out = np.full(333, -5, np.int32)
    out[7] = 2**31 - 1                                       # an existing value that must survive the max
    prog('b200_wcr_cols_max_i32')(A=x, B=y, out=out, N=4099, M=333)
    exp = np.maximum((x - y).max(0), -5)
    exp[7] = 2**31 - 1
    np.testing.assert_array_equal(out, exp)


def test_one_map_with_two_wcr_outputs_is_split_and_fused(cuda):
    """reference tests/multiple_cr_test.py (a sum and a product out of one map) and a rows variant (row sums and row
    maxima of squares from one 2-D map): dace_b200/split_wcr.py gives each output its own map, both become fused tree
    reductions that fold into the existing outputs.  Integers exact (the product wraps modulo 2^64 like the
    reference's), fp32 sums to 1e-5, the max exact."""
    rng = np.random.default_rng(17)
    a = rng.integers(1, 4, 100003).astype(np.int64)
    s, p = np.array([5], np.int64), np.array([3], np.int64)
    prog('b200_multiple_cr')(A=a, s=s, p=p, N=a.size)
    with np.errstate(over='ignore'):
        assert s[0] == 5 + a.sum() and p[0] == np.multiply.reduce(a, dtype=np.int64) * np.int64(3)
    x = rng.standard_normal((777, 4100)).astype(np.float32)
    rs, big = np.full(777, 1.5, np.float32), np.full(777, 0.25, np.float32)
    big[5] = 1e9                                              # an existing value that must survive the max
    prog('b200_two_wcr_rows')(A=x, s=rs, big=big, N=777, M=4100)
    np.testing.assert_allclose(rs, 1.5 + x.astype(np.float64).sum(1), rtol=1e-5, atol=1e-3)
    exp = np.maximum((x * x).max(1), np.float32(0.25))
    exp[5] = 1e9
    np.testing.assert_array_equal(big, exp)


def test_fused_map_reduce_fp16_accumulates_in_fp32(cuda):
    """fp16 producer and fp16 output, fp32 accumulator: the result is the fp32 sum rounded once to fp16."""
    rng = np.random.default_rng(91)
    a = (rng.random((257, 5000)) * 0.5).astype(np.float16)
    s = np.zeros(257, np.float16)
    prog('b200_fused_f16_rows')(A=a, s=s, N=257, M=5000)
    sq = (a * a)                                             # the producer rounds each product to fp16
    ref = sq.astype(np.float64).sum(1)
    assert np.all(np.abs(s.astype(np.float64) - ref) <= 2.0 ** -10 * ref + 1e-3)    # one fp16 rounding of the sum
    naive16 = sq.cumsum(axis=1, dtype=np.float16)[:, -1].astype(np.float64)          # fp16 accumulation drifts far more
    assert np.abs(s.astype(np.float64) - ref).max() <= np.abs(naive16 - ref).max()


def _custom_golden():
    import json
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'custom_golden.npz'))
    meta = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'custom_golden_meta.json')))['custom']
    return g, meta


@pytest.mark.parametrize('tag', ['app_i64_rows', 'app_i64_cols', 'app_i64_full', 'app_i64_lead2', 'absmax_f32_rows',
                                 'absmax_f32_noident', 'refmin_f32_full'])
def test_custom_wcr_lambdas_match_the_reference_pure_expansion(tag, cuda):
    """ReductionType.Custom Reduce nodes expanded with 'B200' (generated functor + header kernels) replay the golden
    vectors the unmodified reference produced with its `pure` expansion - bit for bit (integer arithmetic; the
    float lambdas only select)."""
    import torch
    g, meta = _custom_golden()
    m = next(x for x in meta if x['tag'] == tag)
    a = g[tag + '__in']
    want = g[tag + '__out']
    b0 = g[tag + '__init'] if m['identity'] is None else np.full(want.shape, 77, want.dtype)
    ta, tb = torch.from_numpy(a).to(cuda), torch.from_numpy(b0.copy()).to(cuda)
    syms = {'D%d' % i: s for i, s in enumerate(a.shape)}
    prog('b200_custom_' + tag)(A=ta.data_ptr(), B=tb.data_ptr(), **syms)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(tb.cpu().numpy(), want)


def test_custom_wcr_lambda_large_shapes_same_program(cuda):
    """The same DaCe-built programs at sizes that split rows over CTAs / row slices over the workspace."""
    import torch
    g = torch.Generator(device=cuda)
    g.manual_seed(3)
    for name, shape, dim in (('b200_custom_app_i64_rows', (7, 3000001), 1), ('b200_custom_app_i64_cols', (300001, 70), 0),
                             ('b200_custom_app_i64_full', (5000003, ), 0)):
        x = torch.randint(-3, 4, shape, device=cuda, dtype=torch.int64, generator=g) * 2
        kept = [s for i, s in enumerate(shape) if i != dim] if len(shape) > 1 else [1]
        out = torch.zeros(kept, dtype=torch.int64, device=cuda)
        prog(name)(A=x.data_ptr(), B=out.data_ptr(), **{'D%d' % i: s for i, s in enumerate(shape)})
        torch.cuda.synchronize()
        f = (x.cpu().numpy() + 1).astype(np.uint64)
        with np.errstate(over='ignore'):
            ref = (np.multiply.reduce(f, axis=dim, dtype=np.uint64) - np.uint64(1)).astype(np.int64)
        np.testing.assert_array_equal(out.cpu().numpy().reshape(ref.shape), ref)


def test_shared_tile_wcr_flush_through_the_b200_header(cuda):
    """AccumulateTransient + shared-memory tile; the generated dace::ResetShared / SharedToGlobal1D::Accum
    call sites are served by include/dace_b200_shared_accum.cuh.  Integer histogram bit-exact; the fp32 variant adds
    small integers (exact in fp32), so it is exact too."""
    rng = np.random.default_rng(12)
    n = 1024 * 3000
    keys = rng.integers(0, 256, n).astype(np.int32)
    keys[: n // 2] = rng.integers(3, 9, n // 2)            # most of the tile stays untouched in many blocks
    bins = rng.integers(0, 5, 256).astype(np.int32)
    want = bins + np.bincount(keys, minlength=256).astype(np.int32)
    prog('b200_shared_tile_hist')(keys=keys, bins=bins, N=n)
    np.testing.assert_array_equal(bins, want)
    n2 = 512 * 5000
    k2 = rng.integers(0, 512, n2).astype(np.int32)
    w2 = rng.integers(0, 4, n2).astype(np.float32)
    b2 = np.zeros(512, np.float32)
    prog('b200_shared_tile_weighted_f32')(keys=k2, w=w2, bins=b2, N=n2)
    np.testing.assert_array_equal(b2, np.bincount(k2, weights=w2, minlength=512).astype(np.float32))
