"""GPU parity: the sm_100a kernels (through the C ABI) against the CPU oracle.

Cases follow the reference's own tests: the shape/axes list of reference
tests/library/reduce_test.py:52-79, the dtypes/ops of reference
tests/numpy/reductions_test.py, exact max with explicit identity
(reference tests/wcr_cudatest.py:31-65), sub-range / offset memlets
(reference tests/reduce_offsets_test.py:26-66, tests/library/reduce_test.py:31-49).
Bar: bit-exact for integers, <= 1e-5 relative (fp32) against the fp64-accumulated
oracle, exact for min/max.
"""
import numpy as np
import pytest
import torch

import dace_b200 as db
from oracle import oracle as O

pytestmark = pytest.mark.gpu

RTOL_F32 = 1e-5   # north_star: fp32 <= 1e-5 against an fp64 host reference
RTOL_F64 = 1e-12

_OPS = {'sum': (O.SUM, 'lambda a, b: a + b', 0), 'prod': (O.PRODUCT, 'lambda a, b: a * b', 1),
        'max': (O.MAX, 'lambda a, b: max(a, b)', None), 'min': (O.MIN, 'lambda a, b: min(a, b)', None),
        'band': (O.BAND, 'lambda a, b: a & b', -1), 'bor': (O.BOR, 'lambda a, b: a | b', 0),
        'bxor': (O.BXOR, 'lambda a, b: a ^ b', 0), 'land': (O.LAND, 'lambda a, b: a and b', 1),
        'lor': (O.LOR, 'lambda a, b: a or b', 0), 'lxor': (O.LXOR, 'lambda a, b: a != b', 0)}


def _identity(op, dtype):
    ident = _OPS[op][2]
    if ident is not None:
        return ident
    if np.issubdtype(dtype, np.floating):
        info = np.finfo(dtype)
    else:
        info = np.iinfo(dtype)
    return info.min if op == 'max' else info.max   # dace.dtypes.min_value/max_value (reference dtypes.py:426-449)


def _data(rng, shape, dtype, op='sum'):
    if np.issubdtype(dtype, np.floating):
        if op == 'prod':
            return rng.uniform(0.9, 1.1, shape).astype(dtype)
        return rng.uniform(-10, 10, shape).astype(dtype)   # reference tests/numpy/common.py:16-120
    if op == 'prod':
        return rng.integers(-3, 4, shape).astype(dtype)
    lo = 0 if np.issubdtype(dtype, np.unsignedinteger) else -10
    return rng.integers(lo, 10, shape).astype(dtype)


def _check(got, ref, dtype, a=None, axes=None, op='sum'):
    if np.issubdtype(dtype, np.floating) and op in ('sum', 'prod'):
        rtol = RTOL_F32 if dtype == np.float32 else RTOL_F64
        if op == 'sum' and a is not None:
            # relative to the magnitude of the summed data (cancellation-safe)
            scale = np.abs(a.astype(np.float64)).sum(axis=axes)
            assert np.all(np.abs(got.astype(np.float64) - ref.astype(np.float64)) <= rtol * np.maximum(scale, 1e-30))
        else:
            np.testing.assert_allclose(got, ref, rtol=rtol * 10)
    else:
        np.testing.assert_array_equal(got, ref)


def _run(a, axes, op, cuda, identity='default', out_init=None):
    red, lam, _ = _OPS[op]
    ident = _identity(op, a.dtype) if identity == 'default' else identity
    ta = torch.from_numpy(a).to(cuda)
    if out_init is None:
        got = db.reduce(lam, ta, axis=axes, identity=ident).cpu().numpy()
        ref = O.reduce(red, a, axes, identity=ident, wide=True)
    else:
        to = torch.from_numpy(out_init.copy()).to(cuda)
        db.reduce(lam, ta, to, axis=axes, identity=ident)
        got = to.cpu().numpy()
        ref = out_init.copy()
        O.reduce(red, a, axes, identity=ident, out=ref, wide=True)
    return got, ref


# The parametrize list of reference tests/library/reduce_test.py:53-58 (sizes kept).
REF_CASES = [
    ([8, 512, 4096], (0, 1), np.float32), ([8, 512, 4096], (0, 1), np.float64),
    ([1024, 8], (0, ), np.float32), ([1000000], (0, ), np.float64), ([1111111], (0, ), np.float64),
    ([123, 21, 26, 8], (1, 2), np.float32), ([2, 512, 2], (0, 2), np.float32),
    ([512, 555, 257], (0, 2), np.float64), ([17, 255, 12], (1, ), np.float32),
    ([64, 1024], (1, ), np.float32), ([512, 1024], (1, ), np.float64),
]


@pytest.mark.parametrize('shape,axes,dtype', REF_CASES)
def test_reference_case_list_sum(shape, axes, dtype, cuda):
    rng = np.random.default_rng(42)
    a = rng.random(shape).astype(dtype)   # reference reduce_test.py:64 uses np.random.rand
    got, ref = _run(a, axes, 'sum', cuda)
    _check(got, ref, dtype, a, axes)
    np.testing.assert_allclose(got, np.sum(a, axis=axes), rtol=1e-5 if dtype == np.float32 else 1e-10)


@pytest.mark.parametrize('dtype', [np.float32, np.float64, np.int32, np.int64, np.uint32, np.uint64, np.int8,
                                   np.uint8, np.int16, np.uint16])
@pytest.mark.parametrize('op', ['sum', 'prod', 'max', 'min'])
@pytest.mark.parametrize('shape,axes', [([4097], None), ([33, 1000], (1, )), ([1000, 33], (0, )),
                                        ([5, 300, 7], (1, )), ([6, 50, 9], (0, 2))])
def test_ops_dtypes_layouts(dtype, op, shape, axes, cuda):
    rng = np.random.default_rng(1)
    a = _data(rng, shape, dtype, op)
    got, ref = _run(a, axes, op, cuda)
    _check(got, ref, dtype, a, axes, op)


@pytest.mark.parametrize('dtype', [np.int32, np.int64, np.uint32, np.uint8])
@pytest.mark.parametrize('op', ['band', 'bor', 'bxor', 'land', 'lor', 'lxor'])
def test_bitwise_logical(dtype, op, cuda):
    rng = np.random.default_rng(2)
    for shape, axes in [([5000], None), ([37, 129], (1, )), ([129, 37], (0, ))]:
        if op.startswith('l'):
            a = rng.integers(0, 2, shape).astype(dtype)
        else:
            a = rng.integers(0, 127, shape).astype(dtype)
        ident = _OPS[op][2]
        if ident == -1:
            ident = np.array(-1).astype(dtype, casting='unsafe').item() if np.issubdtype(dtype, np.signedinteger) \
                else np.iinfo(dtype).max
        got, ref = _run(a, axes, op, cuda, identity=ident)
        np.testing.assert_array_equal(got, ref)


@pytest.mark.parametrize('n', [0, 1, 2, 7, 8, 9, 31, 255, 256, 257, 4095, 65536 + 3, (1 << 20) + 5, (1 << 24) + 1])
def test_full_sum_sizes_and_alignment(n, cuda):
    """Ragged sizes and every start alignment (memlet offsets move `_in` anywhere)."""
    rng = np.random.default_rng(3)
    base = rng.uniform(-1, 1, n + 16).astype(np.float32)
    tb = torch.from_numpy(base).to(cuda)
    for off in (0, 1, 3, 5, 8):
        a = base[off:off + n]
        got = db.reduce('lambda a, b: a + b', tb[off:off + n], axis=None, identity=0).cpu().numpy()
        ref = O.reduce(O.SUM, a, None, identity=0, wide=True)
        _check(got, ref, np.float32, a, None)


def test_full_int_exact_large(cuda):
    rng = np.random.default_rng(4)
    a = rng.integers(-1000, 1000, (1 << 24) + 77, dtype=np.int32)
    got, ref = _run(a, None, 'sum', cuda)
    np.testing.assert_array_equal(got, ref)
    a64 = rng.integers(-10**12, 10**12, (1 << 22) + 3, dtype=np.int64)
    for op in ('sum', 'max', 'min'):
        got, ref = _run(a64, None, op, cuda)
        np.testing.assert_array_equal(got, ref)


def test_wcr_cudatest_exact_max(cuda):
    """reference tests/wcr_cudatest.py:31-65: max with identity -999999, exact."""
    a = np.random.default_rng(5).random(10).astype(np.float32)
    got, ref = _run(a, None, 'max', cuda, identity=-999999)
    assert got == ref == a.max()


def test_identity_none_accumulates(cuda):
    """identity=None folds into the existing output (reference operations.py:251)."""
    rng = np.random.default_rng(6)
    a = rng.integers(-10, 10, (40, 50)).astype(np.int64)
    init = rng.integers(-5, 5, 40).astype(np.int64)
    got, ref = _run(a, (1, ), 'sum', cuda, identity=None, out_init=init)
    np.testing.assert_array_equal(got, ref)
    np.testing.assert_array_equal(got, init + a.sum(1))
    got, ref = _run(a, (0, ), 'max', cuda, identity=None, out_init=np.full(50, 3, np.int64))
    np.testing.assert_array_equal(got, np.maximum(3, a.max(0)))
    got, ref = _run(a.reshape(-1), None, 'sum', cuda, identity=None, out_init=np.array(7, np.int64))
    assert got == 7 + a.sum()


def test_nonneutral_identity(cuda):
    a = np.arange(24, dtype=np.int32).reshape(4, 6)
    got, ref = _run(a, (1, ), 'sum', cuda, identity=100)
    np.testing.assert_array_equal(got, 100 + a.sum(1))
    np.testing.assert_array_equal(got, ref)


def test_offsets_and_subranges(cuda):
    """reference tests/reduce_offsets_test.py:26-66: A[25:50, 13:20] -> B[4:11]."""
    rng = np.random.default_rng(7)
    A = rng.random((50, 50))
    B = rng.random(25)
    tA, tB = torch.from_numpy(A).to(cuda), torch.from_numpy(B.copy()).to(cuda)
    db.reduce('lambda a, b: a + b', tA[25:50, 13:20], tB[4:11], axis=0, identity=0)
    exp = B.copy()
    exp[4:11] = A[25:50, 13:20].sum(0)
    np.testing.assert_allclose(tB.cpu().numpy(), exp, rtol=1e-12)
    ref = B.copy()
    O.reduce(O.SUM, A[25:50, 13:20], (0, ), identity=0, out=ref[4:11], wide=True)
    np.testing.assert_allclose(tB.cpu().numpy(), ref, rtol=1e-12)
    # no-identity max into a sub-range
    tB2 = torch.from_numpy(B.copy()).to(cuda)
    db.reduce('lambda a, b: max(a, b)', tA[25:50, 13:20], tB2[4:11], axis=0, identity=None)
    exp = B.copy()
    exp[4:11] = np.maximum(B[4:11], A[25:50, 13:20].max(0))
    np.testing.assert_array_equal(tB2.cpu().numpy(), exp)


def test_strided_memlet_5d(cuda):
    """reference tests/library/reduce_test.py:31-49: A[0:N,0,0,0:Cin,0:Cout] -> C[0:N,5,5,0:Cout]."""
    rng = np.random.default_rng(8)
    A = rng.random((3, 4, 4, 6, 5)).astype(np.float32)
    C = np.zeros((3, 8, 8, 5), np.float32)
    tA, tC = torch.from_numpy(A).to(cuda), torch.from_numpy(C).to(cuda)
    db.reduce('lambda a, b: a + b', tA[:, 0, 0, :, :], tC[:, 5, 5, :], axis=1, identity=0)
    exp = C.copy()
    exp[:, 5, 5, :] = A[:, 0, 0].sum(1)
    np.testing.assert_allclose(tC.cpu().numpy(), exp, rtol=1e-5)


def test_transposed_and_stepped_views(cuda):
    rng = np.random.default_rng(9)
    A = rng.integers(-50, 50, (64, 48, 10)).astype(np.int32)
    tA = torch.from_numpy(A).to(cuda)
    for view_np, view_t, axes in [(A.transpose(2, 0, 1), tA.permute(2, 0, 1), (1, )),
                                  (A[::2, ::3], tA[::2, ::3], (0, 1)),
                                  (A[:, :, ::2], tA[:, :, ::2], (2, )),
                                  (A.transpose(1, 0, 2), tA.permute(1, 0, 2), None)]:
        got = db.reduce('lambda a, b: a + b', view_t, axis=axes, identity=0).cpu().numpy()
        ref = O.reduce(O.SUM, view_np, axes, identity=0)
        np.testing.assert_array_equal(got, ref)
        np.testing.assert_array_equal(got, view_np.sum(axis=axes))


def test_degenerate_and_empty(cuda):
    a = np.arange(12, dtype=np.float32).reshape(3, 1, 4)
    ta = torch.from_numpy(a).to(cuda)
    got = db.reduce('lambda a, b: a + b', ta, axis=1, identity=5).cpu().numpy()   # run-time unit axis: normal fold
    np.testing.assert_array_equal(got, 5 + a[:, 0, :])
    np.testing.assert_array_equal(got, O.reduce(O.SUM, a, (1, ), identity=5))
    got = db.reduce('lambda a, b: a + b', ta[:, 0, :], axis=(), identity=5).cpu().numpy()   # no axis: copy
    np.testing.assert_array_equal(got, a[:, 0, :])
    np.testing.assert_array_equal(got, O.reduce(O.SUM, a[:, 0, :], (), identity=5))
    e = torch.empty((4, 0), dtype=torch.float32, device=cuda)
    got = db.reduce('lambda a, b: a + b', e, axis=1, identity=2.5).cpu().numpy()
    np.testing.assert_array_equal(got, np.full(4, 2.5, np.float32))
    e2 = torch.empty((0, 4), dtype=torch.float32, device=cuda)
    assert db.reduce('lambda a, b: a + b', e2, axis=1, identity=0).numel() == 0


@pytest.mark.parametrize('rows,n', [(1, 1 << 22), (3, (1 << 20) + 17), (70, 40000), (500, 9000), (4096, 4096),
                                    (10000, 100), (100000, 8), (7, 3), (65536, 33)])
def test_rows_all_decompositions(rows, n, cuda):
    """Every row-kernel regime: sub-warp groups, warp, CTA per row, split rows."""
    rng = np.random.default_rng(10)
    a = rng.uniform(-1, 1, (rows, n)).astype(np.float32)
    for op in ('sum', 'max'):
        got, ref = _run(a, (1, ), op, cuda)
        _check(got, ref, np.float32, a, (1, ), op)
    ai = rng.integers(-100, 100, (rows, n)).astype(np.int32)
    got, ref = _run(ai, (1, ), 'sum', cuda)
    np.testing.assert_array_equal(got, ref)


@pytest.mark.parametrize('outer,red,inner', [(1, 1 << 16, 4096), (1, 100000, 8), (1, 5000, 1), (3, 1000, 24),
                                             (123, 546, 8), (1, 64, 100000), (2, 7, 5), (1, 33000, 257),
                                             (9, 2000, 264), (5000, 16, 16)])
def test_cols_all_decompositions(outer, red, inner, cuda):
    rng = np.random.default_rng(11)
    a = rng.uniform(-1, 1, (outer, red, inner)).astype(np.float32)
    got, ref = _run(a, (1, ), 'sum', cuda)
    _check(got, ref, np.float32, a, (1, ))
    ai = rng.integers(-100, 100, (outer, red, inner)).astype(np.int64)
    for op in ('sum', 'min'):
        got, ref = _run(ai, (1, ), op, cuda)
        np.testing.assert_array_equal(got, ref)


@pytest.mark.parametrize('outer,red,inner', [(1, 7, 300004), (1, 7, 300001), (2, 5, 400000), (1, 3, 1000002),
                                             (1, 64, 80000), (3, 1, 90000), (1, 9, 77777 * 4)])
def test_few_rows_many_columns(outer, red, inner, cuda):
    """[R <= 64, K] column folds (thread-per-column kernel, CTAs walk 1024-column chunks): whole chunks and the ragged
    last one, 1 ... 64 rows (groups of four + tail), several outer slices, a view that starts one element into the
    allocation, folding into an existing output."""
    rng = np.random.default_rng(outer * 100 + red)
    a = rng.uniform(-1, 1, (outer, red, inner)).astype(np.float32)
    got, ref = _run(a, (1, ), 'sum', cuda)
    _check(got, ref, np.float32, a, (1, ))
    for dt in (np.int32, np.int64, np.float64):
        ai = rng.integers(-100, 100, (outer, red, inner)).astype(dt)
        for op in ('sum', 'max'):
            got, ref = _run(ai, (1, ), op, cuda)
            np.testing.assert_array_equal(got, ref)
    # a view one element into the allocation (same pitch): not 16-byte aligned any more
    buf = torch.from_numpy(rng.integers(-100, 100, outer * red * inner + 4).astype(np.int32)).to(cuda)
    v = buf[1:1 + outer * red * inner].view(outer, red, inner)
    out = torch.empty((outer, inner), dtype=torch.int32, device=cuda)
    db.reduce('lambda a, b: a + b', v, out, axis=1, identity=0)
    assert torch.equal(out, v.sum(1, dtype=torch.int32))
    # identity None: fold into the existing output
    out2 = torch.full((outer, inner), 5, dtype=torch.int32, device=cuda)
    w = buf[:outer * red * inner].view(outer, red, inner)
    db.reduce('lambda a, b: a + b', w, out2, axis=1, identity=None)
    assert torch.equal(out2, w.sum(1, dtype=torch.int32) + 5)


def test_bf16_f16_fp32_accumulate(cuda):
    """bf16/f16 inputs, fp32 accumulation and fp32 outputs; sum of 128 ones == 128
    (reference tests/fp8_bf16_dtypes_test.py:130-150)."""
    rng = np.random.default_rng(12)
    ones = torch.ones(128, dtype=torch.bfloat16, device=cuda)
    out = torch.zeros((), dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: a + b', ones, out, axis=None, identity=0)
    assert out.item() == 128.0
    assert db.reduce('lambda a, b: a + b', ones, axis=None, identity=0).float().item() == 128.0
    x = rng.standard_normal((300, 1000)).astype(np.float32)
    bits = O.f32_to_bf16_bits(x)
    tx = torch.from_numpy(bits.view(np.int16)).to(cuda).view(torch.bfloat16)
    for axes, shape in (((1, ), 300), ((0, ), 1000), (None, ())):
        for red, lam, ident in ((O.SUM, 'lambda a, b: a + b', 0), (O.MAX, 'lambda a, b: max(a, b)', -3e38)):
            out = torch.empty(shape, dtype=torch.float32, device=cuda)
            db.reduce(lam, tx, out, axis=axes, identity=ident)
            ref = O.reduce(red, bits, axes, identity=ident, in_bf16=True, out_dtype=np.float32, wide=True)
            if red == O.SUM:
                scale = np.abs(O.bf16_bits_to_f32(bits).astype(np.float64)).sum(axis=axes)
                assert np.all(np.abs(out.cpu().numpy().astype(np.float64) - ref) <= RTOL_F32 * scale)
            else:
                np.testing.assert_array_equal(out.cpu().numpy(), ref)
    h = rng.standard_normal((64, 512)).astype(np.float16)
    th = torch.from_numpy(h).to(cuda)
    out = torch.empty(64, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: a + b', th, out, axis=1, identity=0)
    np.testing.assert_allclose(out.cpu().numpy(), h.astype(np.float64).sum(1), rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize('rows,n,dtype', [(512, 8192, torch.bfloat16), (100, 1000, torch.float32),
                                          (33, 77, torch.float16), (4096, 64, torch.bfloat16), (2, 100000, torch.float32),
                                          # tiny dense rows through the shared-memory tile (1 ... 32 lanes per row)
                                          (500001, 10, torch.float32), (140000, 33, torch.bfloat16), (70001, 64, torch.float16),
                                          (45000, 100, torch.float32), (40000, 128, torch.bfloat16), (2100000, 2, torch.float32),
                                          # few long rows: clusters of 1 ... 8 CTAs per row
                                          (1, 3000001, torch.float32), (8, 1 << 20, torch.bfloat16), (100, 70001, torch.float32),
                                          (700, 33000, torch.float16), (3000, 16400, torch.float32)])
def test_rows_max_sum_fused(rows, n, dtype, cuda):
    g = torch.Generator(device='cpu').manual_seed(0)
    x = torch.randn(rows, n, generator=g).to(dtype)
    tx = x.to(cuda)
    mx, sm = db.rows_max_sum(tx)
    xd = x.double().numpy()
    np.testing.assert_array_equal(mx.cpu().numpy(), xd.max(1).astype(np.float32))
    scale = np.abs(xd).sum(1)
    assert np.all(np.abs(sm.cpu().numpy().astype(np.float64) - xd.sum(1)) <= RTOL_F32 * scale)
    # and identical to running the two Reduce nodes separately
    o1 = torch.empty(rows, dtype=torch.float32, device=cuda)
    if dtype != torch.float32:
        db.reduce('lambda a, b: max(a, b)', tx, o1, axis=1, identity=-3e38)
        np.testing.assert_array_equal(o1.cpu().numpy(), mx.cpu().numpy())


def test_custom_minmax_lambda_and_errors(cuda):
    a = torch.arange(100, dtype=torch.float64, device=cuda) - 50
    got = db.reduce('lambda a, b: a if a < b else b', a, axis=None, identity=1e30)   # reference custom_reduce_test.py
    assert got.item() == -50
    with pytest.raises(NotImplementedError):
        db.reduce('lambda a, b: a * a + b', a, axis=None, identity=0)
    with pytest.raises(ValueError):
        db.reduce('lambda a, b: a + b', a.cpu(), axis=None, identity=0)


def test_deterministic_and_workspace(cuda):
    """Fixed-order cross-CTA combine: bitwise repeatable float sums; explicit workspace."""
    g = torch.Generator(device='cpu').manual_seed(1)
    x = torch.rand(1 << 24, generator=g).to(cuda)
    ws = db.Workspace(cuda)
    r = [db.reduce('lambda a, b: a + b', x, axis=None, identity=0, workspace=ws).item() for _ in range(5)]
    assert len(set(r)) == 1
    r2 = db.reduce('lambda a, b: a + b', x, axis=None, identity=0).item()
    assert r2 == r[0]
    m = x.view(4096, 4096)
    c = [db.reduce('lambda a, b: a + b', m, axis=0, identity=0, workspace=ws) for _ in range(3)]
    assert torch.equal(c[0], c[1]) and torch.equal(c[1], c[2])


@pytest.mark.parametrize('outer,red,inner,dtype', [
    (1, 5000, 1024, np.float32), (1, 3001, 4096, np.float32), (1, 20000, 2048 + 8, np.float32),
    (1, 4000, 512, np.float64), (3, 2000, 1024, np.int32), (1, 70000, 1024, np.int64), (2, 999, 8192, np.float32),
    (1, 1 << 15, 1024, np.uint32), (300, 400, 2048, np.float32),
])
def test_cols_wide_rows_kernel(outer, red, inner, dtype, cuda):
    """Rows of >= 4 KiB take cols_wide_kernel (1024-thread CTAs, warps side by side along the
    columns, two-level fixed-order combine through the workspace): all ops, partial column chunks,
    outer > 1, split counts from 1 to the SM count, identity None, repeatability."""
    rng = np.random.default_rng(red)
    a = _data(rng, (outer, red, inner), dtype)
    for op in ('sum', 'max'):
        got, ref = _run(a, (1, ), op, cuda)
        _check(got, ref, dtype, a, (1, ), op)
    init = _data(rng, (outer, inner), dtype)
    got, ref = _run(a, (1, ), 'sum', cuda, identity=None, out_init=init)
    if np.issubdtype(dtype, np.floating):
        scale = np.abs(a.astype(np.float64)).sum(axis=1) + np.abs(init)
        assert np.all(np.abs(got.astype(np.float64) - ref.astype(np.float64)) <= 1e-5 * np.maximum(scale, 1e-30))
    else:
        np.testing.assert_array_equal(got, ref)
    ta = torch.from_numpy(a).to(cuda)
    r1 = db.reduce('lambda a, b: a + b', ta, axis=(1, ), identity=0)
    r2 = db.reduce('lambda a, b: a + b', ta, axis=(1, ), identity=0)
    assert torch.equal(r1, r2)                       # fixed fold order: bit-identical from run to run


def test_cols_wide_rows_bf16_and_2d(cuda):
    x = torch.randn(6000, 4096, device=cuda).to(torch.bfloat16)
    got = torch.zeros(4096, device=cuda)
    db.reduce('lambda a, b: a + b', x, got, axis=(0, ), identity=0)   # bf16 in, fp32 out, fp32 accumulate
    ref = x.double().sum(0)
    assert bool(((got.double() - ref).abs() <= 1e-5 * x.double().abs().sum(0)).all())
    gmax = db.reduce('lambda a, b: max(a, b)', x, axis=(0, ), identity=-3e38)
    assert torch.equal(gmax.float(), x.float().max(0).values)


# ---- [R1, K, R2] with contiguous R2: two streaming passes (rows -> scratch -> cols), api.cu plan_two_pass ----

@pytest.mark.parametrize('dtype', [np.float32, np.float64, np.int32, np.int64, np.uint32, np.uint64, np.int8])
@pytest.mark.parametrize('op', ['sum', 'max', 'min', 'prod'])
@pytest.mark.parametrize('shape', [[6, 50, 64], [33, 7, 1000], [3, 300, 32], [2, 5, 40000],
                                   [300, 7, 33], [2049, 64, 8], [70, 5, 70], [5000, 3, 2]])   # R1 >= R2: outer axis first
def test_outer_inner_axes_two_pass(dtype, op, shape, cuda):
    """Per-channel statistics layout (axes (0, 2) of [N, C, H*W]); every type through the two-pass plans and must
    still be right through the generic kernel."""
    rng = np.random.default_rng(3)
    a = _data(rng, shape, dtype, op)
    got, ref = _run(a, (0, 2), op, cuda)
    _check(got, ref, dtype, a, (0, 2), op)
    if op in ('sum', 'max'):                       # identity None: fold into what the output holds
        init = _data(rng, [shape[1]], dtype, op)
        got, ref = _run(a, (0, 2), op, cuda, identity=None, out_init=init)
        _check(got, ref, dtype, a, (0, 2), op)


def test_two_pass_bitwise_views_and_bf16(cuda):
    rng = np.random.default_rng(4)
    a = rng.integers(0, 2**31 - 1, (9, 40, 77)).astype(np.int32)
    for op, ident in (('band', -1), ('bor', 0), ('bxor', 0)):
        got, ref = _run(a, (0, 2), op, cuda, identity=ident)
        np.testing.assert_array_equal(got, ref)
    # merged kept dims / 4-D: [N, C1, C2, HW] over (0, 3) canonicalises to [R1, K, R2]
    b = rng.random((5, 4, 6, 130)).astype(np.float32)
    got, ref = _run(b, (0, 3), 'sum', cuda)
    _check(got, ref, np.float32, b, (0, 3))
    # padded rows (row stride > R2) still qualify; a transposed outer pair does not (generic kernel) - same numbers
    big = torch.from_numpy(rng.random((7, 12, 200)).astype(np.float32)).to(cuda)
    view = big[:, :, :150]
    got = db.reduce('lambda a, b: a + b', view, axis=(0, 2), identity=0).cpu().numpy()
    np.testing.assert_allclose(got, view.cpu().numpy().astype(np.float64).sum((0, 2)), rtol=1e-5)
    tr = big.permute(1, 0, 2)                      # [12, 7, 200] with swapped strides, reduce (0, 2)
    got = db.reduce('lambda a, b: a + b', tr, axis=(0, 2), identity=0).cpu().numpy()
    np.testing.assert_allclose(got, tr.cpu().numpy().astype(np.float64).sum((0, 2)), rtol=1e-5)
    # bf16 -> f32 output: fp32 scratch between the passes
    x = torch.randn(16, 24, 512, device=cuda).to(torch.bfloat16)
    out = torch.zeros(24, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: a + b', x, out, axis=(0, 2), identity=0)
    ref = x.double().sum((0, 2)).cpu().numpy()
    scale = x.double().abs().sum((0, 2)).cpu().numpy()
    assert np.all(np.abs(out.cpu().numpy() - ref) <= 1e-5 * scale)
    mx = torch.zeros(24, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: max(a, b)', x, mx, axis=(0, 2), identity=-3e38)
    assert torch.equal(mx, x.float().amax((0, 2)))


def test_two_pass_two_streams_share_the_scratch(cuda):
    rng = np.random.default_rng(5)
    a = torch.from_numpy(rng.integers(-100, 100, (40, 300, 500)).astype(np.int32)).to(cuda)
    b = torch.from_numpy(rng.integers(-100, 100, (10, 2000, 64)).astype(np.int32)).to(cuda)
    oa, ob = torch.zeros(300, dtype=torch.int32, device=cuda), torch.zeros(2000, dtype=torch.int32, device=cuda)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    for _ in range(5):
        with torch.cuda.stream(s1):
            db.reduce('lambda a, b: a + b', a, oa, axis=(0, 2), identity=0)
        with torch.cuda.stream(s2):
            db.reduce('lambda a, b: a + b', b, ob, axis=(0, 2), identity=0)
    torch.cuda.synchronize()
    assert torch.equal(oa, a.sum((0, 2)).to(torch.int32)) and torch.equal(ob, b.sum((0, 2)).to(torch.int32))


# ---- permuted / transposed views: axes are put in memory order before a kernel is picked ----

@pytest.mark.parametrize('dtype', [np.float32, np.int64])
@pytest.mark.parametrize('perm,axes', [((1, 0), (1, )), ((1, 0), (0, )), ((2, 0, 1), (0, )), ((2, 0, 1), (1, 2)),
                                       ((0, 2, 1), (1, )), ((0, 3, 1, 2), (0, 2, 3)), ((0, 3, 1, 2), (1, )),
                                       ((3, 2, 1, 0), (0, 2)), ((1, 0, 2), None)])
def test_permuted_views_match_the_contiguous_result(dtype, perm, axes, cuda):
    """np.sum(A.T, axis=1), channels-last tensors reduced over (N, H, W), arbitrary permutations: the view is
    passed as it is (strides, no copy); results must equal the reduction of a contiguous copy."""
    rng = np.random.default_rng(6)
    base_shape = [37, 41, 130, 9][:len(perm)] if len(perm) < 4 else [6, 37, 41, 130]
    base = rng.integers(-50, 50, base_shape).astype(dtype) if dtype == np.int64 else rng.random(base_shape).astype(dtype)
    view = torch.from_numpy(base).to(cuda).permute(*perm)
    assert not view.is_contiguous()
    host = np.ascontiguousarray(np.transpose(base, perm))
    for op in ('sum', 'max'):
        red, lam, _ = _OPS[op]
        ident = _identity(op, np.dtype(dtype))
        got = db.reduce(lam, view, axis=axes, identity=ident).cpu().numpy()
        ref = O.reduce(red, host, axes, identity=ident, wide=True)
        _check(got, ref, dtype, host, axes, op)
    # strided output view as well (transposed destination)
    if axes == (1, ) and len(perm) == 3:
        out = torch.zeros(host.shape[2], host.shape[0], dtype=view.dtype, device=cuda).t()
        db.reduce('lambda a, b: a + b', view, out, axis=axes, identity=0)
        _check(out.cpu().numpy(), O.reduce(O.SUM, host, axes, identity=0, wide=True), dtype, host, axes, 'sum')


@pytest.mark.parametrize('tdt', ['bfloat16', 'float16'])
def test_two_pass_sixteen_bit_outputs(tdt, cuda):
    """[R1, K, R2] with a 16-bit float output: fp32 between the passes, one rounding at the end, identity folded
    once (given, or the value already in the output)."""
    dt = getattr(torch, tdt)
    eps = 2.0 ** -8 if tdt == 'bfloat16' else 2.0 ** -11
    _two_pass_sixteen_bit(torch.rand(12, 40, 3000, device=cuda), dt, eps)
    _two_pass_sixteen_bit(torch.rand(3000, 40, 12, device=cuda), dt, eps)      # long outer axis: folded first


def _two_pass_sixteen_bit(u, dt, eps):
    x = (u - 0.3).to(dt)
    truth = x.double().sum((0, 2))
    scale = x.double().abs().sum((0, 2))
    got = db.reduce('lambda a, b: a + b', x, axis=(0, 2), identity=0)
    assert got.dtype == dt
    assert bool(((got.double() - truth).abs() <= eps * truth.abs() + 1e-5 * scale).all())
    out = torch.full((40, ), 2.0, device=x.device, dtype=dt)
    db.reduce('lambda a, b: a + b', x, out, axis=(0, 2), identity=None)
    assert bool(((out.double() - (truth + 2.0)).abs() <= eps * (truth + 2.0).abs() + 1e-5 * scale).all())
    mx = db.reduce('lambda a, b: max(a, b)', x, axis=(0, 2), identity=-3e38)
    assert torch.equal(mx, x.amax((0, 2)))
    mn = db.reduce('lambda a, b: min(a, b)', x, axis=(0, 2), identity=0.25)
    assert torch.equal(mn.float(), torch.clamp(x.amin((0, 2)).float(), max=0.25))


@pytest.mark.parametrize('dtype', [np.float32, np.int64])
@pytest.mark.parametrize('shape,axes', [([5, 7, 9, 130], (0, 2)), ([6, 5, 8, 300], (1, 3)), ([3, 4, 5, 6, 64], (1, 3)),
                                        ([2, 3, 4, 5, 6, 40], (0, 2, 4)), ([64, 3, 50, 33], (0, 2)), ([65, 3, 50, 33], (0, 2))])
def test_interleaved_axes_host_peel(dtype, shape, axes, cuda):
    """[R1, K1, R2, K2], [K1, R1, K2, R2], ...: the outermost axis is peeled on the host (<= 64 slices) and each slice
    goes to a streaming kernel; a reduced outer axis folds slice after slice, so the identity must count once."""
    rng = np.random.default_rng(8)
    a = _data(rng, shape, dtype, 'sum')
    for op in ('sum', 'max'):
        got, ref = _run(a, axes, op, cuda)
        _check(got, ref, dtype, a, axes, op)
    kept = [s for i, s in enumerate(shape) if i not in axes]
    init = _data(rng, kept, dtype, 'sum')
    got, ref = _run(a, axes, 'sum', cuda, identity=None, out_init=init)       # fold into the existing output
    _check(got, ref, dtype, a, axes, 'sum')
    got, ref = _run(a, axes, 'sum', cuda, identity=3)                          # a non-neutral identity, once
    _check(got, ref, dtype, a, axes, 'sum')


@pytest.mark.parametrize('dtype', [np.float32, np.int32, np.int64, np.uint8, np.float64])
@pytest.mark.parametrize('shape,axes', [
    ([100, 7, 9, 130], (0, 2)),          # [R1, K1, R2, K2] with an outer axis too long for the host peel
    ([70, 5, 8, 3], (0, 2)),             # tiny innermost kept axis
    ([9, 6, 7, 8, 4], (1, 3)),           # [K0, R1, K2, R3, K4]
    ([66, 3, 4, 5, 6, 20], (0, 2, 4)),   # three reduced axes: two intermediates
    ([70, 6, 5, 40], (1, 3)),            # [K0, R1, K2, R3]: innermost axis reduced (rows first), long outer kept axis
    ([3, 2, 2, 2, 2, 2, 2, 33], (0, 2, 4, 6)),
])
def test_interleaved_axes_inner_first(dtype, shape, axes, cuda):
    """Interleaved kept / reduced axes without a direct kernel: the innermost reduced axis is folded first into a
    contiguous intermediate (all outer axes as one index), the rest is planned on the intermediate - one streaming
    pass per reduced axis group instead of the strided kernel.  Identity counts once; identity None folds into the
    output; integers bit-exact."""
    rng = np.random.default_rng(18)
    a = _data(rng, shape, dtype, 'sum')
    for op in ('sum', 'max', 'min'):
        got, ref = _run(a, axes, op, cuda)
        _check(got, ref, dtype, a, axes, op)
    kept = [s for i, s in enumerate(shape) if i not in axes]
    init = _data(rng, kept, dtype, 'sum')
    got, ref = _run(a, axes, 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, axes, 'sum')
    got, ref = _run(a, axes, 'sum', cuda, identity=3)
    _check(got, ref, dtype, a, axes, 'sum')


@pytest.mark.parametrize('dtype', [np.float32, np.int32, np.int64, np.uint8, np.float64])
@pytest.mark.parametrize('shape,axes', [([20000, 16, 8], (1, )), ([10000, 100, 2], (1, )), ([12345, 7, 32], (1, )),
                                        ([9500, 511, 4], (1, )), ([40, 6, 50, 8, 4], (1, 3)), ([10000, 3, 16], (1, ))])
def test_many_small_dense_blocks(dtype, shape, axes, cuda):
    """[O, R, K] with K a power of two <= 32 and small R * K: a warp streams one block as a flat run
    (cols_small_kernel); also as the first pass of an inner-first plan."""
    rng = np.random.default_rng(19)
    a = _data(rng, shape, dtype, 'sum')
    for op in ('sum', 'max'):
        got, ref = _run(a, axes, op, cuda)
        _check(got, ref, dtype, a, axes, op)
    kept = [s for i, s in enumerate(shape) if i not in axes]
    init = _data(rng, kept, dtype, 'sum')
    got, ref = _run(a, axes, 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, axes, 'sum')
    # unaligned start (a view one element into the allocation) and a strided output
    flat = np.concatenate([np.zeros(1, dtype), a.reshape(-1)])
    tv = torch.from_numpy(flat).to(cuda)[1:].view(*shape)
    out = torch.zeros([2 * k for k in kept], dtype=tv.dtype, device=cuda)[tuple(slice(None, None, 2) for _ in kept)]
    db.reduce('lambda a, b: a + b', tv, out, axis=axes, identity=0)
    _check(out.cpu().numpy(), O.reduce(O.SUM, a, axes, identity=0, wide=True), dtype, a, axes, 'sum')


def test_inner_first_views_bf16_and_workspace_query(cuda):
    """The same plan on a bf16 input (fp32 intermediates), on a strided output, with a base-only workspace (falls
    back to the scratch-free kernels, same result) and the scratch the query asks for."""
    import ctypes
    from dace_b200 import _lib
    from dace_b200.api import Workspace
    lib = _lib.load()
    g = torch.Generator(device=cuda)
    g.manual_seed(21)
    x = torch.randn(80, 6, 12, 96, device=cuda, generator=g).to(torch.bfloat16)
    out = torch.empty(6, 96, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: a + b', x, out, axis=(0, 2), identity=0)
    ref = x.double().sum((0, 2))
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=1e-3)
    mx = torch.empty(6, 96, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: max(a, b)', x, mx, axis=(0, 2), identity=-3e38)
    assert torch.equal(mx, x.float().amax((0, 2)))
    xi = torch.randint(-50, 50, (90, 5, 11, 70), dtype=torch.int32, device=cuda, generator=g)
    want = xi.sum((0, 2), dtype=torch.int32)
    shape = (ctypes.c_int64 * 4)(*xi.shape)
    istr = (ctypes.c_int64 * 4)(*xi.stride())
    need = lib.dace_b200_reduce_nd_workspace_bytes(2, 2, 4, shape, istr, 0b0101)
    base = lib.dace_b200_workspace_bytes()
    assert need >= base + 90 * 5 * 70 * 4          # the first intermediate [R1, K1, K2]
    for ws in (Workspace(cuda), Workspace(cuda, nbytes=need)):
        o = torch.zeros(70, 5, dtype=torch.int32, device=cuda).t()          # transposed destination
        db.reduce('lambda a, b: a + b', xi, o, axis=(0, 2), identity=0, workspace=ws)
        assert torch.equal(o, want)
    # a padded view (outer axes not dense): no intermediate plan, still exact
    big = torch.randint(-50, 50, (90, 5, 11, 80), dtype=torch.int32, device=cuda, generator=g)
    v = big[:, :, :, :70]
    assert torch.equal(db.reduce('lambda a, b: a + b', v, axis=(0, 2), identity=0), v.sum((0, 2), dtype=torch.int32))


@pytest.mark.parametrize('dtype', [np.float32, np.int32, np.float64])
@pytest.mark.parametrize('shape', [[6, 5, 8, 300], [3, 70, 4, 33], [2, 2, 2, 2], [40, 3, 100, 1000]])
def test_outer_kept_two_pass(dtype, shape, cuda):
    """[O, R1, K, R2] over axes (1, 3) (per-sample, per-channel statistics): one rows pass over O*R1*K runs and one
    column pass with O as the outer dimension."""
    rng = np.random.default_rng(9)
    a = _data(rng, shape, dtype, 'sum')
    for op in ('sum', 'min'):
        got, ref = _run(a, (1, 3), op, cuda)
        _check(got, ref, dtype, a, (1, 3), op)
    init = _data(rng, [shape[0], shape[2]], dtype, 'sum')
    got, ref = _run(a, (1, 3), 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, (1, 3), 'sum')
    # strided output (a transposed destination)
    ta = torch.from_numpy(a).to(cuda)
    out = torch.zeros(shape[2], shape[0], dtype=ta.dtype, device=cuda).t()
    db.reduce('lambda a, b: a + b', ta, out, axis=(1, 3), identity=0)
    _check(out.cpu().numpy(), O.reduce(O.SUM, a, (1, 3), identity=0, wide=True), dtype, a, (1, 3), 'sum')


@pytest.mark.parametrize('K', [2, 3, 17, 100, 255, 256, 257, 1000, 1024])
@pytest.mark.parametrize('dtype', [np.float32, np.int32, np.float64])
def test_narrow_dense_column_folds(K, dtype, cuda):
    """[R, K <= 1024] dense, axis 0: the flat-stream kernel (a thread's column never changes); deterministic."""
    rng = np.random.default_rng(K)
    R = (1 << 22) // K + 37
    a = _data(rng, [R, K], dtype, 'sum')
    for op in ('sum', 'max'):
        got, ref = _run(a, (0, ), op, cuda)
        _check(got, ref, dtype, a, (0, ), op)
    init = _data(rng, [K], dtype, 'sum')
    got, ref = _run(a, (0, ), 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, (0, ), 'sum')
    ta = torch.from_numpy(a).to(cuda)
    o1 = db.reduce('lambda a, b: a + b', ta, axis=(0, ), identity=0)
    o2 = db.reduce('lambda a, b: a + b', ta, axis=(0, ), identity=0)
    assert torch.equal(o1, o2)


@pytest.mark.parametrize('K', [2, 4, 6, 8, 100, 128, 250, 256, 1000, 1024])
@pytest.mark.parametrize('dtype', [np.uint8, np.int8, np.int16])
def test_narrow_dense_column_folds_small_integer_types(K, dtype, cuda):
    """1- and 2-byte elements: a thread takes a 32-bit word of adjacent columns when K is a multiple of the word
    (4 / 2 elements) and the base is word-aligned, else one element - bit-exact either way, also for a view that
    starts one element into the allocation."""
    rng = np.random.default_rng(K + 1)
    R = (1 << 22) // K + 37
    a = _data(rng, [R, K], dtype, 'sum')
    for op in ('sum', 'max', 'min'):
        got, ref = _run(a, (0, ), op, cuda)
        np.testing.assert_array_equal(got, ref)
    buf = torch.from_numpy(np.concatenate([a.reshape(-1)[:1], a.reshape(-1)])).to(cuda)
    v = buf[1:].view(R, K)                                         # same data, base off by one element
    got = db.reduce('lambda a, b: max(a, b)', v, axis=(0, ), identity=int(np.iinfo(dtype).min))
    np.testing.assert_array_equal(got.cpu().numpy(), a.max(0))


@pytest.mark.parametrize('K', [2, 6, 128, 250, 1024])
@pytest.mark.parametrize('tdt', ['bfloat16', 'float16'])
def test_narrow_dense_column_folds_half_types(K, tdt, cuda):
    dt = getattr(torch, tdt)
    eps = 2.0 ** -8 if tdt == 'bfloat16' else 2.0 ** -11
    R = (1 << 22) // K + 37
    x = (torch.rand(R, K, device=cuda) - 0.3).to(dt)
    out = torch.zeros(K, dtype=torch.float32, device=cuda)
    db.reduce('lambda a, b: a + b', x, out, axis=(0, ), identity=0)             # fp32 output: no final rounding
    ref, scale = x.double().sum(0), x.double().abs().sum(0)
    assert bool(((out.double() - ref).abs() <= 1e-5 * scale).all())
    if float(ref.abs().max()) < 3e4:                                            # (fp16 holds nothing past 65504)
        got = db.reduce('lambda a, b: a + b', x, axis=(0, ), identity=0)        # 16-bit output: one rounding
        assert bool(((got.double() - ref).abs() <= eps * ref.abs() + 1e-5 * scale).all())
    assert torch.equal(db.reduce('lambda a, b: max(a, b)', x, axis=(0, ), identity=-60000.0), x.amax(0))


@pytest.mark.parametrize('n', [2, 3, 5, 8, 16, 17, 33, 100, 128, 129, 250, 333, 511, 512, 513])   # > 128: the lanes-per-row kernel, rows not whole 32-byte vectors
@pytest.mark.parametrize('dtype', [np.float32, np.int64, np.uint8, np.float64])
def test_tiny_dense_rows(n, dtype, cuda):
    """[R, n <= 128] dense, axis 1: rows staged through shared memory as one flat tile per CTA."""
    rng = np.random.default_rng(n)
    R = (1 << 22) // n + 41
    a = _data(rng, [R, n], dtype, 'sum')
    for op in ('sum', 'max'):
        got, ref = _run(a, (1, ), op, cuda)
        _check(got, ref, dtype, a, (1, ), op)
    init = _data(rng, [R], dtype, 'sum')
    got, ref = _run(a, (1, ), 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, (1, ), 'sum')


def test_tiny_rows_bf16_to_f32_and_strided_out(cuda):
    x = torch.randn((1 << 22) // 48 + 3, 48, device=cuda).to(torch.bfloat16)
    out = torch.zeros(x.shape[0] * 2, dtype=torch.float32, device=cuda)[::2]
    db.reduce('lambda a, b: a + b', x, out, axis=1, identity=0)
    ref = x.double().sum(1)
    scale = x.double().abs().sum(1)
    assert bool(((out.double() - ref).abs() <= 1e-5 * scale + 1e-12).all())
    mx = db.reduce('lambda a, b: max(a, b)', x, axis=1, identity=-3e38)
    assert torch.equal(mx, x.amax(1))


def test_signed_zero_max_is_a_zero(cuda):
    """Float max / min use fmaxf / fminf where the reference uses `(a > b) ? a : b` (reference
    dace/runtime/include/dace/math.h:51-55).  On NaN-free data they agree except for the SIGN of a zero result,
    where the reference itself depends on the fold order (max(-0.0, +0.0) keeps whichever came second).  The
    decision: the value is pinned (a zero), the sign is not."""
    x = torch.tensor([-0.0, 0.0] * 5000 + [-0.0], device=cuda)
    for lam, ident in (('lambda a, b: max(a, b)', -1.0), ('lambda a, b: min(a, b)', 1.0)):
        got = db.reduce(lam, x, axis=None, identity=ident)
        assert got.item() == 0.0
        rows = db.reduce(lam, x[:10000].view(100, 100), axis=1, identity=ident)
        assert torch.all(rows == 0.0)
    neg = torch.full((4097, ), -0.0, device=cuda)
    got = db.reduce('lambda a, b: max(a, b)', neg, axis=None, identity=-1.0)
    assert got.item() == 0.0 and torch.signbit(got).item()      # all inputs -0.0: no choice to make


@pytest.mark.parametrize('dtype', [np.float32, np.int32, np.float64, np.int16])
def test_all_elements_of_padded_and_strided_views(dtype, cuda):
    """Every element of a window of a larger array (a pitch between the rows): rows into an intermediate, then one
    contiguous run (api.cu plan_inner_first with nothing kept) - not one warp of the strided kernel; also a 3-D
    window padded on two levels, sum and max, identity None."""
    rng = np.random.default_rng(21)
    base = _data(rng, [3000, 2500], dtype)
    tb = torch.from_numpy(base).to(cuda)
    for view, host in ((tb[:, :2400], base[:, :2400]), (tb[5:2000, 17:999], base[5:2000, 17:999]), (tb[::2, :2400], base[::2, :2400])):
        for op, lam, red in (('sum', 'lambda a, b: a + b', O.SUM), ('max', 'lambda a, b: max(a, b)', O.MAX)):
            out = torch.zeros((), dtype=tb.dtype, device=cuda)
            ident = 0 if op == 'sum' else _identity('max', dtype)
            db.reduce(lam, view, out, axis=None, identity=ident)
            ref = O.reduce(red, np.ascontiguousarray(host), None, identity=ident, wide=True)
            _check(out.cpu().numpy().reshape(ref.shape), ref, dtype, np.ascontiguousarray(host), None, op)
    cube = _data(rng, [40, 300, 500], dtype)
    tc = torch.from_numpy(cube).to(cuda)
    v, h = tc[:, :250, :450], np.ascontiguousarray(cube[:, :250, :450])
    out = torch.zeros((), dtype=tc.dtype, device=cuda)
    db.reduce('lambda a, b: a + b', v, out, axis=None, identity=0)
    _check(out.cpu().numpy().reshape(()), O.reduce(O.SUM, h, None, identity=0, wide=True), dtype, h, None, 'sum')


@pytest.mark.parametrize('axes', [(0, ), (1, ), (2, ), (0, 1), (0, 2), (1, 2), (0, 1, 2)])
@pytest.mark.parametrize('dtype', [np.float32, np.int64, np.int16])
def test_windows_of_a_larger_3d_array(axes, dtype, cuda):
    """[A, B, C] windows with a pitch on two levels (stencil interiors, sub-blocks): every axis combination; the ones
    that fold the innermost axis run the row kernels on two levels of rows (RowsArgs::hi_*), directly or as the
    first pass of the inner-first plan.  Sum and max, identity None, long and short innermost axes."""
    rng = np.random.default_rng(sum(axes) + 7)
    for full, win in (((60, 70, 900), (slice(2, 58), slice(3, 66), slice(5, 880))),
                      ((300, 40, 64), (slice(0, 300), slice(0, 37), slice(0, 60))),
                      ((5, 600, 4100), (slice(1, 5), slice(0, 590), slice(4, 4100)))):
        base = _data(rng, list(full), dtype)
        host = np.ascontiguousarray(base[win])
        view = torch.from_numpy(base).to(cuda)[win]
        for op in ('sum', 'max'):
            got, ref = _run(host, axes, op, cuda)                 # the oracle on the contiguous copy
            red, lam, _ = _OPS[op]
            ident = _identity(op, dtype)
            out = db.reduce(lam, view, axis=axes, identity=ident)
            _check(out.cpu().numpy().reshape(ref.shape), ref, dtype, host, axes, op)
        kept = [s for i, s in enumerate(host.shape) if i not in axes] or [1]
        init = _data(rng, kept, dtype)
        out = torch.from_numpy(init.copy()).to(cuda).reshape(kept if len(axes) < 3 else ())
        db.reduce('lambda a, b: a + b', view, out, axis=axes, identity=None)
        ref = O.reduce(O.SUM, host, axes, identity=None, out_init=init.reshape(out.shape if out.dim() else ()).copy(), wide=True) \
            if False else None
        want = host.astype(np.float64).sum(axes) + init.reshape(np.shape(host.astype(np.float64).sum(axes))).astype(np.float64)
        if np.issubdtype(dtype, np.floating):
            scale = np.abs(host.astype(np.float64)).sum(axes) + np.abs(init.astype(np.float64)).reshape(want.shape)
            assert np.all(np.abs(out.cpu().numpy().astype(np.float64).reshape(want.shape) - want) <= RTOL_F32 * np.maximum(scale, 1e-30))
        else:
            np.testing.assert_array_equal(out.cpu().numpy().reshape(want.shape), want.astype(dtype))


@pytest.mark.parametrize('axes', [(2, 3), (0, 2, 3), (3, ), (1, 2, 3), (0, 1, 2, 3), (0, 1), (1, ), (0, 3)])
@pytest.mark.parametrize('dtype', [np.float32, np.int32])
def test_interior_of_a_4d_array(axes, dtype, cuda):
    """[N, C, H, W] interiors `A[:, :, 1:-1, 1:-1]` (per-image, per-channel and global statistics of a stencil
    domain): N and C stay dense, a pitch after every row and after every image plane."""
    rng = np.random.default_rng(len(axes) * 3 + axes[0])
    base = _data(rng, [6, 10, 70, 300], dtype)
    host = np.ascontiguousarray(base[:, :, 1:-1, 1:-1])
    view = torch.from_numpy(base).to(cuda)[:, :, 1:-1, 1:-1]
    for op in ('sum', 'min'):
        red, lam, _ = _OPS[op]
        ident = _identity(op, dtype)
        out = db.reduce(lam, view, axis=axes, identity=ident)
        ref = O.reduce(red, host, axes, identity=ident, wide=True)
        _check(out.cpu().numpy().reshape(ref.shape), ref, dtype, host, axes, op)


@pytest.mark.parametrize('dtype', [np.float32, np.int32])
def test_window_padded_on_three_levels(dtype, cuda):
    """[A, B, C, D] windows with a pitch after every row, plane and cube: the planner peels the outermost axis and
    plans each slice with everything it has (rows on two levels, inner-first), instead of giving the whole view to
    the strided kernel.  Slices above the peel thresholds (1 MiB folded, 32 MiB kept)."""
    rng = np.random.default_rng(11)
    base = _data(rng, [4, 54, 64, 3072], dtype)
    win = (slice(0, 4), slice(2, 52), slice(1, 61), slice(8, 3008))
    host = np.ascontiguousarray(base[win])
    view = torch.from_numpy(base).to(cuda)[win]
    for axes in ((0, 1, 2, 3), (1, 2, 3), (1, 3), (2, 3), (0, 1, 3), (0, 2, 3), (1, ), (0, 1), (3, )):
        for op in ('sum', 'max'):
            red, lam, _ = _OPS[op]
            ident = _identity(op, dtype)
            out = db.reduce(lam, view, axis=axes, identity=ident)
            ref = O.reduce(red, host, axes, identity=ident, wide=True)
            _check(out.cpu().numpy().reshape(ref.shape), ref, dtype, host, axes, op)
    # identity None: folded into what the output holds, once, also when the folded outermost axis is peeled
    want = host.astype(np.float64).sum((0, 1, 3))
    init = _data(rng, [60], dtype)
    out = torch.from_numpy(init.copy()).to(cuda)
    db.reduce('lambda a, b: a + b', view, out, axis=(0, 1, 3), identity=None)
    want = want + init.astype(np.float64)
    if np.issubdtype(dtype, np.floating):
        scale = np.abs(host.astype(np.float64)).sum((0, 1, 3)) + np.abs(init.astype(np.float64))
        assert np.all(np.abs(out.cpu().numpy().astype(np.float64) - want) <= RTOL_F32 * np.maximum(scale, 1e-30))
    else:
        np.testing.assert_array_equal(out.cpu().numpy(), want.astype(np.int64).astype(dtype))


@pytest.mark.parametrize('tdtype', ['bfloat16', 'float16'])
def test_windows_with_16_bit_output(tdtype, cuda):
    """bf16 -> bf16 / f16 -> f16 over windows: the multi-pass plans keep fp32 between their passes and there is no
    fp32 -> 16-bit kernel pair, so the last pass writes fp32 behind the intermediate and one small launch folds the
    identity and rounds once.  (Before: these layouts fell to one warp of the strided kernel.)  Exact small integers
    whose sums stay exactly representable in fp32; the result is the correctly rounded 16-bit value."""
    dt = getattr(torch, tdtype)
    g = torch.Generator(device='cuda').manual_seed(5)
    base = torch.randint(-3, 4, (60, 70, 900), device=cuda, generator=g).to(dt)
    view = base[2:58, 3:66, 5:880]
    for axes in ((0, 1, 2), (0, 2), (1, 2), (0, 1), (2, ), (1, ), (0, )):
        ref64 = view.double().sum(axes)
        out = db.reduce('lambda a, b: a + b', view, axis=axes, identity=0)
        assert out.dtype == dt
        assert torch.equal(out.reshape(ref64.shape), ref64.to(dt)), axes
        mref = view.float().amax(axes).to(dt)
        out = db.reduce('lambda a, b: max(a, b)', view, axis=axes, identity=-100)
        assert torch.equal(out.reshape(mref.shape), mref), axes
    # identity None: folded into what the output holds
    init = torch.randint(-5, 6, (63, ), device=cuda, generator=g).to(dt)
    out = init.clone()
    db.reduce('lambda a, b: a + b', view, out, axis=(0, 2), identity=None)
    assert torch.equal(out, (view.double().sum((0, 2)) + init.double()).to(dt))
    # every element of a padded 2-D window
    base2 = torch.randint(-3, 4, (3000, 4096), device=cuda, generator=g).to(dt)
    out = db.reduce('lambda a, b: a + b', base2[:, :4000], axis=None, identity=0)
    assert torch.equal(out.reshape(()), base2[:, :4000].double().sum().to(dt))


@pytest.mark.parametrize('axes', [(0, ), (1, ), None])
def test_broadcast_views_stride_zero(axes, cuda):
    """Inputs with a zero stride (a row or a column broadcast over the other axis, `expand`): every element is
    counted as often as the view shows it."""
    g = torch.Generator(device='cuda').manual_seed(4)
    row = torch.randint(-20, 20, (1, 3000), device=cuda, dtype=torch.int32, generator=g)
    col = torch.randint(-20, 20, (2500, 1), device=cuda, dtype=torch.int32, generator=g)
    for x in (row.expand(2500, 3000), col.expand(2500, 3000), row.expand(2500, 3000)[::2, 5:2999]):
        ref = x.contiguous().sum(axes, dtype=torch.int32) if axes is not None else x.contiguous().sum(dtype=torch.int32)
        out = db.reduce('lambda a, b: a + b', x, axis=axes, identity=0)
        assert torch.equal(out.reshape(ref.shape), ref)
        mref = x.contiguous().amax(axes) if axes is not None else x.contiguous().max()
        out = db.reduce('lambda a, b: max(a, b)', x, axis=axes, identity=-1000)
        assert torch.equal(out.reshape(mref.shape), mref)


@pytest.mark.parametrize('shape', [(3000, 3, 700), (200000, 4, 64), (70000, 2, 33), (5000, 16, 300), (64, 5, 9000), (100000, 1, 40)])
@pytest.mark.parametrize('dtype', [np.float32, np.int64, np.int16])
def test_short_axis_between_many_slices(shape, dtype, cuda):
    """[N, R <= 16, M] over the middle axis (channel folds): the thread-per-output kernel; sum / max, identity None,
    a strided output, a window of a larger array."""
    rng = np.random.default_rng(shape[0] % 97)
    a = _data(rng, list(shape), dtype)
    for op in ('sum', 'max'):
        got, ref = _run(a, (1, ), op, cuda)
        _check(got, ref, dtype, a, (1, ), op)
    init = _data(rng, [shape[0], shape[2]], dtype)
    got, ref = _run(a, (1, ), 'sum', cuda, identity=None, out_init=init)
    _check(got, ref, dtype, a, (1, ), 'sum')
    ta = torch.from_numpy(a).to(cuda)
    out = torch.zeros(shape[2], shape[0], dtype=ta.dtype, device=cuda).t()          # transposed destination
    db.reduce('lambda a, b: a + b', ta, out, axis=(1, ), identity=0)
    _check(out.cpu().numpy(), O.reduce(O.SUM, a, (1, ), identity=0, wide=True), dtype, a, (1, ), 'sum')
    if shape[2] > 40:
        w = ta[:, :, 3:shape[2] - 5]
        out = db.reduce('lambda a, b: a + b', w, axis=(1, ), identity=0)
        h = np.ascontiguousarray(a[:, :, 3:shape[2] - 5])
        _check(out.cpu().numpy(), O.reduce(O.SUM, h, (1, ), identity=0, wide=True), dtype, h, (1, ), 'sum')
