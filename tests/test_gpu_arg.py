"""GPU parity: argmin / argmax (Min_Location / Max_Location) through the C ABI against the oracle
and the reference's own golden vectors.

Cases follow reference tests/numpy/reductions_test.py:58-84 (argmax / argmin along axis 1 of
[10, 5, 3] for floats and integers) and tests/numpy/ndarray_attributes_methods_test.py:95-111
(flattened), plus the layouts the kernels distinguish: short / long contiguous rows, one huge row
split over the machine, strided reduced axis with and without vector loads, [outer, reduced, inner].
Bar: indices and values bit-exact (first occurrence wins; (init, -1) start pair).
"""
import json
import os
import zlib

import numpy as np
import pytest
import torch

import dace_b200 as db
from oracle import oracle as O

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden')
ARG_META = json.load(open(os.path.join(GOLDEN, 'arg_golden_meta.json')))['arg']
ARG = np.load(os.path.join(GOLDEN, 'arg_golden.npz'))

_FN = {'max': (db.argmax, O.MAX_LOCATION), 'min': (db.argmin, O.MIN_LOCATION)}


def _dev(a):
    return torch.from_numpy(a).cuda()


@pytest.mark.parametrize('m', ARG_META, ids=[m['tag'] for m in ARG_META])
def test_golden_argminmax(m):
    a, expected = ARG[m['tag'] + '__in'], ARG[m['tag'] + '__out']
    fn, _ = _FN[m['func']]
    got = fn(_dev(a), axis=m['axis']).cpu().numpy()
    np.testing.assert_array_equal(got.reshape(expected.shape), expected)


def _data(rng, shape, dtype, ties):
    if np.issubdtype(dtype, np.floating):
        a = rng.uniform(-10, 10, shape).astype(dtype)
        if ties:
            a = np.round(a)          # many repeated values -> first occurrence matters
        return a
    info = np.iinfo(dtype)
    # stay off the type's extreme values: an element equal to the start value (min_value / max_value of the
    # type) never replaces the (init, -1) pair in the reference, which NumPy's argmin/argmax do not model
    if ties:
        return rng.integers(max(info.min + 1, -4), min(info.max - 1, 4) + 1, shape).astype(dtype)
    return rng.integers(max(info.min + 1, -30000), min(info.max - 1, 30000) + 1, shape).astype(dtype)


_SHAPES = [
    ((10, 5, 3), 1), ((10, 5, 3), 0), ((10, 5, 3), 2), ((7, 9, 11), None),
    ((1000, 37), 1),            # short ragged rows, G lanes per row
    ((513, 1000), 1),           # 4 KB rows
    ((40, 20000), 1),           # long rows, CTA per row
    ((3, 1 << 20), 1),          # few long rows -> row splits + ticket
    ((1 << 22) + 13, None),     # one row over the whole machine, unaligned length
    ((5000, 256), 0),           # strided reduced axis, vector columns
    ((100000, 24), 0),          # narrow, many rows -> splits
    ((3000, 77), 0),            # columns not a vector multiple -> scalar path
    ((6, 2000, 128), 1),        # [outer, reduced, inner]
    ((4, 300, 5), 1),
    ((8, 200000), 0),           # few rows, many columns: a thread per column vector, all rows
    ((3, 5, 77001), 1),         # the same with an outer axis and a width that is not a vector multiple
    ((64, 40004), 0),
    ((5000, 5000), 1),          # rows longer than 32 KiB (8-byte types) but thousands of them: still a warp per row
    ((1, 1), 1), ((5, 1), 1), ((1, 9), 0),
]


@pytest.mark.parametrize('func', ['max', 'min'])
@pytest.mark.parametrize('dtype', [np.float32, np.float64, np.int32, np.int64, np.uint32, np.uint64, np.int8,
                                   np.uint8, np.int16], ids=lambda d: np.dtype(d).name)
@pytest.mark.parametrize('ties', [False, True], ids=['spread', 'ties'])
def test_arg_matches_oracle(func, dtype, ties):
    rng = np.random.default_rng(zlib.crc32(('%s %s %s' % (func, np.dtype(dtype).name, ties)).encode()))
    fn, red = _FN[func]
    for shape, axis in _SHAPES:
        if isinstance(shape, int):
            shape = (shape, )
        a = _data(rng, shape, dtype, ties)
        ta = _dev(a)
        ov = torch.empty([s for i, s in enumerate(a.shape) if axis is not None and i != axis % a.ndim],
                         dtype=ta.dtype, device='cuda')
        for idx_dtype in (torch.int32, torch.int64):
            got = fn(ta, axis=axis, out_val=ov, idx_dtype=idx_dtype)
            ridx, rval = O.reduce_arg(red, a, axis)
            np.testing.assert_array_equal(got.cpu().numpy().reshape(ridx.shape), ridx,
                                          err_msg='%s %s axis=%s' % (func, shape, axis))
            np.testing.assert_array_equal(ov.cpu().numpy().reshape(rval.shape), rval)
            fnp = np.argmax if func == 'max' else np.argmin
            np.testing.assert_array_equal(got.cpu().numpy().reshape(ridx.shape), fnp(a, axis=axis))


@pytest.mark.parametrize('func', ['max', 'min'])
@pytest.mark.parametrize('tdt', [torch.bfloat16, torch.float16], ids=['bf16', 'f16'])
def test_arg_half_types(func, tdt):
    fn, red = _FN[func]
    g = torch.Generator(device='cpu').manual_seed(3)
    for shape, axis in [((300, 4096), 1), ((4096, 320), 0), ((1 << 21, ), None), ((5, 700, 16), 1)]:
        t = torch.randn(shape, generator=g).to(tdt)
        bits = t.view(torch.int16).numpy().view(np.uint16)
        if tdt == torch.float16:
            a, bf = bits.view(np.float16), False
        else:
            a, bf = bits, True
        ridx, rval = O.reduce_arg(red, a, axis, in_bf16=bf)
        ov = torch.empty(ridx.shape, dtype=tdt, device='cuda')
        got = fn(t.cuda(), axis=axis, out_val=ov)
        np.testing.assert_array_equal(got.cpu().numpy().reshape(ridx.shape), ridx)
        np.testing.assert_array_equal(ov.view(torch.int16).cpu().numpy().view(np.uint16).reshape(rval.shape),
                                      rval.view(np.uint16))


def test_arg_fold_into_existing_pair_and_explicit_init():
    rng = np.random.default_rng(9)
    a = rng.integers(-50, 50, (64, 5000)).astype(np.int32)
    ta = _dev(a)
    idx = torch.empty(64, dtype=torch.int64, device='cuda')
    val = torch.empty(64, dtype=torch.int32, device='cuda')
    db.argmax(ta[:, :3000], axis=1, out_idx=idx, out_val=val)
    db.argmax(ta[:, 3000:], axis=1, out_idx=idx, out_val=val, init=None)
    i1, v1 = O.reduce_arg(O.MAX_LOCATION, a[:, :3000], 1, idx_dtype=np.int64)
    i2, v2 = O.reduce_arg(O.MAX_LOCATION, a[:, 3000:], 1, init=None, out_idx=i1, out_val=v1)
    np.testing.assert_array_equal(idx.cpu().numpy(), i2)
    np.testing.assert_array_equal(val.cpu().numpy(), v2)
    # explicit init above every element: nothing replaces the start pair -> -1 everywhere
    got = db.argmax(ta, axis=1, init=1000)
    assert (got.cpu().numpy() == -1).all()
    ridx, _ = O.reduce_arg(O.MAX_LOCATION, a, 1, init=10)
    np.testing.assert_array_equal(db.argmax(ta, axis=1, init=10).cpu().numpy(), ridx)


def test_arg_strided_views_and_offsets():
    rng = np.random.default_rng(21)
    a = rng.uniform(-1, 1, (300, 1111)).astype(np.float32)
    ta = _dev(a)
    # unaligned row starts (offset memlet), row stride larger than the range
    sub = ta[7:200, 3:1000]
    ridx, _ = O.reduce_arg(O.MIN_LOCATION, a[7:200, 3:1000], 1)
    np.testing.assert_array_equal(db.argmin(sub, axis=1).cpu().numpy(), ridx)
    ridx, _ = O.reduce_arg(O.MIN_LOCATION, a[7:200, 3:1000], 0)
    np.testing.assert_array_equal(db.argmin(sub, axis=0).cpu().numpy(), ridx)
    # transposed view: reduced axis has the large stride, kept axis is contiguous
    ridx, _ = O.reduce_arg(O.MAX_LOCATION, np.ascontiguousarray(a.T), 1)
    np.testing.assert_array_equal(db.argmax(ta.t(), axis=1).cpu().numpy(), ridx)


def test_arg_full_size_properties():
    """2^28 f32 (1 GiB): planted extremum found; determinism; value == max of the array."""
    n = 1 << 28
    g = torch.Generator(device='cuda').manual_seed(1)
    t = torch.rand(n, device='cuda', generator=g)
    pos = [123456789, 200000001]
    t[pos[0]] = 2.0
    t[pos[1]] = 2.0             # tie: the first one must win
    val = torch.empty((), dtype=torch.float32, device='cuda')
    i1 = db.argmax(t, out_val=val, idx_dtype=torch.int64)
    i2 = db.argmax(t, idx_dtype=torch.int64)
    assert int(i1) == pos[0] and int(i2) == pos[0] and float(val) == 2.0
    t[5] = -1.0
    assert int(db.argmin(t)) == 5


def test_arg_over_windows_of_larger_arrays():
    """Kept axes that merge into two dense groups: 3-D windows folded over their last / first / middle axis, the
    interior of a 4-D array over its last axis; three groups are refused."""
    g = torch.Generator(device='cuda').manual_seed(11)
    base = torch.rand(60, 70, 900, device='cuda', generator=g)
    w = base[2:58, 3:66, 5:880]
    for axis in (0, 1, 2):
        assert torch.equal(db.argmax(w, axis=axis).long(), w.argmax(dim=axis))
        assert torch.equal(db.argmin(w, axis=axis).long(), w.argmin(dim=axis))
    b4 = torch.rand(6, 10, 70, 300, device='cuda', generator=g)[:, :, 1:-1, 1:-1]
    assert torch.equal(db.argmax(b4, axis=3).long(), b4.argmax(dim=3))
    assert torch.equal(db.argmax(b4, axis=2).long(), b4.argmax(dim=2))
    with pytest.raises(NotImplementedError):
        db.argmax(b4, axis=0)
