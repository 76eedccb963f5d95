"""GPU: no kernel writes outside its output.  compute-sanitizer is closed on this GPU pool, so every kernel family is
run with its outputs embedded in a larger buffer of sentinel values (128 elements of guard on either side) and the
guards are compared afterwards; results are checked against torch on the way (integer data, bit-exact)."""
import pytest
import torch

import dace_b200 as db

pytestmark = pytest.mark.gpu
ADD, MAX = 'lambda a, b: a + b', 'lambda a, b: max(a, b)'
GUARD = 128


def _ints(shape, dt=torch.int32, lo=-50, hi=50, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    return torch.randint(lo, hi, shape, device='cuda', dtype=dt, generator=g)


def _guarded(numel, dtype, sentinel):
    buf = torch.full((numel + 2 * GUARD, ), sentinel, dtype=dtype, device='cuda')
    return buf, buf[GUARD:GUARD + numel]


def _guards_intact(buf, numel, sentinel):
    return bool((buf[:GUARD] == sentinel).all()) and bool((buf[GUARD + numel:] == sentinel).all())


REDUCE_CASES = [
    ('full', ((1 << 22) + 5, ), None), ('rows group', (4096, 4096), (1, )), ('rows odd', (5000, 1001), (1, )),
    ('rows block', (70, 40000), (1, )), ('rows split', (3, 1 << 20), (1, )), ('rows tiny', (130000, 33), (1, )),
    ('rows tiny 100', (45000, 100), (1, )), ('rows 64', (70000, 64), (1, )), ('cols', (65536, 256), (0, )),
    ('cols wide', (4096, 8192), (0, )), ('cols few rows', (7, 600004), (0, )), ('cols few rows ragged', (2, 5, 400001), (1, )),
    ('cols narrow', ((1 << 21) + 3, 3), (0, )), ('cols mid', (64, 512, 256), (1, )),
    ('cols small + inner first', (16, 16, 64, 64, 4), (1, 3)), ('inner first', (32, 64, 32, 128), (0, 2)),
    ('outer first', (4096, 64, 8), (0, 2)), ('two pass', (6, 50, 4096), (0, 2)),
]


@pytest.mark.parametrize('name,shape,axes', REDUCE_CASES, ids=[c[0] for c in REDUCE_CASES])
@pytest.mark.parametrize('dt', [torch.int32, torch.int64, torch.int16], ids=['i32', 'i64', 'i16'])
def test_reduce_writes_only_its_output(name, shape, axes, dt):
    x = _ints(shape, dt, -20, 20, seed=len(shape))
    ref = x.sum(axes, dtype=dt) if axes is not None else x.sum(dtype=dt).reshape(1)
    buf, out = _guarded(ref.numel(), dt, 12345)
    db.reduce(ADD, x, out.view(ref.shape) if axes is not None else out.view(()), axis=axes, identity=0)
    torch.cuda.synchronize()
    assert torch.equal(out.view(ref.shape), ref)
    assert _guards_intact(buf, ref.numel(), 12345)


def test_generic_strided_view_and_fill():
    x = _ints((300, 70, 50))[::2, :, ::3]
    ref = x.sum((0, 2), dtype=torch.int32)
    buf, out = _guarded(ref.numel(), torch.int32, 777)
    db.reduce(ADD, x, out, axis=(0, 2), identity=0)
    assert torch.equal(out, ref) and _guards_intact(buf, ref.numel(), 777)
    e = torch.empty((0, 9), dtype=torch.int32, device='cuda')            # empty reduced range: identity fill
    buf, out = _guarded(9, torch.int32, 777)
    db.reduce(ADD, e, out, axis=(0, ), identity=5)
    assert bool((out == 5).all()) and _guards_intact(buf, 9, 777)


@pytest.mark.parametrize('shape,axis', [((5000, 256), 1), ((3, 1 << 20), 1), ((5000, 256), 0), ((8, 200000), 0), ((5000, 9000), 1)])
def test_arg_reduce_writes_only_its_outputs(shape, axis):
    g = torch.Generator(device='cuda').manual_seed(3)
    x = torch.rand(shape, device='cuda', generator=g)
    n = shape[1 - axis]
    ibuf, idx = _guarded(n, torch.int32, -99)
    vbuf, val = _guarded(n, torch.float32, -7.0)
    db.argmax(x, axis=axis, out_idx=idx, out_val=val)
    torch.cuda.synchronize()
    assert torch.equal(idx.long(), x.argmax(dim=axis)) and torch.equal(val, x.amax(dim=axis))
    assert _guards_intact(ibuf, n, -99) and _guards_intact(vbuf, n, -7.0)


@pytest.mark.parametrize('shape', [(512, 8192), (140000, 33), (70001, 64), (8, 1 << 19), (700, 33000)])
def test_rows_max_sum_writes_only_its_outputs(shape):
    x = _ints(shape).float()
    mbuf, mx = _guarded(shape[0], torch.float32, -1e30)
    sbuf, sm = _guarded(shape[0], torch.float32, -1e30)
    db.rows_max_sum(x, mx, sm)
    torch.cuda.synchronize()
    assert torch.equal(mx, x.amax(1)) and torch.equal(sm, x.sum(1))
    assert _guards_intact(mbuf, shape[0], -1e30) and _guards_intact(sbuf, shape[0], -1e30)


@pytest.mark.parametrize('nbins,weighted,dist', [(256, False, 'uniform'), (5000, True, 'uniform'), (1 << 16, True, 'uniform'),
                                                 (1 << 16, False, 'uniform'), (1 << 20, False, 'uniform'),
                                                 ((1 << 20) - 3, False, 'uniform'), (1 << 20, True, 'uniform'),
                                                 (1 << 20, False, 'equal'), (1 << 20, False, 'sorted')])
def test_scatter_writes_only_its_bins(nbins, weighted, dist):
    n = (1 << 22) + 77
    if dist == 'equal':
        keys = torch.full((n, ), nbins - 1, dtype=torch.int32, device='cuda')        # the LAST bin, next to the guard
    else:
        keys = _ints((n, ), torch.int32, -2, nbins + 2, seed=nbins % 97)              # a few keys out of range
        if dist == 'sorted':
            keys = torch.sort(keys).values
    w = _ints((n, ), torch.int32, 0, 5, seed=1) if weighted else None
    buf, bins = _guarded(nbins, torch.int32, 424242)
    bins.zero_()
    db.wcr_scatter(ADD, bins, keys, w)
    torch.cuda.synchronize()
    ok = (keys >= 0) & (keys < nbins)
    ref = torch.zeros(nbins, dtype=torch.int32, device='cuda')
    ref.index_add_(0, keys[ok].long(), w[ok] if weighted else torch.ones(int(ok.sum()), dtype=torch.int32, device='cuda'))
    assert torch.equal(bins, ref)
    assert _guards_intact(buf, nbins, 424242)
