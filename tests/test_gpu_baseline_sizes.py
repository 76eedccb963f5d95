"""Parity at the sizes BASELINE.json quotes: the kernels behind the bench lines,
through the C ABI, against the CPU oracle on the SAME values.

* C2  full sum of 2^30 float32 (uniform [0,1) and the mixed-sign variant): the >= 256 MiB
      ``split_reduce_kernel<.., 32, 8, 1024>`` variant; <= 1e-5 relative to the RESULT for the
      all-positive data, <= 1e-5 of sum |x| for the mixed-sign data, against the oracle's fp64 fold.
* >= 256 MiB inputs of every dtype family reach the same 1024-thread variant: integers bit-exact.
* C3  [65536, 8192] bf16 -> fp32: fused max+sum and the single max; max bit-exact.
* C4  2^28 int32 keys -> 2^20 int32 bins: uniform, Zipf(1.1) (clipped and wrapped), all keys equal;
      bit-exact against the oracle's WCR loop and np.bincount.
* C5  one [2^19, 4096] float32 row shard, column sum (the wide-row column kernel).

Host copies of the inputs (4-8 GiB) are made once per test; the oracle's single-threaded C loops take
1-4 s at these sizes.
"""
import numpy as np
import pytest
import torch

import dace_b200 as db
from oracle import oracle as O

pytestmark = pytest.mark.gpu

ADD = 'lambda a, b: a + b'
MAX = 'lambda a, b: max(a, b)'
MIN = 'lambda a, b: min(a, b)'


def _free():
    torch.cuda.empty_cache()


@pytest.mark.parametrize('variant', ['uniform01', 'mixed_sign'])
def test_c2_full_sum_2p30_f32(variant, cuda):
    g = torch.Generator(device=cuda)
    g.manual_seed(1234)
    x = torch.rand(1 << 30, dtype=torch.float32, device=cuda, generator=g)
    if variant == 'mixed_sign':
        x.mul_(2).sub_(1)
    out = torch.full((), 123.0, dtype=torch.float32, device=cuda)
    db.reduce(ADD, x, out, axis=None, identity=0)
    got = float(out.item())
    again = torch.zeros((), dtype=torch.float32, device=cuda)
    db.reduce(ADD, x, again, axis=None, identity=0)
    assert float(again.item()) == got, 'the fixed-order combine must be run-to-run deterministic'
    host = x.cpu().numpy()
    del x
    _free()
    ref = float(O.reduce(O.SUM, host, None, identity=0, wide=True))   # fp64 fold, stored as fp32
    if variant == 'uniform01':
        assert abs(got - ref) <= 1e-5 * abs(ref), (got, ref)
    else:
        scale = float(np.abs(host, dtype=np.float32).sum(dtype=np.float64))
        assert abs(got - ref) <= 1e-5 * scale, (got, ref, scale)
    # identity=None folds into the value already there
    x2 = torch.from_numpy(host[:1 << 28]).to(cuda)
    acc = torch.full((), 1000.0, dtype=torch.float32, device=cuda)
    db.reduce(ADD, x2, acc, axis=None, identity=None)
    ref2 = 1000.0 + float(O.reduce(O.SUM, host[:1 << 28], None, identity=0, wide=True))
    assert abs(float(acc.item()) - ref2) <= 1e-5 * max(abs(ref2), float(np.abs(host[:1 << 28]).sum(dtype=np.float64)))


BIG = [  # (numpy dtype, elements) - every input is >= 256 MiB + a ragged tail, so the 1024-thread variant runs
    (np.int32, (1 << 26) + 12345), (np.uint32, (1 << 26) + 7), (np.int64, (1 << 25) + 333), (np.uint64, (1 << 25) + 1),
    (np.int16, (1 << 27) + 99), (np.uint8, (1 << 28) + 5), (np.int8, (1 << 28) + 3), (np.float64, (1 << 25) + 17),
]


@pytest.mark.parametrize('dtype,n', BIG)
def test_big_split_variant_every_dtype_family(dtype, n, cuda):
    rng = np.random.default_rng(n % 1000)
    if np.issubdtype(dtype, np.floating):
        a = rng.random(n).astype(dtype)
    else:
        info = np.iinfo(dtype)
        a = rng.integers(max(info.min, -1000), min(info.max, 1000), n).astype(dtype)
    ta = torch.from_numpy(a.view(np.int32 if dtype == np.uint32 else np.int64 if dtype == np.uint64 else dtype)).to(cuda)
    if dtype == np.uint32:
        ta = ta.view(torch.uint32)
    elif dtype == np.uint64:
        ta = ta.view(torch.uint64)
    for lam, red in ((ADD, O.SUM), (MAX, O.MAX), (MIN, O.MIN)):
        if np.issubdtype(dtype, np.floating):
            ident = 0 if red == O.SUM else (np.finfo(dtype).min if red == O.MAX else np.finfo(dtype).max)
        else:
            ident = 0 if red == O.SUM else (np.iinfo(dtype).min if red == O.MAX else np.iinfo(dtype).max)
        got = db.reduce(lam, ta, axis=None, identity=ident)
        got = got.cpu().view(ta.dtype).numpy().view(dtype) if dtype in (np.uint32, np.uint64) else got.cpu().numpy()
        ref = O.reduce(red, a, None, identity=ident, wide=True)
        if np.issubdtype(dtype, np.floating) and red == O.SUM:
            np.testing.assert_allclose(got, ref, rtol=1e-12)
        else:
            np.testing.assert_array_equal(got, ref)   # integer sums wrap like the reference's (mod 2^k)


def test_big_split_variant_several_long_rows_i32(cuda):
    """gridDim.y = rows with >= 2 MiB rows and >= 256 MiB in total: the 1024-thread variant over rows."""
    rng = np.random.default_rng(5)
    a = rng.integers(-100, 100, (3, (1 << 25) + 11)).astype(np.int32)
    got = db.reduce(ADD, torch.from_numpy(a).to(cuda), axis=1, identity=0).cpu().numpy()
    np.testing.assert_array_equal(got, O.reduce(O.SUM, a, (1, ), identity=0))


@pytest.mark.parametrize('positive', [False, True])
def test_c3_rows_max_sum_65536x8192_bf16(positive, cuda):
    g = torch.Generator(device=cuda)
    g.manual_seed(0)
    x = torch.randn(65536, 8192, dtype=torch.float32, device=cuda, generator=g)
    if positive:
        x.abs_()
    x = x.to(torch.bfloat16)
    mx, sm = db.rows_max_sum(x)
    mx1 = torch.empty(65536, dtype=torch.float32, device=cuda)
    db.reduce(MAX, x, mx1, axis=1, identity=-3.0e38)
    sm1 = torch.empty(65536, dtype=torch.float32, device=cuda)
    db.reduce(ADD, x, sm1, axis=1, identity=0)
    bits = x.view(torch.int16).cpu().numpy().view(np.uint16)
    del x
    _free()
    ref_max = O.reduce(O.MAX, bits, (1, ), identity=-3.0e38, in_bf16=True, out_dtype=np.float32)
    np.testing.assert_array_equal(mx.cpu().numpy(), ref_max)
    np.testing.assert_array_equal(mx1.cpu().numpy(), ref_max)
    ref_sum = O.reduce(O.SUM, bits, (1, ), identity=0, in_bf16=True, out_dtype=np.float32, wide=True).astype(np.float64)
    vals = O.bf16_bits_to_f32(bits)
    scale = np.abs(vals).sum(axis=1, dtype=np.float64)
    for got in (sm.cpu().numpy(), sm1.cpu().numpy()):
        assert np.all(np.abs(got.astype(np.float64) - ref_sum) <= 1e-5 * scale)
        if positive:
            np.testing.assert_allclose(got, ref_sum, rtol=1e-5)


def _c4_keys(dist, n, nbins):
    rng = np.random.default_rng(0)
    if dist == 'uniform':
        return rng.integers(0, nbins, n, dtype=np.int32)       # default_rng(0).integers(0, 2^20, 2^28, int32)
    if dist == 'all_equal':
        return np.full(n, 12345, np.int32)
    z = rng.zipf(1.1, n // 16)
    if dist == 'zipf_clipped':
        k = np.minimum(z - 1, nbins - 1).astype(np.int32)
    else:
        k = (z % nbins).astype(np.int32)
    return np.tile(k, 16)


@pytest.mark.parametrize('dist', ['uniform', 'zipf_wrapped', 'zipf_clipped', 'all_equal'])
def test_c4_histogram_2p28_keys_2p20_bins(dist, cuda):
    n, nbins = 1 << 28, 1 << 20
    keys = _c4_keys(dist, n, nbins)
    tk = torch.from_numpy(keys).to(cuda)
    bins = torch.zeros(nbins, dtype=torch.int32, device=cuda)
    db.wcr_scatter(ADD, bins, tk)
    got = bins.cpu().numpy()
    assert int(got.astype(np.int64).sum()) == n          # checksum of counts
    ref = O.wcr_scatter(O.SUM, np.zeros(nbins, np.int32), keys)
    np.testing.assert_array_equal(got, ref)
    np.testing.assert_array_equal(got, np.bincount(keys, minlength=nbins).astype(np.int32))
    # WCR accumulates into existing contents: a second pass doubles every count
    db.wcr_scatter(ADD, bins, tk)
    np.testing.assert_array_equal(bins.cpu().numpy(), 2 * ref)


def test_c5_shard_column_sum_2p19x4096_f32(cuda):
    g = torch.Generator(device=cuda)
    g.manual_seed(3)
    x = torch.rand(1 << 19, 4096, dtype=torch.float32, device=cuda, generator=g)
    out = torch.empty(4096, dtype=torch.float32, device=cuda)
    db.reduce(ADD, x, out, axis=0, identity=0)
    out2 = torch.empty(4096, dtype=torch.float32, device=cuda)
    db.reduce(ADD, x, out2, axis=0, identity=0)
    assert torch.equal(out, out2), 'fixed-order two-level combine must be deterministic'
    host = x.cpu().numpy()
    del x
    _free()
    got = out.cpu().numpy()
    # the oracle walks one output at a time (16 KiB strides over 8 GiB): run it on 64 columns, NumPy's fp64
    # column sums cover all 4096
    sl = np.ascontiguousarray(host[:, 1000:1064])
    ref = O.reduce(O.SUM, sl, (0, ), identity=0, wide=True)
    np.testing.assert_allclose(got[1000:1064], ref, rtol=1e-5)        # all-positive: relative to the result
    np.testing.assert_allclose(got, host.sum(axis=0, dtype=np.float64), rtol=1e-5)
    # max over the same shard: exact
    tx = torch.from_numpy(host).to(cuda)
    mx = db.reduce(MAX, tx, axis=0, identity=-1.0)
    np.testing.assert_array_equal(mx.cpu().numpy(), host.max(axis=0))
