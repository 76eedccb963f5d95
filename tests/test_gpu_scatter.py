"""GPU parity for WCR map reductions (scatter-add / histogram) against the oracle.

Workloads follow reference samples/explicit/histogram.py:12-26 (-> uint32[256]),
reference tests/mapreduce_test.py:274-291 (uint8 image -> uint32[256] histogram),
reference tests/npbench/misc/azimint_hist_test.py:39-59 (hist[bin] += 1 / += w)
and BASELINE config C4 (int32 keys -> 2^20 int32 bins, bit-exact).
"""
import numpy as np
import pytest
import torch

import dace_b200 as db
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _hist(keys_np, nbins, cuda, bins_dtype=np.int32):
    tk = torch.from_numpy(keys_np).to(cuda)
    bins = torch.zeros(nbins, dtype=getattr(torch, np.dtype(bins_dtype).name), device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, tk)
    return bins.cpu().numpy()


@pytest.mark.parametrize('nbins', [1, 2, 7, 256, 1000, 4096, 20000, 24576, 24577, 100000, 1 << 20])
@pytest.mark.parametrize('n', [1, 1000, (1 << 20) + 13])
def test_histogram_int32_exact(nbins, n, cuda):
    rng = np.random.default_rng(nbins * 7 + n)
    keys = rng.integers(0, nbins, n).astype(np.int32)
    got = _hist(keys, nbins, cuda)
    ref = O.wcr_scatter(O.SUM, np.zeros(nbins, np.int32), keys)
    np.testing.assert_array_equal(got, ref)
    np.testing.assert_array_equal(got, np.bincount(keys, minlength=nbins).astype(np.int32))


@pytest.mark.parametrize('dist', ['uniform', 'zipf', 'equal', 'sorted'])
@pytest.mark.parametrize('nbins', [256, 1 << 20])
def test_histogram_distributions(dist, nbins, cuda):
    rng = np.random.default_rng(0)
    n = (1 << 22) + 5
    if dist == 'uniform':
        keys = rng.integers(0, nbins, n)
    elif dist == 'zipf':
        keys = np.minimum(rng.zipf(1.1, n) - 1, nbins - 1)
    elif dist == 'equal':
        keys = np.full(n, nbins // 3)
    else:
        keys = np.sort(rng.integers(0, nbins, n))
    keys = keys.astype(np.int32)
    got = _hist(keys, nbins, cuda)
    np.testing.assert_array_equal(got, np.bincount(keys, minlength=nbins).astype(np.int32))
    assert int(got.astype(np.int64).sum()) == n   # checksum of the counts


def test_uint8_image_histogram_uint32(cuda):
    """reference tests/mapreduce_test.py:274-291."""
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (480, 640)).astype(np.uint8)
    got = _hist(img.reshape(-1), 256, cuda, np.uint32)
    ref = O.wcr_scatter(O.SUM, np.zeros(256, np.uint32), img.reshape(-1))
    np.testing.assert_array_equal(got, ref)
    np.testing.assert_array_equal(got, np.histogram(img, bins=256, range=(0, 256))[0].astype(np.uint32))


def test_accumulates_into_existing_bins_and_skips_oob(cuda):
    rng = np.random.default_rng(2)
    keys = rng.integers(-5, 105, 100000).astype(np.int64)
    start = rng.integers(0, 9, 100).astype(np.int64)
    bins = torch.from_numpy(start.copy()).to(cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, torch.from_numpy(keys).to(cuda), 3)
    ref = O.wcr_scatter(O.SUM, start.copy(), keys, 3)
    np.testing.assert_array_equal(bins.cpu().numpy(), ref)
    ok = keys[(keys >= 0) & (keys < 100)]
    np.testing.assert_array_equal(ref, start + 3 * np.bincount(ok, minlength=100))


@pytest.mark.parametrize('nbins', [64, 5000, 300000])
def test_weighted_scatter_add(nbins, cuda):
    """hist[bin] += weights[i]: ints exact, floats to 1e-5 of the per-bin magnitude."""
    rng = np.random.default_rng(3)
    n = 500000
    keys = rng.integers(0, nbins, n).astype(np.int32)
    tk = torch.from_numpy(keys).to(cuda)
    wi = rng.integers(-100, 100, n).astype(np.int64)
    bins = torch.zeros(nbins, dtype=torch.int64, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, tk, torch.from_numpy(wi).to(cuda))
    np.testing.assert_array_equal(bins.cpu().numpy(), O.wcr_scatter(O.SUM, np.zeros(nbins, np.int64), keys, wi))
    for dt, tdt, tol in ((np.float32, torch.float32, 1e-5), (np.float64, torch.float64, 1e-12)):
        wf = rng.uniform(0, 1, n).astype(dt)
        bins = torch.zeros(nbins, dtype=tdt, device=cuda)
        db.wcr_scatter('lambda a, b: a + b', bins, tk, torch.from_numpy(wf).to(cuda))
        ref = np.bincount(keys, weights=wf.astype(np.float64), minlength=nbins)
        np.testing.assert_allclose(bins.cpu().numpy(), ref, rtol=tol * 10, atol=tol)


@pytest.mark.parametrize('op,red', [('lambda a, b: max(a, b)', O.MAX), ('lambda a, b: min(a, b)', O.MIN)])
@pytest.mark.parametrize('dt,tdt', [(np.int32, torch.int32), (np.int64, torch.int64), (np.float32, torch.float32),
                                    (np.float64, torch.float64), (np.uint32, torch.uint32)])
def test_scatter_min_max_exact(op, red, dt, tdt, cuda):
    rng = np.random.default_rng(4)
    n, nbins = 200000, 777
    keys = rng.integers(0, nbins, n).astype(np.int32)
    if np.issubdtype(dt, np.floating):
        vals = rng.uniform(-100, 100, n).astype(dt)
        start = np.zeros(nbins, dt)
    elif np.issubdtype(dt, np.unsignedinteger):
        vals = rng.integers(0, 1 << 31, n).astype(dt)
        start = np.full(nbins, 1 << 20, dt)
    else:
        vals = rng.integers(-1 << 30, 1 << 30, n).astype(dt)
        start = np.zeros(nbins, dt)
    bins = torch.from_numpy(start.copy()).to(cuda)
    db.wcr_scatter(op, bins, torch.from_numpy(keys).to(cuda), torch.from_numpy(vals).to(cuda))
    ref = O.wcr_scatter(red, start.copy(), keys, vals)
    np.testing.assert_array_equal(bins.cpu().numpy(), ref)


def test_scatter_bitwise_and_product(cuda):
    rng = np.random.default_rng(5)
    n, nbins = 100000, 100
    keys = rng.integers(0, nbins, n).astype(np.int32)
    tk = torch.from_numpy(keys).to(cuda)
    vals = rng.integers(0, 1 << 30, n).astype(np.int32)
    for op, red, init in (('lambda a, b: a | b', O.BOR, 0), ('lambda a, b: a ^ b', O.BXOR, 0),
                          ('lambda a, b: a & b', O.BAND, -1)):
        start = np.full(nbins, init, np.int32)
        bins = torch.from_numpy(start.copy()).to(cuda)
        db.wcr_scatter(op, bins, tk, torch.from_numpy(vals).to(cuda))
        np.testing.assert_array_equal(bins.cpu().numpy(), O.wcr_scatter(red, start.copy(), keys, vals))
    pv = rng.choice(np.array([1, 1, 1, -1, 2], np.int64), 3000)
    pk = rng.integers(0, nbins, 3000).astype(np.int32)
    bins = torch.ones(nbins, dtype=torch.int64, device=cuda)
    db.wcr_scatter('lambda a, b: a * b', bins, torch.from_numpy(pk).to(cuda), torch.from_numpy(pv).to(cuda))
    np.testing.assert_array_equal(bins.cpu().numpy(), O.wcr_scatter(O.PRODUCT, np.ones(nbins, np.int64), pk, pv))


def test_unaligned_index_pointer(cuda):
    rng = np.random.default_rng(6)
    keys = rng.integers(0, 1 << 17, 100003).astype(np.int32)
    tk = torch.from_numpy(keys).to(cuda)
    for off in (1, 2, 3):
        bins = torch.zeros(1 << 17, dtype=torch.int32, device=cuda)
        db.wcr_scatter('lambda a, b: a + b', bins, tk[off:])
        np.testing.assert_array_equal(bins.cpu().numpy(), np.bincount(keys[off:], minlength=1 << 17))


@pytest.mark.parametrize('nbins', [24577, 33000, 65536, 100000, 300001, 1 << 19, (1 << 19) + 1, 1 << 20, (1 << 21) - 5])   # 17 ... 64 buckets of 2^10 ... 2^15 bins
@pytest.mark.parametrize('dist', ['uniform', 'zipf', 'halfhot'])
def test_partitioned_histogram_large_bins(nbins, dist, cuda):
    """Radix-partition + shared-memory histogram path (wcr_partition.cu): several 2^23-key chunks,
    partial tiles, skewed keys overflowing the bucket regions, out-of-range keys, existing contents."""
    rng = np.random.default_rng(nbins % 1000)
    n = (1 << 24) + (1 << 22) + 4321
    if dist == 'uniform':
        keys = rng.integers(-3, nbins + 3, n)           # a few keys out of range on both sides
    elif dist == 'zipf':
        keys = np.minimum(rng.zipf(1.2, n) - 1, nbins - 1)
    else:
        keys = rng.integers(0, nbins, n)
        keys[::2] = nbins // 2                          # half of all keys hit one bin
    keys = keys.astype(np.int32)
    start = rng.integers(0, 5, nbins).astype(np.int32)
    bins = torch.from_numpy(start.copy()).to(cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, torch.from_numpy(keys).to(cuda))
    ok = keys[(keys >= 0) & (keys < nbins)]
    np.testing.assert_array_equal(bins.cpu().numpy(), start + np.bincount(ok, minlength=nbins).astype(np.int32))


def test_partitioned_histogram_variants(cuda):
    rng = np.random.default_rng(9)
    n, nbins = (1 << 23) + 999, 300000
    keys = rng.integers(0, nbins, n).astype(np.int32)
    tk = torch.from_numpy(keys).to(cuda)
    ref = np.bincount(keys, minlength=nbins)
    # unaligned key pointer
    for off in (1, 2, 3):
        bins = torch.zeros(nbins, dtype=torch.int32, device=cuda)
        db.wcr_scatter('lambda a, b: a + b', bins, tk[off:])
        np.testing.assert_array_equal(bins.cpu().numpy(), np.bincount(keys[off:], minlength=nbins))
    # constant increment other than 1, uint32 and float32 bins, uint32 keys
    bins = torch.zeros(nbins, dtype=torch.int32, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, tk, 3)
    np.testing.assert_array_equal(bins.cpu().numpy(), 3 * ref)
    bu = torch.zeros(nbins, dtype=torch.uint32, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bu, torch.from_numpy(keys.astype(np.uint32)).to(cuda))
    np.testing.assert_array_equal(bu.cpu().numpy(), ref.astype(np.uint32))
    bf = torch.zeros(nbins, dtype=torch.float32, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bf, tk, 0.5)
    np.testing.assert_array_equal(bf.cpu().numpy(), (0.5 * ref).astype(np.float32))   # small counts: exact in fp32
    # repeated calls reuse the per-device scratch; two streams are ordered through it
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    b1 = torch.zeros(nbins, dtype=torch.int32, device=cuda)
    b2 = torch.zeros(nbins, dtype=torch.int32, device=cuda)
    torch.cuda.synchronize()
    for _ in range(3):
        db.wcr_scatter('lambda a, b: a + b', b1, tk, stream=s1)
        db.wcr_scatter('lambda a, b: a + b', b2, tk, stream=s2)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(b1.cpu().numpy(), 3 * ref)
    np.testing.assert_array_equal(b2.cpu().numpy(), 3 * ref)


@pytest.mark.parametrize('dist', ['all_equal', 'sorted', 'zipf', 'uniform'])
def test_global_path_merges_equal_keys_in_a_warp(dist, cuda):
    """Large int64 / min / max bin arrays take the global-atomic kernel, which switches to
    warp-merged atomics when a warp samples duplicate keys (skewed or sorted inputs)."""
    rng = np.random.default_rng(11)
    n, nbins = (1 << 22) + 77, 1 << 18
    if dist == 'all_equal':
        keys = np.full(n, 4242)
    elif dist == 'sorted':
        keys = np.sort(rng.integers(0, nbins, n))
    elif dist == 'zipf':
        keys = np.minimum(rng.zipf(1.2, n) - 1, nbins - 1)
    else:
        keys = rng.integers(-2, nbins + 2, n)
    keys = keys.astype(np.int32)
    tk = torch.from_numpy(keys).to(cuda)
    ok = keys[(keys >= 0) & (keys < nbins)]
    b64 = torch.zeros(nbins, dtype=torch.int64, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', b64, tk, 5)
    np.testing.assert_array_equal(b64.cpu().numpy(), 5 * np.bincount(ok, minlength=nbins))
    bmax = torch.full((nbins, ), -7, dtype=torch.int32, device=cuda)
    db.wcr_scatter('lambda a, b: max(a, b)', bmax, tk, 9)
    exp = np.full(nbins, -7, np.int32)
    exp[np.unique(ok)] = 9
    np.testing.assert_array_equal(bmax.cpu().numpy(), exp)
    bf = torch.zeros(nbins, dtype=torch.float64, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bf, tk, 0.25)
    np.testing.assert_array_equal(bf.cpu().numpy(), 0.25 * np.bincount(ok, minlength=nbins))


@pytest.mark.parametrize('dist', ['sorted', 'runs', 'all_equal', 'uniform_then_sorted'])
def test_device_side_path_selection(dist, cuda):
    """int32 Sum into a large bin array: a device-side key sample routes sorted / long-run inputs to the
    warp-merged global-atomic kernel and everything else to the partitioned kernel (exactly one of
    the two runs).  Whatever is chosen, counts are bit-exact and back-to-back calls on two streams with
    different decisions do not disturb each other."""
    rng = np.random.default_rng(23)
    n, nbins = (1 << 23) + 1234, 1 << 20
    uniform = rng.integers(0, nbins, n).astype(np.int32)
    if dist == 'sorted':
        keys = np.sort(uniform)
    elif dist == 'runs':
        keys = np.repeat(rng.integers(0, nbins, n // 64 + 1), 64)[:n].astype(np.int32)
    elif dist == 'all_equal':
        keys = np.full(n, 77777, np.int32)
    else:
        keys = uniform
    start = rng.integers(0, 3, nbins).astype(np.int32)
    bins = torch.from_numpy(start.copy()).to(cuda)
    tk = torch.from_numpy(keys).to(cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, tk)
    np.testing.assert_array_equal(bins.cpu().numpy(), start + np.bincount(keys, minlength=nbins).astype(np.int32))
    if dist == 'uniform_then_sorted':
        ts = torch.from_numpy(np.sort(uniform)).to(cuda)
        b1 = torch.zeros(nbins, dtype=torch.int32, device=cuda)
        b2 = torch.zeros(nbins, dtype=torch.int32, device=cuda)
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        torch.cuda.synchronize()
        for _ in range(3):
            with torch.cuda.stream(s1):
                db.wcr_scatter('lambda a, b: a + b', b1, tk)
            with torch.cuda.stream(s2):
                db.wcr_scatter('lambda a, b: a + b', b2, ts)
        torch.cuda.synchronize()
        ref = 3 * np.bincount(uniform, minlength=nbins).astype(np.int32)
        np.testing.assert_array_equal(b1.cpu().numpy(), ref)
        np.testing.assert_array_equal(b2.cpu().numpy(), ref)


@pytest.mark.parametrize('dist', ['half_hot', 'zipf', 'uniform', 'all_equal'])
def test_weighted_scatter_large_bins_skew_goes_through_the_hot_key_cache(dist, cuda):
    """Per-element values into a large bin array: no partition, but the key sampler still routes skewed keys to the
    shared-memory hot-key cache (values path) instead of letting hot bins serialise in L2.  Integer values exact;
    float values are small integers, so their sums are exact too whatever the order."""
    rng = np.random.default_rng(31)
    n, nbins = (1 << 22) + 1234, 1 << 20
    if dist == 'half_hot':
        keys = rng.integers(0, nbins, n)
        keys[rng.random(n) < 0.5] = 31337
    elif dist == 'zipf':
        keys = rng.zipf(1.2, n) % nbins
    elif dist == 'all_equal':
        keys = np.full(n, 99)
    else:
        keys = rng.integers(-2, nbins + 2, n)          # a few out of range
    keys = keys.astype(np.int32)
    ok = (keys >= 0) & (keys < nbins)
    tk = torch.from_numpy(keys).to(cuda)
    for dt, tdt in ((np.int32, torch.int32), (np.float32, torch.float32), (np.int64, torch.int64), (np.float64, torch.float64)):
        w = rng.integers(0, 4, n).astype(dt)
        start = rng.integers(0, 3, nbins).astype(dt)
        bins = torch.from_numpy(start.copy()).to(cuda)
        db.wcr_scatter('lambda a, b: a + b', bins, tk, torch.from_numpy(w).to(cuda))
        want = start + np.bincount(keys[ok], weights=w[ok].astype(np.float64), minlength=nbins).astype(dt)
        np.testing.assert_array_equal(bins.cpu().numpy(), want)
    lib = __import__('dace_b200')._lib.load()
    assert lib.dace_b200_wcr_scatter_workspace_bytes(3, 2, 0, 1, n, nbins) == 256      # values: the sampler's flag only
    assert lib.dace_b200_wcr_scatter_workspace_bytes(3, 2, 0, 1, 1000, nbins) == 0


@pytest.mark.parametrize('nbins', [7, 256, 5000, 24576])
@pytest.mark.parametrize('op,npop', [('lambda a, b: max(a, b)', np.maximum), ('lambda a, b: min(a, b)', np.minimum)])
def test_privatised_min_max_small_bins(nbins, op, npop, cuda):
    """Min / Max into bins that fit shared memory: per-warp private copies seeded with the neutral element, flushed
    with one atomic per touched bin - exact for every type, negative floats included, existing contents kept,
    constant value and per-element values, skewed keys (the warp merge)."""
    rng = np.random.default_rng(nbins)
    n = (1 << 20) + 77
    keys = rng.integers(-2, nbins + 2, n).astype(np.int32)
    keys[: n // 3] = nbins // 2                                     # a hot bin
    ok = (keys >= 0) & (keys < nbins)
    tk = torch.from_numpy(keys).to(cuda)
    for dt in (np.int32, np.int64, np.float32, np.float64, np.uint32):
        if np.issubdtype(dt, np.floating):
            vals = rng.uniform(-100, 100, n).astype(dt)
            start = rng.uniform(-50, 50, nbins).astype(dt)
        elif dt == np.uint32:
            vals = rng.integers(0, 1 << 31, n).astype(dt)
            start = rng.integers(1 << 29, 1 << 30, nbins).astype(dt)
        else:
            vals = rng.integers(-1 << 30, 1 << 30, n).astype(dt)
            start = rng.integers(-1 << 28, 1 << 28, nbins).astype(dt)
        bins = torch.from_numpy(start.copy()).to(cuda)
        db.wcr_scatter(op, bins, tk, torch.from_numpy(vals).to(cuda))
        want = start.copy()
        npop.at(want, keys[ok], vals[ok])
        np.testing.assert_array_equal(bins.cpu().numpy(), want)
        # constant value: every touched bin meets the constant once
        bins = torch.from_numpy(start.copy()).to(cuda)
        const = vals[:1].copy()
        db.wcr_scatter(op, bins, tk, const[0].item())
        want = start.copy()
        touched = np.zeros(nbins, bool)
        touched[keys[ok]] = True
        want[touched] = npop(want[touched], const[0])
        np.testing.assert_array_equal(bins.cpu().numpy(), want)


@pytest.mark.parametrize('nbins', [65536, 300001, 1 << 20])
@pytest.mark.parametrize('dist', ['uniform', 'zipf', 'equal', 'sorted'])
def test_partitioned_histogram_int64_keys(nbins, dist, cuda):
    """int64 keys (DaCe's default index type) through the radix partition: a key is in range only if its high word
    is zero - negative keys and keys >= 2^32 whose LOW word happens to be a valid bin are dropped."""
    rng = np.random.default_rng(nbins % 89)
    n = (1 << 23) + 4321
    if dist == 'uniform':
        keys = rng.integers(-3, nbins + 3, n)
        keys[::1001] += 1 << 32                          # low word in range, high word not
        keys[5::2003] = -keys[5::2003] - 1
    elif dist == 'zipf':
        keys = np.minimum(rng.zipf(1.2, n) - 1, nbins - 1)
    elif dist == 'equal':
        keys = np.full(n, nbins - 1)
    else:
        keys = np.sort(rng.integers(0, nbins, n))
    keys = keys.astype(np.int64)
    ok = keys[(keys >= 0) & (keys < nbins)]
    for dt, tdt in ((np.int32, torch.int32), (np.int64, torch.int64)):
        start = rng.integers(0, 5, nbins).astype(dt)
        bins = torch.from_numpy(start.copy()).to(cuda)
        db.wcr_scatter('lambda a, b: a + b', bins, torch.from_numpy(keys).to(cuda))
        np.testing.assert_array_equal(bins.cpu().numpy(), start + np.bincount(ok, minlength=nbins).astype(dt))
    # a key pointer 8 bytes off a 16-byte boundary: the scalar loads
    tk = torch.from_numpy(np.concatenate([keys[:1], keys])).to(cuda)[1:]
    bins = torch.zeros(nbins, dtype=torch.int32, device=cuda)
    db.wcr_scatter('lambda a, b: a + b', bins, tk)
    np.testing.assert_array_equal(bins.cpu().numpy(), np.bincount(ok, minlength=nbins).astype(np.int32))
