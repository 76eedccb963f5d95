"""GPU: include/dace_b200_block_reduce.cuh (in-kernel block / warp reductions, the replacement of
cub::BlockReduce in the reference's `CUDA (block)` expansion, reference
dace/libraries/standard/nodes/reduce.py:698-712) against NumPy / the oracle, through the native harness
tests/native/block_reduce_harness.cu.  Integers bit-exact; float sums within 1e-5 * sum|x| of fp64."""
import ctypes
import os

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

HARNESS = os.path.join(os.path.dirname(__file__), 'native', '_built', 'libblock_reduce_harness.so')

# case id -> (dtype, threads per block, items per thread, operator)   [same table as the harness]
CASES = {
    0: (np.float32, 32, 2, 'sum'), 1: (np.float32, 128, 1, 'sum'), 2: (np.int32, 96, 3, 'sum'),
    3: (np.int32, 50, 2, 'max'), 4: (np.float64, 1024, 1, 'min'), 5: (np.uint32, 32, 4, 'bxor'),
    6: (np.int64, 7, 2, 'sum'), 7: (np.int64, 1000, 1, 'max'), 8: (np.int32, 256, 1, 'min'),
    9: (np.uint32, 33, 1, 'band'), 10: (np.uint32, 480, 2, 'bor'), 11: (np.float64, 65, 3, 'prod'),
    12: (np.int32, 192, 1, 'land'), 13: (np.int32, 192, 1, 'lor'), 14: (np.int32, 100, 1, 'lxor'),
    15: (np.int64, 72, 2, 'affine'), 16: (np.float32, 1, 5, 'sum'), 17: (np.float32, 1024, 2, 'max'),
}
ORACLE = {'sum': O.SUM, 'max': O.MAX, 'min': O.MIN, 'bxor': O.BXOR, 'band': O.BAND, 'bor': O.BOR, 'prod': O.PRODUCT,
          'land': O.LAND, 'lor': O.LOR, 'lxor': O.LXOR}


@pytest.fixture(scope='module')
def lib():
    if not os.path.exists(HARNESS):
        pytest.fail('tests/native/_built/libblock_reduce_harness.so missing: run __graft_entry__.build()')
    h = ctypes.CDLL(HARNESS)
    h.block_reduce_harness_run.restype = ctypes.c_int
    h.block_reduce_harness_run.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_int, ctypes.c_void_p]
    h.block_reduce_harness_warp.restype = ctypes.c_int
    h.block_reduce_harness_warp.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int, ctypes.c_void_p]
    return h


def _data(rng, dt, n, opname):
    if opname in ('land', 'lor', 'lxor'):
        return (rng.random(n) < (0.98 if opname == 'land' else 0.02 if opname == 'lor' else 0.5)).astype(dt) * 3
    if opname == 'prod':
        return (1.0 + 0.01 * rng.standard_normal(n)).astype(dt)
    if opname == 'affine':
        return rng.integers(-3, 4, n).astype(dt)
    if opname == 'band':
        return (rng.integers(0, 2**32, n, dtype=np.uint64) | rng.integers(0, 2**32, n, dtype=np.uint64)).astype(dt)
    if np.issubdtype(dt, np.floating):
        return rng.standard_normal(n).astype(dt)
    info = np.iinfo(dt)
    lo, hi = (max(info.min, -2**40), min(info.max, 2**40)) if opname != 'sum' else (-1000, 1000)
    return rng.integers(lo, hi, n, dtype=np.int64 if info.min < 0 else np.uint64).astype(dt)


def _fold(opname, x, axis):
    if opname == 'affine':   # a + b + a*b == (1+a)(1+b) - 1, exact in wrap-around int64
        with np.errstate(over='ignore'):
            return np.prod(1 + x, axis=axis, dtype=np.int64) - 1
    ident = {'sum': 0, 'prod': 1, 'bxor': 0, 'bor': 0, 'lor': 0, 'lxor': 0, 'land': 1, 'band': ~np.uint32(0)}.get(opname)
    if ident is None:
        info = np.finfo(x.dtype) if np.issubdtype(x.dtype, np.floating) else np.iinfo(x.dtype)
        ident = info.max if opname == 'min' else info.min
    return O.reduce(ORACLE[opname], x, (axis, ), identity=ident, wide=opname in ('sum', 'prod'))


@pytest.mark.parametrize('case', sorted(CASES))
def test_block_reduce_cases(cuda, lib, case):
    dt, nt, items, opname = CASES[case]
    nblocks = 37
    rng = np.random.default_rng(case)
    x = _data(rng, dt, nblocks * nt * items, opname)
    tin = torch.from_numpy(x.view(np.uint8)).to(cuda)
    out = torch.zeros(nblocks * x.itemsize, dtype=torch.uint8, device=cuda)
    out_all = torch.zeros(nblocks * nt * x.itemsize, dtype=torch.uint8, device=cuda)
    rc = lib.block_reduce_harness_run(case, tin.data_ptr(), out.data_ptr(), out_all.data_ptr(), nblocks,
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(dt)
    got_all = out_all.cpu().numpy().view(dt).reshape(nblocks, nt)
    blocks = x.reshape(nblocks, nt * items)
    ref = _fold(opname, blocks, 1)
    first = x.reshape(nblocks, nt, items)[:, :, 0]
    ref_all = _fold(opname, first, 1)
    if np.issubdtype(dt, np.floating) and opname in ('sum', 'prod'):
        tol = 1e-5 if dt == np.float32 else 1e-12
        scale = np.abs(blocks.astype(np.float64)).sum(1) if opname == 'sum' else np.abs(ref.astype(np.float64))
        assert np.all(np.abs(got.astype(np.float64) - ref) <= tol * scale)
        scale_all = np.abs(first.astype(np.float64)).sum(1) if opname == 'sum' else np.abs(ref_all.astype(np.float64))
        assert np.all(np.abs(got_all.astype(np.float64) - np.asarray(ref_all)[:, None]) <= tol * scale_all[:, None])
    else:
        np.testing.assert_array_equal(got, np.asarray(ref).astype(dt))
        np.testing.assert_array_equal(got_all, np.broadcast_to(np.asarray(ref_all).astype(dt)[:, None], (nblocks, nt)))
    # every thread of a block holds the same all-reduce value, bit for bit
    assert np.all(got_all.view(np.uint8).reshape(nblocks, nt, -1) == got_all.view(np.uint8).reshape(nblocks, nt, -1)[:, :1])


def test_block_reduce_custom_struct_functor(cuda, lib):
    """16-byte value/index pairs with a custom functor (argmax, first index wins ties)."""
    nblocks, nt = 21, 200
    rng = np.random.default_rng(5)
    x = rng.integers(0, 50, nblocks * nt).astype(np.float64)   # many ties
    tin = torch.from_numpy(x).to(cuda)
    out = torch.zeros(nblocks * 2, dtype=torch.int64, device=cuda)
    out_all = torch.zeros(nblocks * nt * 2, dtype=torch.int64, device=cuda)
    assert lib.block_reduce_harness_run(100, tin.data_ptr(), out.data_ptr(), out_all.data_ptr(), nblocks,
                                        torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    o = out.cpu().numpy().reshape(nblocks, 2)
    oa = out_all.cpu().numpy().reshape(nblocks, nt, 2)
    xb = x.reshape(nblocks, nt)
    np.testing.assert_array_equal(o[:, 0].copy().view(np.float64), xb.max(1))
    np.testing.assert_array_equal(o[:, 1], xb.argmax(1))
    assert np.all(oa == o[:, None, :])


def test_warp_allreduce(cuda, lib):
    nwarps = 64
    rng = np.random.default_rng(6)
    xi = rng.integers(-10**6, 10**6, nwarps * 32).astype(np.int32)
    xf = rng.standard_normal(nwarps * 32).astype(np.float32)
    ti, tf = torch.from_numpy(xi).to(cuda), torch.from_numpy(xf).to(cuda)
    oi, of = torch.zeros_like(ti), torch.zeros_like(tf)
    assert lib.block_reduce_harness_warp(ti.data_ptr(), oi.data_ptr(), tf.data_ptr(), of.data_ptr(), nwarps,
                                         torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    np.testing.assert_array_equal(oi.cpu().numpy().reshape(nwarps, 32),
                                  np.broadcast_to(xi.reshape(nwarps, 32).sum(1, dtype=np.int32)[:, None], (nwarps, 32)))
    np.testing.assert_array_equal(of.cpu().numpy().reshape(nwarps, 32),
                                  np.broadcast_to(xf.reshape(nwarps, 32).max(1)[:, None], (nwarps, 32)))
