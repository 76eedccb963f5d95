"""GPU: include/dace_b200_shared_accum.cuh (shared tile -> global accumulate / reset) through the native
harness tests/native/shared_accum_harness.cu: every operator / element type of the header, contiguous and strided
destinations (the contiguous fp32 sum takes the red.global.add.v4.f32 path), mostly-neutral tiles (skipped
entries).  Integers bit-exact against NumPy; float sums of small integers are exact too."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

HARNESS = os.path.join(os.path.dirname(__file__), 'native', '_built', 'libshared_accum_harness.so')

# case id -> (numpy dtype, fold over axis 0, neutral, value generator)
_I32, _I64 = np.iinfo(np.int32), np.iinfo(np.int64)
CASES = {
    0: (np.int32, np.add.reduce, 0, lambda r, s: r.integers(-50, 50, s)),
    1: (np.float32, np.add.reduce, 0, lambda r, s: r.integers(-8, 9, s)),
    2: (np.int32, np.maximum.reduce, _I32.min, lambda r, s: r.integers(-10**6, 10**6, s)),
    3: (np.uint64, np.add.reduce, 0, lambda r, s: r.integers(0, 2**40, s)),
    4: (np.float32, np.maximum.reduce, -np.inf, lambda r, s: r.uniform(-100, 100, s)),
    5: (np.float64, np.minimum.reduce, np.inf, lambda r, s: r.uniform(-100, 100, s)),
    6: (np.int32, np.bitwise_xor.reduce, 0, lambda r, s: r.integers(0, 2**31 - 1, s)),
    7: (np.float32, np.multiply.reduce, 1, lambda r, s: r.choice([0.5, 1.0, 2.0, -1.0], s)),
    8: (np.int64, np.minimum.reduce, _I64.max, lambda r, s: r.integers(-10**12, 10**12, s)),
    9: (np.float64, np.add.reduce, 0, lambda r, s: r.integers(-1000, 1000, s)),
}


@pytest.fixture(scope='module')
def lib():
    if not os.path.exists(HARNESS):
        pytest.fail('tests/native/_built/libshared_accum_harness.so missing: run __graft_entry__.build()')
    h = ctypes.CDLL(HARNESS)
    h.shared_accum_harness_len.restype = ctypes.c_int
    h.shared_accum_harness_run.restype = ctypes.c_int
    h.shared_accum_harness_run.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_int, ctypes.c_void_p]
    return h


@pytest.mark.parametrize('case', sorted(CASES))
@pytest.mark.parametrize('gstride', [1, 3])
@pytest.mark.parametrize('sparse', [False, True])
def test_shared_tile_flush(cuda, lib, case, gstride, sparse):
    dt, fold, neutral, gen = CASES[case]
    n = lib.shared_accum_harness_len(case)
    nblocks, items = 37, 5
    rng = np.random.default_rng(case * 10 + gstride)
    src = gen(rng, (nblocks, items, n)).astype(dt)
    if sparse:      # most entries neutral: those tile slots must be skipped, and the result must not change
        mask = rng.random((nblocks, items, n)) < 0.9
        src[mask] = neutral
    out0 = gen(rng, (n * gstride, )).astype(dt)
    t_src = torch.from_numpy(src.view(np.int64) if dt == np.uint64 else src).to(cuda)
    t_out = torch.from_numpy((out0.view(np.int64) if dt == np.uint64 else out0).copy()).to(cuda)
    rc = lib.shared_accum_harness_run(case, t_src.data_ptr(), t_out.data_ptr(), nblocks, items, gstride,
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    got = t_out.cpu().numpy()
    got = got.view(np.uint64) if dt == np.uint64 else got
    want = out0.copy()
    with np.errstate(over='ignore'):
        folded = fold(src.reshape(nblocks * items, n), axis=0)
        want[::gstride][:n] = fold(np.stack([out0[::gstride][:n], folded.astype(dt)]), axis=0)
    np.testing.assert_array_equal(got, want)     # untouched slots of a strided destination included
