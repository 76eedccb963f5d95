"""Allocation-free entry point ``dace_b200_reduce_rows_noalloc`` (scratch in a file-scope
``__device__`` array) against the regular ``dace_b200_reduce_rows`` on the same device buffers.

Collected last on purpose: a failure here must not hide other tests.
Both entry points run the same kernel with the same split geometry, so integer results must be
bit-identical and float results agree within the suite's fp32 tolerance (1e-5 of sum |x|); the
test calls the new entry point twice in a row because the static scratch keeps the previous
call's contents.
"""
import ctypes

import numpy as np
import pytest
import torch

from dace_b200 import _lib
from dace_b200.api import Workspace

pytestmark = pytest.mark.gpu

SUM, MAX = 3, 2
F32, I32 = 0, 2


def _call_noalloc(lib, red, dt, x, out, rows, n, stride, identity_bytes):
    ident = ctypes.create_string_buffer(identity_bytes, 8)
    st = lib.dace_b200_reduce_rows_noalloc(red, dt, dt, x.data_ptr(), out.data_ptr(), rows, n, stride, ident,
                                           torch.cuda.current_stream().cuda_stream)
    assert st == 0, lib.dace_b200_status_string(st)


_WS = []


def _call_regular(lib, red, dt, x, out, rows, n, stride, identity_bytes):
    if not _WS:
        _WS.append(Workspace(x.device))     # the regular entry point works in a caller-owned workspace
    ident = ctypes.create_string_buffer(identity_bytes, 8)
    st = lib.dace_b200_reduce_rows(red, dt, dt, x.data_ptr(), out.data_ptr(), rows, n, stride, ident, _WS[0].ptr,
                                   _WS[0].nbytes, torch.cuda.current_stream().cuda_stream)
    assert st == 0, lib.dace_b200_status_string(st)


@pytest.mark.parametrize('rows,n', [(1, (1 << 22) + 77), (3, 1 << 20), (7, 300001), (256, 5000), (2, 100)])
def test_noalloc_matches_regular_entry_point_twice(rows, n, cuda):
    lib = _lib.load()
    g = torch.Generator(device=cuda)
    g.manual_seed(7)
    stride = n + 5          # segments need not be packed
    xf = torch.rand(rows * stride, device=cuda, generator=g, dtype=torch.float32) * 2 - 1
    xi = torch.randint(-1000, 1000, (rows * stride, ), device=cuda, generator=g, dtype=torch.int32)
    for call in range(2):   # the second call finds the scratch as the first one left it
        want_f = torch.full((rows, ), -7.0, device=cuda, dtype=torch.float32)
        got_f = torch.full((rows, ), 9.0, device=cuda, dtype=torch.float32)
        _call_regular(lib, SUM, F32, xf, want_f, rows, n, stride, np.float32(0).tobytes())
        _call_noalloc(lib, SUM, F32, xf, got_f, rows, n, stride, np.float32(0).tobytes())
        scale = xf.view(rows, stride)[:, :n].abs().double().sum(dim=1)
        assert torch.all((got_f.double() - want_f.double()).abs() <= 1e-5 * scale), (call, got_f, want_f)
        ref64 = xf.view(rows, stride)[:, :n].double().sum(dim=1)
        assert torch.all((got_f.double() - ref64).abs() <= 1e-5 * scale)

        for red, ident in ((SUM, 0), (MAX, -2**31)):
            want_i = torch.zeros(rows, device=cuda, dtype=torch.int32)
            got_i = torch.ones(rows, device=cuda, dtype=torch.int32)
            _call_regular(lib, red, I32, xi, want_i, rows, n, stride, np.int32(ident).tobytes())
            _call_noalloc(lib, red, I32, xi, got_i, rows, n, stride, np.int32(ident).tobytes())
            assert torch.equal(got_i, want_i), (call, red)
    torch.cuda.synchronize()


def test_noalloc_rejects_more_segments_than_the_static_scratch_serves(cuda, capfd):
    lib = _lib.load()
    x = torch.zeros(257 * 8, device=cuda)
    out = torch.zeros(257, device=cuda)
    ident = ctypes.create_string_buffer(np.float32(0).tobytes(), 8)
    st = lib.dace_b200_reduce_rows_noalloc(SUM, F32, F32, x.data_ptr(), out.data_ptr(), 257, 8, 8, ident,
                                           torch.cuda.current_stream().cuda_stream)
    assert st == 1      # cudaErrorInvalidValue
    assert 'exceed the maximum of 256' in capfd.readouterr().err
    # the limit itself is served
    st = lib.dace_b200_reduce_rows_noalloc(SUM, F32, F32, x.data_ptr(), out.data_ptr(), 256, 8, 8, ident,
                                           torch.cuda.current_stream().cuda_stream)
    assert st == 0
    torch.cuda.synchronize()
