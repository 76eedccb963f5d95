"""GPU: include/dace_b200_fused_reduce.cuh (map -> reduce in one kernel with a load functor) through the
native harness tests/native/fused_reduce_harness.cu, against NumPy in fp64 / exact integers.  Shapes cover the
warp-per-row kernel, CTA-per-row, split rows (+ finish kernel through the workspace) and > 65535 rows."""
import ctypes
import os

import numpy as np
import pytest
import torch

import dace_b200 as db

pytestmark = pytest.mark.gpu

HARNESS = os.path.join(os.path.dirname(__file__), 'native', '_built', 'libfused_reduce_harness.so')
SHAPES = [(1, 1), (1, 5000001), (3, 100000), (7, 4097), (600, 5000), (5000, 300), (70000, 40), (300, 2047), (2, 0)]


@pytest.fixture(scope='module')
def lib():
    if not os.path.exists(HARNESS):
        pytest.fail('tests/native/_built/libfused_reduce_harness.so missing: run __graft_entry__.build()')
    h = ctypes.CDLL(HARNESS)
    h.fused_reduce_harness_run.restype = ctypes.c_int
    h.fused_reduce_harness_run.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong,
                                           ctypes.c_ulonglong, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_void_p,
                                           ctypes.c_ulonglong, ctypes.c_void_p]
    return h


def _run(lib, case, in0, in1, out, rows, n, use_identity, ws):
    rc = lib.fused_reduce_harness_run(case, in0.data_ptr(), in1.data_ptr() if in1 is not None else None, out.data_ptr(),
                                      out.stride(0) if out.dim() else 1, rows, n, 1 if use_identity else 0,
                                      ws.ptr if ws is not None else None, ws.nbytes if ws is not None else 0,
                                      torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()


@pytest.mark.parametrize('rows,n', SHAPES)
def test_fused_exp_shift_sum_f32(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows * 31 + n)
    x = torch.randn(rows, max(n, 1), device=cuda, generator=g)[:, :n].contiguous()
    m = x.max(dim=1).values if n else torch.zeros(rows, device=cuda)
    out = torch.full((rows, ), 3.0, device=cuda)
    _run(lib, 0, x, m, out, rows, n, True, ws)
    ref = torch.exp(x.double() - m.double()[:, None]).sum(1)
    assert torch.allclose(out.double(), ref, rtol=2e-5, atol=1e-30)
    # identity None: fold into the existing output; and no workspace: still right (no row splitting)
    out2 = torch.full((rows, ), 3.0, device=cuda)
    _run(lib, 0, x, m, out2, rows, n, False, None)
    assert torch.allclose(out2.double(), ref + 3.0, rtol=2e-5)


@pytest.mark.parametrize('rows,n', SHAPES)
def test_fused_integer_cases_exact(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(n + 5)
    a = torch.randint(-1000, 1000, (rows, n), device=cuda, dtype=torch.int32, generator=g)
    b = torch.randint(-1000, 1000, (rows, n), device=cuda, dtype=torch.int32, generator=g)
    out = torch.zeros(rows, dtype=torch.int32, device=cuda)
    _run(lib, 2, a, b, out, rows, n, True, ws)
    ref = (a - b).abs().max(dim=1).values.clamp(min=-5) if n else torch.full((rows, ), -5, dtype=torch.int32, device=cuda)
    assert torch.equal(out, ref.to(torch.int32))
    al = torch.randint(-10**6, 10**6, (rows, n), device=cuda, dtype=torch.int64, generator=g)
    w = torch.randint(-10**6, 10**6, (n, ), device=cuda, dtype=torch.int64, generator=g)
    outl = torch.zeros(rows * 2, dtype=torch.int64, device=cuda)[::2]           # strided output
    _run(lib, 3, al, w, outl, rows, n, True, ws)
    assert torch.equal(outl, (al * w[None, :]).sum(1) + 11)


def test_fused_row_norms_f64_deterministic(cuda, lib):
    ws = db.Workspace(cuda)
    x = torch.randn(5, 3000017, device=cuda, dtype=torch.float64)
    o1, o2 = torch.zeros(5, dtype=torch.float64, device=cuda), torch.zeros(5, dtype=torch.float64, device=cuda)
    _run(lib, 1, x, None, o1, 5, x.shape[1], True, ws)
    _run(lib, 1, x, None, o2, 5, x.shape[1], True, ws)
    assert torch.equal(o1, o2)                                   # fixed combine order
    assert torch.allclose(o1, (x * x).sum(1), rtol=1e-12)


# ---- column folds: out[c] = fold over r of load(r, c) ----

COL_SHAPES = [(1, 1), (5000001, 1), (100000, 3), (4097, 40), (5000, 600), (300, 5000), (64, 70000), (1000, 4096), (0, 5)]


@pytest.mark.parametrize('rows,n', COL_SHAPES)
def test_fused_column_folds(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows * 7 + n)
    x = torch.randn(max(rows, 1), n, device=cuda, generator=g)[:rows].contiguous()
    m = torch.randn(max(rows, 1), device=cuda, generator=g)[:rows].contiguous() * 0.1
    out = torch.full((n, ), 3.0, device=cuda)
    _run(lib, 10, x, m, out, rows, n, True, ws)
    ref = torch.exp(x.double() - m.double()[:, None]).sum(0)
    assert torch.allclose(out.double(), ref, rtol=2e-5, atol=1e-30)
    out2 = torch.full((n, ), 3.0, device=cuda)
    _run(lib, 10, x, m, out2, rows, n, False, None)                 # identity None, no workspace (no row slices)
    assert torch.allclose(out2.double(), ref + 3.0, rtol=2e-5)
    a = torch.randint(-1000, 1000, (rows, n), device=cuda, dtype=torch.int32, generator=g)
    b = torch.randint(-1000, 1000, (rows, n), device=cuda, dtype=torch.int32, generator=g)
    oi = torch.zeros(n, dtype=torch.int32, device=cuda)
    _run(lib, 12, a, b, oi, rows, n, True, ws)
    refi = (a - b).abs().max(dim=0).values.clamp(min=-5) if rows else torch.full((n, ), -5, dtype=torch.int32, device=cuda)
    assert torch.equal(oi, refi.to(torch.int32))
    al = torch.randint(-10**6, 10**6, (rows, n), device=cuda, dtype=torch.int64, generator=g)
    w = torch.randint(-10**6, 10**6, (n, ), device=cuda, dtype=torch.int64, generator=g)
    ol = torch.zeros(n * 2, dtype=torch.int64, device=cuda)[::2]
    _run(lib, 13, al, w, ol, rows, n, True, ws)
    assert torch.equal(ol, (al * w[None, :]).sum(0) + 11)


# ---- custom WCR functors (no neutral element: accumulators carry a "holds a value" flag) ----

def _add_plus_product_ref(x: torch.Tensor, dim: int, identity: int) -> np.ndarray:
    """fold of a + b + a*b = prod(1 + x_i) - 1, exact modulo 2^64 like the device's int64 arithmetic."""
    f = (x.cpu().numpy().astype(np.int64) + 1).astype(np.uint64)
    with np.errstate(over='ignore'):
        p = np.multiply.reduce(f, axis=dim, dtype=np.uint64) * np.uint64(identity + 1)
        return (p - np.uint64(1)).astype(np.int64)


@pytest.mark.parametrize('rows,n', [(1, 20), (1, 3000001), (3, 100000), (600, 5000), (5000, 300), (70000, 40), (2, 0)])
def test_fused_custom_functors_rows(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows + n)
    x = torch.rand(rows, n, device=cuda, generator=g) * 100 + 1
    out = torch.full((rows, ), -1.0, device=cuda)
    _run(lib, 20, x, None, out, rows, n, True, ws)                     # lambda a, b: a if a < b else b, identity 9999999
    ref = x.min(dim=1).values if n else torch.full((rows, ), 9999999.0, device=cuda)
    assert torch.equal(out, ref)
    out2 = torch.full((rows, ), 50.0, device=cuda)
    _run(lib, 20, x, None, out2, rows, n, False, None)                 # identity None: fold into the output
    assert torch.equal(out2, torch.minimum(ref, torch.tensor(50.0, device=cuda)))
    xi = torch.randint(-3, 4, (rows, n), device=cuda, dtype=torch.int64, generator=g) * 2     # 1 + x is odd
    oi = torch.zeros(rows, dtype=torch.int64, device=cuda)
    _run(lib, 21, xi, None, oi, rows, n, True, ws)                     # lambda a, b: a + b + a * b, identity 0
    np.testing.assert_array_equal(oi.cpu().numpy(), _add_plus_product_ref(xi, 1, 0))


@pytest.mark.parametrize('rows,n', [(20, 1), (3000001, 1), (100000, 3), (5000, 600), (300, 5000), (64, 70000)])
def test_fused_custom_functors_cols(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows * 3 + n)
    x = torch.rand(rows, n, device=cuda, generator=g) * 100 + 1
    out = torch.full((n, ), -1.0, device=cuda)
    _run(lib, 30, x, None, out, rows, n, True, ws)
    assert torch.equal(out, x.min(dim=0).values)
    xi = torch.randint(-3, 4, (rows, n), device=cuda, dtype=torch.int64, generator=g) * 2
    oi = torch.zeros(n, dtype=torch.int64, device=cuda)
    _run(lib, 31, xi, None, oi, rows, n, True, ws)
    np.testing.assert_array_equal(oi.cpu().numpy(), _add_plus_product_ref(xi, 0, 0))


# ---- vector-load functors (kVec / vec(), see VECTOR LOADS in the header) ----

VEC_ROW_SHAPES = [(1, 8), (1, 5000000), (3, 100000), (7, 4104), (600, 5000), (5000, 304), (70000, 40), (300000, 128),
                  (40000, 512), (3000, 2040), (300, 2047), (11, 64)]
VEC_COL_SHAPES = [(1, 8), (100000, 8), (4097, 40), (300, 5000), (64, 70000), (1000, 4096), (5, 19000), (33, 1 << 20),
                  (700, 40000), (9, 1200001)]


@pytest.mark.parametrize('rows,n', VEC_ROW_SHAPES)
@pytest.mark.parametrize('case', [40, 41])
def test_fused_vector_loads_rows_f32(cuda, lib, case, rows, n):
    """Same results as the scalar functor, to fp32 rounding (the grouping of the additions differs); odd n and
    misaligned operands fall back to the scalar kernels inside the same call."""
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows * 31 + n)
    buf = torch.randn(rows * n + 8, device=cuda, generator=g)
    for off in (0, 1):                                              # off = 1: 4 bytes off the vector alignment
        x = buf[off:off + rows * n].view(rows, n)
        m = x.max(dim=1).values
        out = torch.full((rows, ), 3.0, device=cuda)
        _run(lib, case, x, m, out, rows, n, True, ws)
        ref = torch.exp(x.double() - m.double()[:, None]).sum(1)
        assert torch.allclose(out.double(), ref, rtol=2e-5, atol=1e-30)
        out2 = torch.full((rows, ), 3.0, device=cuda)
        _run(lib, case, x, m, out2, rows, n, False, None)
        assert torch.allclose(out2.double(), ref + 3.0, rtol=2e-5)


@pytest.mark.parametrize('rows,n', VEC_ROW_SHAPES)
def test_fused_vector_loads_rows_int64_exact(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(n + 5)
    al = torch.randint(-10**6, 10**6, (rows, n), device=cuda, dtype=torch.int64, generator=g)
    w = torch.randint(-10**6, 10**6, (n, ), device=cuda, dtype=torch.int64, generator=g)
    outl = torch.zeros(rows * 2, dtype=torch.int64, device=cuda)[::2]
    _run(lib, 42, al, w, outl, rows, n, True, ws)
    assert torch.equal(outl, (al * w[None, :]).sum(1) + 11)
    xi = torch.randint(-3, 4, (rows, n), device=cuda, dtype=torch.int64, generator=g) * 2
    oi = torch.zeros(rows, dtype=torch.int64, device=cuda)
    _run(lib, 43, xi, None, oi, rows, n, True, ws)                     # custom functor + vector loads
    np.testing.assert_array_equal(oi.cpu().numpy(), _add_plus_product_ref(xi, 1, 0))


@pytest.mark.parametrize('rows,n', VEC_COL_SHAPES)
def test_fused_vector_loads_cols(cuda, lib, rows, n):
    ws = db.Workspace(cuda)
    g = torch.Generator(device=cuda).manual_seed(rows * 7 + n)
    x = torch.randn(rows, n, device=cuda, generator=g)
    m = torch.randn(rows, device=cuda, generator=g) * 0.1
    ref = torch.exp(x.double() - m.double()[:, None]).sum(0)
    for case in (50, 51):
        out = torch.full((n, ), 3.0, device=cuda)
        _run(lib, case, x, m, out, rows, n, True, ws)
        assert torch.allclose(out.double(), ref, rtol=2e-5, atol=1e-30)
        out2 = torch.full((n, ), 3.0, device=cuda)
        _run(lib, case, x, m, out2, rows, n, False, None)
        assert torch.allclose(out2.double(), ref + 3.0, rtol=2e-5)
    al = torch.randint(-10**6, 10**6, (rows, n), device=cuda, dtype=torch.int64, generator=g)
    w = torch.randint(-10**6, 10**6, (n, ), device=cuda, dtype=torch.int64, generator=g)
    ol = torch.zeros(n * 2, dtype=torch.int64, device=cuda)[::2]
    _run(lib, 52, al, w, ol, rows, n, True, ws)
    assert torch.equal(ol, (al * w[None, :]).sum(0) + 11)
    xi = torch.randint(-3, 4, (rows, n), device=cuda, dtype=torch.int64, generator=g) * 2
    oi = torch.zeros(n, dtype=torch.int64, device=cuda)
    _run(lib, 53, xi, None, oi, rows, n, True, ws)
    np.testing.assert_array_equal(oi.cpu().numpy(), _add_plus_product_ref(xi, 0, 0))
