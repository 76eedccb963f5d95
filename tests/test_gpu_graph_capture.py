"""The compute entry points neither allocate nor synchronise (include/dace_b200_reduce.h): in a FRESH
process - nothing warmed up, no scratch grown by an earlier call - ``dace_b200_reduce_nd`` on a two-pass
shape and ``dace_b200_wcr_scatter`` on the radix-partition path (2^24 keys, 2^20 bins) are captured into a
CUDA graph (torch's capture runs in cudaStreamCaptureModeGlobal: any cudaMalloc / cudaFree / synchronise
from any thread invalidates it) and replayed; results are checked bit for bit against torch.
"""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = textwrap.dedent('''
    import sys
    sys.path.insert(0, %r)
    import torch
    import dace_b200 as db
    from dace_b200 import _lib
    from dace_b200.api import Workspace
    import ctypes

    dev = torch.device('cuda:0')
    lib = _lib.load()
    g = torch.Generator(device=dev); g.manual_seed(5)
    # [R1, K, R2] = [64, 48, 4099] int32, axes (0, 2): rows pass into mid[R1, K], column pass over mid
    x = torch.randint(-50, 50, (64, 48, 4099), dtype=torch.int32, device=dev, generator=g)
    out = torch.zeros(48, dtype=torch.int32, device=dev)
    shape = (ctypes.c_int64 * 3)(*x.shape); istr = (ctypes.c_int64 * 3)(*x.stride())
    need = lib.dace_b200_reduce_nd_workspace_bytes(2, 2, 3, shape, istr, 0b101)
    assert need > lib.dace_b200_workspace_bytes(), 'the shape must have a two-pass plan'
    ws = Workspace(dev, nbytes=need)
    keys = torch.randint(0, 1 << 20, (1 << 24, ), dtype=torch.int32, device=dev, generator=g)
    bins = torch.zeros(1 << 20, dtype=torch.int32, device=dev)
    sneed = lib.dace_b200_wcr_scatter_workspace_bytes(3, 2, 2, 0, keys.numel(), bins.numel())
    assert sneed > 0, 'the histogram must take the partitioned path'
    scratch = Workspace(dev, nbytes=sneed, zero_init=False)
    ws2 = Workspace(dev)
    big = torch.randint(-9, 9, ((1 << 24) + 3, ), dtype=torch.int32, device=dev, generator=g)
    tot = torch.zeros((), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(graph, stream=side):          # first use of every kernel happens under capture
        db.reduce('lambda a, b: a + b', x, out, axis=(0, 2), identity=0, workspace=ws)
        db.wcr_scatter('lambda a, b: a + b', bins, keys, workspace=scratch)
        db.reduce('lambda a, b: a + b', big, tot, axis=None, identity=0, workspace=ws2)
    torch.cuda.synchronize()
    assert int(bins.sum()) == 0, 'capture must not execute'
    for rep in (1, 2, 3):
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, x.sum(dim=(0, 2), dtype=torch.int32)), 'two-pass reduce under graph replay'
        assert torch.equal(bins, rep * torch.bincount(keys, minlength=1 << 20).to(torch.int32)), 'histogram replay %%d' %% rep
        assert int(tot) == int(big.sum(dtype=torch.int32)), 'full reduce under graph replay'
    print('graph capture ok')
''') % ROOT


def test_compute_entry_points_capture_into_a_cuda_graph_in_a_fresh_process(cuda):
    res = subprocess.run([sys.executable, '-c', SCRIPT], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and 'graph capture ok' in res.stdout, res.stdout[-2000:] + res.stderr[-4000:]


def test_a_missing_or_small_workspace_is_rejected_not_allocated(cuda):
    """The library never allocates a workspace behind the caller's back (round-1 behaviour: a per-device default)."""
    import ctypes
    import torch
    from dace_b200 import _lib
    lib = _lib.load()
    x = torch.ones(1 << 20, device=cuda)
    out = torch.zeros(1, device=cuda)
    ident = ctypes.create_string_buffer(b'\0' * 8, 8)
    st = lib.dace_b200_reduce_full(3, 0, 0, x.data_ptr(), out.data_ptr(), x.numel(), ident, None, 0, None)
    assert st == 1     # cudaErrorInvalidValue
    small = torch.zeros(1024, dtype=torch.uint8, device=cuda)
    st = lib.dace_b200_reduce_full(3, 0, 0, x.data_ptr(), out.data_ptr(), x.numel(), ident, small.data_ptr(), 1024, None)
    assert st == 1
    # scatter: no workspace = the global-atomic kernel, still correct
    keys = torch.randint(0, 1 << 20, (1 << 22, ), dtype=torch.int32, device=cuda)
    bins = torch.zeros(1 << 20, dtype=torch.int32, device=cuda)
    st = lib.dace_b200_wcr_scatter(3, 2, 2, keys.data_ptr(), None, None, keys.numel(), bins.data_ptr(), bins.numel(),
                                   None, 0, torch.cuda.current_stream().cuda_stream)
    assert st == 0
    assert torch.equal(bins, torch.bincount(keys, minlength=1 << 20).to(torch.int32))
