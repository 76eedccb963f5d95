"""GPU: this library against the REFERENCE'S OWN GPU expansions on the same device buffers.

oracle/_ref/librefgpu_*.so are the unmodified reference's 'CUDA (device)' (CUB,
reference dace/libraries/standard/nodes/reduce.py:391-622), 'GPUAuto'
(reduce.py:886-1343) and generated WCR-map kernel (codegen/targets/cpu.py:965-1013)
compiled for sm_100 in the container where /root/reference exists
(oracle/build_ref.py:_gpu_programs).  Arrays are GPU_Global, so the programs take device
pointers.  Integers must agree bit for bit; floats must both be within 1e-5 of the fp64
truth (SURVEY.md 8(c) parity rules).
"""
import os

import numpy as np
import pytest

from oracle import ref_loader

pytestmark = pytest.mark.gpu

if not ref_loader.available() or 'refgpu_full_sum_f32_cub' not in ref_loader.manifest()['programs']:
    pytest.skip('oracle/_ref GPU comparison programs not present', allow_module_level=True)

ADD, MAX = 'lambda a, b: a + b', 'lambda a, b: max(a, b)'


def ref(name):
    return ref_loader.RefProgram(name)


def _run_ref(name, a, out, **syms):
    import torch
    ref(name)(A=a.data_ptr(), B=out.data_ptr(), **syms)
    torch.cuda.synchronize()


@pytest.mark.parametrize('impl', ['cub', 'gpuauto'])
def test_full_sum_matches_reference_gpu(impl, cuda):
    import torch
    import dace_b200 as db
    x = torch.rand(1 << 24, device=cuda)
    r = torch.zeros(1, device=cuda)
    _run_ref('refgpu_full_sum_f32_%s' % impl, x, r, D0=x.numel())
    mine = db.reduce(ADD, x, axis=None, identity=0)
    truth = x.double().sum().item()
    scale = x.double().abs().sum().item()
    assert abs(mine.item() - truth) <= 1e-5 * scale
    assert abs(r.item() - truth) <= 1e-4 * scale       # the reference's own error (GPUAuto: float atomics)
    assert abs(mine.item() - truth) <= abs(r.item() - truth) + 1e-6 * scale


@pytest.mark.parametrize('name,lam,ident', [('refgpu_row_sum_f32_cub', ADD, 0), ('refgpu_row_sum_f32_gpuauto', ADD, 0),
                                            ('refgpu_row_max_f32_cub', MAX, -3.4028234663852886e38)])
def test_rows_match_reference_gpu(name, lam, ident, cuda):
    import torch
    import dace_b200 as db
    x = torch.randn(3000, 4096, device=cuda)
    r = torch.zeros(3000, device=cuda)
    _run_ref(name, x, r, D0=3000, D1=4096)
    mine = db.reduce(lam, x, axis=(1, ), identity=ident)
    if lam is MAX:
        assert torch.equal(mine, r) and torch.equal(mine, x.max(1).values)
    else:
        truth, scale = x.double().sum(1), x.double().abs().sum(1)
        assert bool(((mine.double() - truth).abs() <= 1e-5 * scale).all())
        assert bool(((r.double() - truth).abs() <= 1e-5 * scale).all())


def test_column_sum_matches_reference_gpuauto(cuda):
    import torch
    import dace_b200 as db
    x = torch.rand(20000, 512, device=cuda)
    r = torch.zeros(512, device=cuda)
    _run_ref('refgpu_col_sum_f32_gpuauto', x, r, D0=20000, D1=512)
    mine = db.reduce(ADD, x, axis=(0, ), identity=0)
    truth, scale = x.double().sum(0), x.double().abs().sum(0)
    assert bool(((mine.double() - truth).abs() <= 1e-5 * scale).all())
    assert bool(((r.double() - truth).abs() <= 1e-4 * scale).all())


def test_int32_rows_bit_exact_with_reference_cub(cuda):
    import torch
    import dace_b200 as db
    x = torch.randint(-2**31, 2**31 - 1, (777, 1025), device=cuda, dtype=torch.int32)   # wraps around
    r = torch.zeros(777, device=cuda, dtype=torch.int32)
    _run_ref('refgpu_row_sum_i32_cub', x, r, D0=777, D1=1025)
    mine = db.reduce(ADD, x, axis=(1, ), identity=0)
    assert torch.equal(mine, r)


def test_bf16_rows_reference_cub_accumulates_in_bf16(cuda):
    """The reference's bf16 -> bf16 sum accumulates in bf16 (reduce.py:206); ours in fp32 and is the
    closer of the two to the fp64 truth."""
    import torch
    import dace_b200 as db
    x = torch.rand(256, 4096, device=cuda).to(torch.bfloat16)
    r = torch.zeros(256, device=cuda, dtype=torch.bfloat16)
    _run_ref('refgpu_row_sum_bf16_cub', x, r, D0=256, D1=4096)
    mine = db.reduce(ADD, x, axis=(1, ), identity=0)
    assert mine.dtype == torch.bfloat16
    truth = x.double().sum(1)
    err_mine = (mine.double() - truth).abs().max().item()
    err_ref = (r.double() - truth).abs().max().item()
    assert err_mine <= 2.0 ** -8 * truth.abs().max().item()      # one bf16 rounding of the fp32 sum
    assert err_mine <= err_ref + 1e-9


@pytest.mark.parametrize('nbins', [256, 1 << 20])
def test_histogram_bit_exact_with_reference_gpu_kernel(nbins, cuda):
    import torch
    import dace_b200 as db
    keys = torch.randint(0, nbins, (1 << 23, ), device=cuda, dtype=torch.int32)
    r = torch.zeros(nbins, device=cuda, dtype=torch.int32)
    ref('refgpu_hist_i32')(keys=keys.data_ptr(), bins=r.data_ptr(), N=keys.numel())
    torch.cuda.synchronize()
    mine = db.histogram(keys, nbins)
    assert torch.equal(mine, r)
    assert int(mine.sum().item()) == keys.numel()


def test_fused_map_reduce_vs_reference_mapreducefusion(cuda):
    """s[i] = sum_j exp(A[i, j] - mx[i]): this backend's FuseMapIntoReduce program (tree reduction in one generated
    kernel) and the reference's own MapReduceFusion program (WCR map, atomics) on the same device buffers; both
    within fp32 tolerance of the fp64 truth, ours at least as close."""
    import torch
    built = os.path.join(os.path.dirname(__file__), 'dace_e2e', '_built')
    if not ref_loader.available(built):
        pytest.skip('tests/dace_e2e/_built not present')
    a = torch.randn(2048, 8192, device=cuda)
    mx = a.max(dim=1).values.contiguous()
    truth = torch.exp(a.double() - mx.double()[:, None]).sum(1)
    s_ref = torch.full((2048, ), 7.0, device=cuda)
    s_mine = torch.full((2048, ), 7.0, device=cuda)
    torch.cuda.synchronize()
    ref('refgpu_softmax_den_mapreducefusion')(A=a.data_ptr(), mx=mx.data_ptr(), s=s_ref.data_ptr(), N=2048, M=8192)
    ref_loader.RefProgram('b200_fused_softmax_den', built)(A=a.data_ptr(), mx=mx.data_ptr(), s=s_mine.data_ptr(), N=2048, M=8192)
    torch.cuda.synchronize()
    err_ref = ((s_ref.double() - truth).abs() / truth).max().item()
    err_mine = ((s_mine.double() - truth).abs() / truth).max().item()
    assert err_mine <= 2e-5 and err_ref <= 1e-3, (err_mine, err_ref)
