"""Generate tests/golden/* by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY.  /root/reference cannot travel to the GPU box, so its
answers do: small seeded inputs + the outputs the reference produced for them.

* reduce_golden.npz - Reduce library node through the reference's Python
  frontend and hand-built SDFGs, CPU expansions `pure` / `OpenMP` / `pure-seq`
  (cases modelled on reference tests/numpy/reductions_test.py,
  tests/library/reduce_test.py:52-79, tests/reduce_offsets_test.py:26-66,
  tests/wcr_cudatest.py, tests/custom_reduce_test.py, tests/fp8_bf16_dtypes_test.py);
* wcr_golden.npz - WCR map reductions (histogram, weighted scatter-add; reference
  samples/explicit/histogram.py, tests/npbench/misc/azimint_hist_test.py);
* redtype_golden.json - dace.frontend.operations.detect_reduction_type on a
  list of lambdas (reference dace/frontend/operations.py:130-185).

* arg_golden.npz - np.argmax / np.argmin through the frontend's struct-WCR lowering
  (reference tests/numpy/reductions_test.py:58-84), generated with one OpenMP thread.

* custom_golden.npz - custom WCR lambdas through the reference's `pure` expansion (reference
  tests/custom_reduce_test.py and associative lambdas that are none of the built-in types).

Usage (from a writable cwd):  PYTHONPATH=/root/repo python -m oracle.make_golden [--arg | --custom]
"""
import json
import os
import shutil
import tempfile

import numpy as np

from .refenv import import_reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

LAMBDAS = [
    'lambda a, b: a + b', 'lambda x, y: y + x', 'lambda a, b: a * b', 'lambda a, b: max(a, b)',
    'lambda a, b: min(a, b)', 'lambda a, b: a & b', 'lambda a, b: a | b', 'lambda a, b: a ^ b',
    'lambda a, b: a and b', 'lambda a, b: a or b', 'lambda a, b: a != b', 'lambda a, b: b',
    'lambda a, b: a - b', 'lambda a, b: a / b', 'lambda a, b: a if a < b else b', 'lambda a, b: a if a > b else b',
    'lambda a, b: a * a + b', 'lambda a, b: a + b + 1', 'lambda a, b: 2 * a + b',
]


def _reduce_cases(dace):
    from oracle.build_ref import _reduce_sdfg
    rng = np.random.default_rng(42)
    add, mul = 'lambda a, b: a + b', 'lambda a, b: a * b'
    mx, mn = 'lambda a, b: max(a, b)', 'lambda a, b: min(a, b)'
    f32, f64, i32, i64, u32, u8, bf16 = (dace.float32, dace.float64, dace.int32, dace.int64, dace.uint32,
                                         dace.uint8, dace.bfloat16)
    npdt = {f32: np.float32, f64: np.float64, i32: np.int32, i64: np.int64, u32: np.uint32, u8: np.uint8}
    out = {}
    meta = []

    def run(tag, din, dout, shape, axes, wcr, identity, impl, data=None, out_init=None):
        name = 'g_%s' % tag
        sdfg = _reduce_sdfg(dace, name, din, dout, len(shape), axes, wcr, identity, impl)
        if data is None:
            if din in (f32, f64):
                data = rng.uniform(-10, 10, shape).astype(npdt[din])
            else:
                lo = 0 if din in (u32, u8) else -10
                data = rng.integers(lo, 10, shape).astype(npdt[din])
        kept = [s for i, s in enumerate(shape) if axes is not None and i not in axes] or [1]
        B = np.zeros(kept, npdt.get(dout, np.float32)) if out_init is None else out_init.copy()
        syms = {'D%d' % i: s for i, s in enumerate(shape)}
        sdfg(A=data, B=B, **syms)
        out[tag + '__in'] = data.view(np.uint16) if data.dtype.name == 'bfloat16' else data
        out[tag + '__out'] = B
        if out_init is not None:
            out[tag + '__init'] = out_init
        meta.append({'tag': tag, 'axes': axes, 'wcr': wcr, 'identity': identity, 'impl': impl,
                     'in_dtype': din.to_string(), 'out_dtype': dout.to_string()})

    fmax = float(np.finfo(np.float32).max)
    i = 0
    for impl in ('pure', 'OpenMP', 'pure-seq'):
        t = impl.replace('-', '').lower()
        run('rowsum_f32_%s' % t, f32, f32, [37, 129], [1], add, 0, impl)
        run('colsum_f32_%s' % t, f32, f32, [129, 37], [0], add, 0, impl)
        run('rowmax_f64_%s' % t, f64, f64, [20, 77], [1], mx, -1e300, impl)
        run('mid_i32_%s' % t, i32, i32, [5, 30, 7], [1], add, 0, impl)
        run('rowmin_i64_%s' % t, i64, i64, [11, 200], [1], mn, int(np.iinfo(np.int64).max), impl)
    for impl in ('pure', 'OpenMP'):
        t = impl.lower()
        run('fullsum_f64_%s' % t, f64, f64, [4097], None, add, 0, impl)
        run('fullsum_i64_%s' % t, i64, i64, [10007], None, add, 0, impl)
        run('fullprod_i64_%s' % t, i64, i64, [23], None, mul, 1, impl,
            data=rng.integers(-3, 4, [23]).astype(np.int64))
        run('fullmax_f32_%s' % t, f32, f32, [10], None, mx, -999999, impl)           # tests/wcr_cudatest.py
    run('outerinner_f64_pure', f64, f64, [6, 50, 9], [0, 2], add, 0, 'pure')           # reduce_test.py axes (0,2)
    run('mid2_f32_pure', f32, f32, [12, 5, 6, 8], [1, 2], add, 0, 'pure')              # reduce_test.py axes (1,2)
    run('lead2_f32_pure', f32, f32, [4, 16, 64], [0, 1], add, 0, 'pure')               # reduce_test.py axes (0,1)
    run('band_i32_pure', i32, i32, [9, 33], [1], 'lambda a, b: a & b', -1, 'pure',
        data=rng.integers(0, 1 << 20, [9, 33]).astype(np.int32))
    run('bxor_u32_pure', u32, u32, [9, 33], [0], 'lambda a, b: a ^ b', 0, 'pure',
        data=rng.integers(0, 1 << 31, [9, 33]).astype(np.uint32))
    run('u8sum_wrap_pure', u8, u8, [3, 300], [1], add, 0, 'pure', data=rng.integers(0, 256, [3, 300]).astype(np.uint8))
    run('nonneutral_identity_pure', i32, i32, [4, 6], [1], add, 100, 'pure')
    run('noidentity_sum_pure', i64, i64, [8, 40], [1], add, None, 'pure',
        out_init=rng.integers(-5, 5, [8]).astype(np.int64))                                # operations.py:251
    run('noidentity_max_pure', f64, f64, [40, 8], [0], mx, None, 'pure', out_init=np.full([8], 3.0))
    run('degenerate_pure', f32, f32, [3, 1, 4], [1], add, 5, 'pure')                   # reduce.py:77-87
    run('custom_min_pure', f64, f64, [100], None, 'lambda a, b: a if a < b else b', 1e30, 'pure')
    # bf16 in, fp32 out: inputs are bf16-rounded normals, stored as uint16 bit patterns.
    import torch
    xb = torch.randn(16, 128, generator=torch.Generator().manual_seed(0)).to(torch.bfloat16)
    bits = xb.view(torch.int16).numpy().view(np.uint16)
    for tag, wcr, ident in (('bf16_rowsum_pure', add, 0), ('bf16_rowmax_pure', mx, -fmax)):
        sdfg = _reduce_sdfg(dace, 'g_' + tag, bf16, f32, 2, [1], wcr, ident, 'pure')
        B = np.zeros([16], np.float32)
        A = np.empty(bits.shape, dace.bfloat16.type)   # ml_dtypes.bfloat16; a fresh array, not a view
        A.view(np.uint16)[...] = bits
        sdfg(A=A, B=B, D0=16, D1=128)
        out[tag + '__in'] = bits
        out[tag + '__out'] = B
        meta.append({'tag': tag, 'axes': [1], 'wcr': wcr, 'identity': ident, 'impl': 'pure',
                     'in_dtype': 'bfloat16', 'out_dtype': 'float32'})
    return out, meta


def _frontend_cases(dace):
    """np.sum / np.max / np.min through @dace.program (reference tests/numpy/reductions_test.py)."""
    rng = np.random.default_rng(7)
    out, meta = {}, []
    N, M = dace.symbol('N'), dace.symbol('M')

    @dace.program
    def fsum1(A: dace.float32[N, M], B: dace.float32[N]):
        B[:] = np.sum(A, axis=1)

    @dace.program
    def fmax0(A: dace.int64[N, M], B: dace.int64[M]):
        B[:] = np.max(A, axis=0)

    @dace.program
    def fminall(A: dace.float64[N, M], B: dace.float64[1]):
        B[0] = np.min(A)

    @dace.program
    def fsumneg(A: dace.int32[N, M], B: dace.int32[N]):
        B[:] = np.sum(A, axis=-1)

    for tag, prog, dt, axes, wcr, shape in (
        ('np_sum_axis1_f32', fsum1, np.float32, [1], 'lambda a, b: a + b', (64, 96)),
        ('np_max_axis0_i64', fmax0, np.int64, [0], 'lambda a, b: max(a, b)', (50, 30)),
        ('np_min_all_f64', fminall, np.float64, None, 'lambda a, b: min(a, b)', (20, 30)),
        ('np_sum_axisneg_i32', fsumneg, np.int32, [1], 'lambda a, b: a + b', (33, 65)),
    ):
        A = (rng.uniform(-10, 10, shape) if np.issubdtype(dt, np.floating) else rng.integers(-10, 10, shape)).astype(dt)
        kept = [s for i, s in enumerate(shape) if axes is not None and i not in axes] or [1]
        B = np.zeros(kept, dt)
        prog(A=A, B=B, N=shape[0], M=shape[1])
        out[tag + '__in'], out[tag + '__out'] = A, B
        if np.issubdtype(dt, np.floating):
            ident = 0 if 'sum' in tag else (float(np.finfo(dt).max) if 'min' in tag else float(np.finfo(dt).min))
        else:
            ident = 0 if 'sum' in tag else (int(np.iinfo(dt).max) if 'min' in tag else int(np.iinfo(dt).min))
        meta.append({'tag': tag, 'axes': axes, 'wcr': wcr, 'identity': ident, 'impl': 'frontend-default',
                     'in_dtype': np.dtype(dt).name, 'out_dtype': np.dtype(dt).name})
    return out, meta


def _wcr_cases(dace):
    rng = np.random.default_rng(3)
    N, NB = dace.symbol('N'), dace.symbol('NB')
    out, meta = {}, []

    @dace.program
    def hist(keys: dace.int32[N], bins: dace.int32[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                o >> bins(1, lambda x, y: x + y)[:]
                o[k] = 1

    @dace.program
    def hist8(keys: dace.uint8[N], bins: dace.uint32[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                o >> bins(1, lambda x, y: x + y)[:]
                o[k] = 1

    @dace.program
    def wadd(keys: dace.int32[N], w: dace.int64[N], bins: dace.int64[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i]
                o >> bins(1, lambda x, y: x + y)[:]
                o[k] = v

    @dace.program
    def wmax(keys: dace.int32[N], w: dace.float64[N], bins: dace.float64[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i]
                o >> bins(1, lambda x, y: max(x, y))[:]
                o[k] = v

    for tag, nb, n in (('hist_i32_256', 256, 20000), ('hist_i32_70000', 70000, 50000)):
        keys = rng.integers(0, nb, n).astype(np.int32)
        bins = rng.integers(0, 3, nb).astype(np.int32)   # WCR accumulates into existing contents
        out[tag + '__keys'], out[tag + '__init'] = keys, bins.copy()
        hist(keys=keys, bins=bins, N=n, NB=nb)
        out[tag + '__out'] = bins
        meta.append({'tag': tag, 'wcr': 'lambda x, y: x + y', 'values': None})
    keys = rng.integers(0, 256, 30000).astype(np.uint8)
    bins = np.zeros(256, np.uint32)
    out['hist_u8_u32__keys'], out['hist_u8_u32__init'] = keys, bins.copy()
    hist8(keys=keys, bins=bins, N=30000, NB=256)
    out['hist_u8_u32__out'] = bins
    meta.append({'tag': 'hist_u8_u32', 'wcr': 'lambda x, y: x + y', 'values': None})
    keys = rng.integers(0, 1000, 40000).astype(np.int32)
    w = rng.integers(-100, 100, 40000).astype(np.int64)
    bins = np.zeros(1000, np.int64)
    out['wadd_i64__keys'], out['wadd_i64__values'], out['wadd_i64__init'] = keys, w, bins.copy()
    wadd(keys=keys, w=w, bins=bins, N=40000, NB=1000)
    out['wadd_i64__out'] = bins
    meta.append({'tag': 'wadd_i64', 'wcr': 'lambda x, y: x + y', 'values': 'array'})
    wf = rng.uniform(-100, 100, 40000)
    bins = np.full(1000, -1e9)
    out['wmax_f64__keys'], out['wmax_f64__values'], out['wmax_f64__init'] = keys, wf, bins.copy()
    wmax(keys=keys, w=wf, bins=bins, N=40000, NB=1000)
    out['wmax_f64__out'] = bins
    meta.append({'tag': 'wmax_f64', 'wcr': 'lambda x, y: max(x, y)', 'values': 'array'})
    return out, meta


def _arg_cases(dace):
    """np.argmax / np.argmin through @dace.program (reference tests/numpy/reductions_test.py:58-84,
    tests/numpy/ndarray_attributes_methods_test.py:95-111).  The frontend lowers them to a struct WCR map
    (reference dace/frontend/python/replacements/reduction.py:261-348) that the CPU code generator runs as an
    OpenMP loop with a critical section per element, so ties are resolved in whatever order the threads arrive.
    The vectors are generated with OMP_NUM_THREADS=1: index order, which is also NumPy's first-occurrence rule."""
    rng = np.random.default_rng(11)
    out, meta = {}, []
    N, M, K = dace.symbol('N'), dace.symbol('M'), dace.symbol('K')

    def make(func, dt, axis, ndim):
        shape = [N, M, K][:ndim]
        fn = np.argmax if func == 'max' else np.argmin
        if axis is None:
            @dace.program
            def prog(A: dt[tuple(shape)]):
                return fn(A)
        else:
            @dace.program
            def prog(A: dt[tuple(shape)]):
                return fn(A, axis=axis)
        return prog

    f32, f64, i32, i64 = dace.float32, dace.float64, dace.int32, dace.int64
    npdt = {f32: np.float32, f64: np.float64, i32: np.int32, i64: np.int64}
    cases = [
        ('argmax_f64_axis1', 'max', f64, 1, (10, 5, 3), 'uniform'),
        ('argmin_f64_axis1', 'min', f64, 1, (10, 5, 3), 'uniform'),
        ('argmin_i32_axis1', 'min', i32, 1, (10, 5, 3), 'ties'),
        ('argmax_i32_axis1', 'max', i32, 1, (10, 5, 3), 'ties'),
        ('argmin_i64_axis1', 'min', i64, 1, (10, 5, 3), 'ties'),
        ('argmax_i64_axis1', 'max', i64, 1, (10, 5, 3), 'ties'),
        ('argmax_f32_flat', 'max', f32, None, (7, 9, 11), 'uniform'),
        ('argmin_f32_flat', 'min', f32, None, (7, 9, 11), 'uniform'),
        ('argmax_f32_axis0', 'max', f32, 0, (300, 40), 'uniform'),
        ('argmin_f32_last', 'min', f32, 1, (40, 300), 'uniform'),
        ('argmax_i32_ties_rows', 'max', i32, 1, (33, 500), 'ties'),
        ('argmax_i32_all_min', 'max', i32, 1, (4, 50), 'allmin'),      # nothing above min_value -> index -1
        ('argmax_f32_neg_inf', 'max', f32, 0, (50, 6), 'neginf'),      # -inf never beats -finfo.max -> -1
        ('argmin_f64_all_equal', 'min', f64, 1, (5, 64), 'equal'),
    ]
    for tag, func, dt, axis, shape, kind in cases:
        nd = npdt[dt]
        if kind == 'uniform':
            A = rng.uniform(-10, 10, shape).astype(nd)
        elif kind == 'ties':
            A = rng.integers(-3, 4, shape).astype(nd)
        elif kind == 'allmin':
            A = np.full(shape, np.iinfo(nd).min, nd)
            A[1, 7] = 5
        elif kind == 'neginf':
            A = np.full(shape, -np.inf, nd)
            A[10:12, 2] = 1.5
            A[3, 4] = np.finfo(nd).min
        else:
            A = np.full(shape, 2.5, nd)
        prog = make(func, dt, axis, len(shape))
        syms = dict(zip('NMK', shape))
        res = prog(A=A, **syms)
        out[tag + '__in'] = A
        out[tag + '__out'] = np.asarray(res).copy()
        meta.append({'tag': tag, 'func': func, 'axis': axis, 'dtype': np.dtype(nd).name, 'kind': kind})
    return out, meta


# Custom WCR lambdas through the reference's `pure` expansion (the only CPU expansion that takes them,
# reference reduce.py:27-154; CPU lowering = wcr_custom CAS loops, reduction.h:56-164).  The lambdas are
# associative and commutative (every DaCe WCR must be), inputs avoid ties, so the results do not depend on the
# thread interleaving.  Shared with tests/dace_e2e/build_programs.py, which builds the same SDFGs with 'B200'.
CUSTOM_CASES = [
    # tag, dtype, shape, axes, lambda, identity
    ('app_i64_rows', 'int64', [33, 257], [1], 'lambda a, b: a + b + a * b', 0),
    ('app_i64_cols', 'int64', [257, 33], [0], 'lambda a, b: a + b + a * b', 0),
    ('app_i64_full', 'int64', [5001], None, 'lambda a, b: a + b + a * b', 0),
    ('app_i64_lead2', 'int64', [9, 31, 40], [0, 1], 'lambda a, b: a + b + a * b', 7),
    ('absmax_f32_rows', 'float32', [20, 300], [1], 'lambda a, b: a if abs(a) > abs(b) else b', 0),
    ('absmax_f32_noident', 'float32', [300, 20], [0], 'lambda a, b: a if abs(a) > abs(b) else b', None),
    ('refmin_f32_full', 'float32', [20], None, 'lambda a, b: a if a < b else b', 9999999),     # tests/custom_reduce_test.py:7-9
]


def custom_inputs(tag, dtype, shape):
    rng = np.random.default_rng(sum(map(ord, tag)))
    if dtype == 'int64':
        return (rng.integers(-3, 4, shape) * 2).astype(np.int64)        # 1 + x is odd: prod(1 + x) never collapses to 0
    return rng.uniform(-100, 100, shape).astype(np.float32)


def _custom_cases(dace):
    from oracle.build_ref import _reduce_sdfg
    out, meta = {}, []
    for tag, dtype, shape, axes, wcr, identity in CUSTOM_CASES:
        dt = {'int64': dace.int64, 'float32': dace.float32}[dtype]
        sdfg = _reduce_sdfg(dace, 'gc_' + tag, dt, dt, len(shape), axes, wcr, identity, 'pure')
        data = custom_inputs(tag, dtype, shape)
        kept = [s for i, s in enumerate(shape) if axes is not None and i not in axes] or [1]
        init = None
        if identity is None:
            init = np.random.default_rng(1).uniform(-50, 50, kept).astype(data.dtype)
        B = np.zeros(kept, data.dtype) if init is None else init.copy()
        sdfg(A=data, B=B, **{'D%d' % i: s for i, s in enumerate(shape)})
        out[tag + '__in'] = data
        out[tag + '__out'] = B
        if init is not None:
            out[tag + '__init'] = init
        meta.append({'tag': tag, 'dtype': dtype, 'shape': shape, 'axes': axes, 'wcr': wcr, 'identity': identity,
                     'impl': 'pure'})
    return out, meta


def main_custom():
    tmp = tempfile.mkdtemp(prefix='dace_b200_golden_')
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        dace = import_reference(os.path.join(tmp, '.dacecache'))
        assert dace is not None, 'reference checkout not available'
        o, m = _custom_cases(dace)
        np.savez_compressed(os.path.join(GOLDEN, 'custom_golden.npz'), **o)
        with open(os.path.join(GOLDEN, 'custom_golden_meta.json'), 'w') as f:
            json.dump({'reference_version': dace.__version__, 'custom': m}, f, indent=1)
        print('wrote custom-lambda golden vectors to', GOLDEN)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp, ignore_errors=True)


def main_arg():
    os.environ['OMP_NUM_THREADS'] = '1'
    tmp = tempfile.mkdtemp(prefix='dace_b200_golden_')
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        dace = import_reference(os.path.join(tmp, '.dacecache'))
        assert dace is not None, 'reference checkout not available'
        o, m = _arg_cases(dace)
        np.savez_compressed(os.path.join(GOLDEN, 'arg_golden.npz'), **o)
        with open(os.path.join(GOLDEN, 'arg_golden_meta.json'), 'w') as f:
            json.dump({'reference_version': dace.__version__, 'omp_num_threads': 1, 'arg': m}, f, indent=1)
        print('wrote arg golden vectors to', GOLDEN)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp, ignore_errors=True)


def main():
    tmp = tempfile.mkdtemp(prefix='dace_b200_golden_')
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        dace = import_reference(os.path.join(tmp, '.dacecache'))
        assert dace is not None, 'reference checkout not available'
        from dace.frontend.operations import detect_reduction_type
        os.makedirs(GOLDEN, exist_ok=True)
        red = {}
        for lam in LAMBDAS:
            red[lam] = {'default': str(detect_reduction_type(lam)).split('.')[-1],
                        'openmp': str(detect_reduction_type(lam, openmp=True)).split('.')[-1]}
        with open(os.path.join(GOLDEN, 'redtype_golden.json'), 'w') as f:
            json.dump({'reference_version': dace.__version__, 'lambdas': red}, f, indent=1)
        o1, m1 = _reduce_cases(dace)
        o2, m2 = _frontend_cases(dace)
        o1.update(o2)
        np.savez_compressed(os.path.join(GOLDEN, 'reduce_golden.npz'), **o1)
        o3, m3 = _wcr_cases(dace)
        np.savez_compressed(os.path.join(GOLDEN, 'wcr_golden.npz'), **o3)
        with open(os.path.join(GOLDEN, 'golden_meta.json'), 'w') as f:
            json.dump({'reference_version': dace.__version__, 'reduce': m1 + m2, 'wcr': m3}, f, indent=1)
        print('wrote golden vectors to', GOLDEN)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == '__main__':
    import sys
    if '--arg' in sys.argv:      # only the argmin/argmax vectors (the other files stay byte-identical)
        main_arg()
    elif '--custom' in sys.argv:  # only the custom-lambda vectors
        main_custom()
    else:
        main()
        main_arg()
        main_custom()
