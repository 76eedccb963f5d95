"""Compile the REFERENCE's own CPU expansions of the hot path into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Runs only where /root/reference exists (this
container).  It imports the unmodified reference (through oracle/refenv.py),
builds the SDFGs of the benchmark configs - an ``@dace.program`` with
``np.sum(A, axis=1)`` (BASELINE C1), hand-built ``state.add_reduce`` SDFGs
(reference tests/library/reduce_test.py:14-27) and the WCR histogram map
(reference samples/explicit/histogram.py:12-26) - forces
``Reduce.implementation`` to the reference's CPU expansions (`pure`, `OpenMP`,
`pure-seq`), lets the reference generate C++ and compile it with its own CMake
recipe, and copies ONLY the resulting shared objects (never sources) into
oracle/_ref/ together with a manifest describing each entry point's arguments.
The .so files depend on libgomp/libstdc++ only, so they travel to the GPU box
and are loaded there with ctypes by oracle/ref_loader.py - as the checker and as
the timed CPU baseline (cpu_baseline.kind == "reference"), never as product code.

Usage:  python -m oracle.build_ref
"""
import json
import os
import shutil
import sys
import tempfile

from .refenv import import_reference, reference_available  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, '_ref')
MANIFEST = os.path.join(REF_DIR, 'manifest.json')
VERSION = 6


def up_to_date() -> bool:
    try:
        with open(MANIFEST) as f:
            m = json.load(f)
        return m.get('version') == VERSION and all(
            os.path.exists(os.path.join(REF_DIR, e['lib'])) for e in m['programs'].values())
    except Exception:
        return False


def _reduce_sdfg(dace, name, in_dtype, out_dtype, ndim, axes, wcr, identity, impl):
    """Hand-built Reduce SDFG with symbolic int64 sizes D0..D{ndim-1}."""
    syms = [dace.symbol('D%d' % i, dace.int64) for i in range(ndim)]
    sdfg = dace.SDFG(name)
    sdfg.add_array('A', syms, in_dtype)
    kept = [s for i, s in enumerate(syms) if axes is not None and i not in axes]
    sdfg.add_array('B', kept if kept else [1], out_dtype)
    state = sdfg.add_state()
    a, b = state.add_read('A'), state.add_write('B')
    red = state.add_reduce(wcr, axes, identity)
    red.implementation = impl
    state.add_nedge(a, red, dace.Memlet.from_array('A', sdfg.arrays['A']))
    state.add_nedge(red, b, dace.Memlet.from_array('B', sdfg.arrays['B']))
    return sdfg


def _programs(dace):
    import numpy as np
    from dace.libraries.standard.nodes.reduce import Reduce
    add, mx, mn = 'lambda a, b: a + b', 'lambda a, b: max(a, b)', 'lambda a, b: min(a, b)'
    progs = []

    # C1: the @dace.program of BASELINE config 0, Reduce forced to each CPU expansion.
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)
    for impl, tag in (('pure', 'pure'), ('OpenMP', 'omp'), ('pure-seq', 'pureseq')):

        @dace.program
        def rowsum(A: dace.float32[N, M], B: dace.float32[N]):
            B[:] = np.sum(A, axis=1)

        sdfg = rowsum.to_sdfg()
        sdfg.name = 'c1_npsum_axis1_f32_' + tag
        for n, _ in sdfg.all_nodes_recursive():
            if isinstance(n, Reduce):
                n.implementation = impl
        progs.append((sdfg, {'kind': 'reduce', 'impl': impl, 'frontend': 'np.sum(A, axis=1)'}))

    f32, f64, i32, i64, bf16 = dace.float32, dace.float64, dace.int32, dace.int64, dace.bfloat16
    fmax = float(np.finfo(np.float32).max)
    spec = [
        # name, in, out, ndim, axes, wcr, identity, impls
        ('full_sum_f32', f32, f32, 1, None, add, 0, ('pure', 'OpenMP')),
        ('full_sum_f64', f64, f64, 1, None, add, 0, ('pure', 'OpenMP')),
        ('full_max_f32', f32, f32, 1, None, mx, -fmax, ('pure', 'OpenMP')),
        ('full_sum_i64', i64, i64, 1, None, add, 0, ('pure', 'OpenMP')),
        ('row_sum_f32', f32, f32, 2, [1], add, 0, ('pure', 'OpenMP', 'pure-seq')),
        ('row_max_f32', f32, f32, 2, [1], mx, -fmax, ('pure', 'OpenMP', 'pure-seq')),
        ('row_sum_i32', i32, i32, 2, [1], add, 0, ('pure', 'OpenMP')),
        ('row_min_i64', i64, i64, 2, [1], mn, int(np.iinfo(np.int64).max), ('pure', 'OpenMP')),
        ('col_sum_f32', f32, f32, 2, [0], add, 0, ('pure', 'OpenMP', 'pure-seq')),
        ('col_sum_i32', i32, i32, 2, [0], add, 0, ('pure', )),
        ('mid_sum_f32', f32, f32, 3, [1], add, 0, ('pure', 'OpenMP')),
        ('outer_inner_sum_f64', f64, f64, 3, [0, 2], add, 0, ('pure', 'pure-seq')),
        ('mid2_sum_f32', f32, f32, 4, [1, 2], add, 0, ('pure', )),
        # C3: bf16 input, fp32 output/accumulation - only `pure` compiles for both ops
        ('row_sum_bf16_f32', bf16, f32, 2, [1], add, 0, ('pure', )),
        ('row_max_bf16_f32', bf16, f32, 2, [1], mx, -fmax, ('pure', )),
    ]
    tags = {'pure': 'pure', 'OpenMP': 'omp', 'pure-seq': 'pureseq'}
    for name, din, dout, nd, axes, wcr, ident, impls in spec:
        for impl in impls:
            sdfg = _reduce_sdfg(dace, '%s_%s' % (name, tags[impl]), din, dout, nd, axes, wcr, ident, impl)
            progs.append((sdfg, {'kind': 'reduce', 'impl': impl, 'axes': axes, 'wcr': wcr, 'identity': ident}))

    # C4: WCR map histogram, keys -> bins.
    NB = dace.symbol('NB', dace.int64)

    @dace.program
    def hist_i32(keys: dace.int32[N], bins: dace.int32[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = 1

    sdfg = hist_i32.to_sdfg()
    sdfg.name = 'c4_hist_i32'
    progs.append((sdfg, {'kind': 'wcr_map', 'impl': 'wcr-omp-atomic'}))

    @dace.program
    def hist_u8_u32(keys: dace.uint8[N], bins: dace.uint32[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = 1

    sdfg = hist_u8_u32.to_sdfg()
    sdfg.name = 'hist_u8_u32'
    progs.append((sdfg, {'kind': 'wcr_map', 'impl': 'wcr-omp-atomic'}))

    @dace.program
    def scatter_add_f64(keys: dace.int32[N], w: dace.float64[N], bins: dace.float64[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = v

    sdfg = scatter_add_f64.to_sdfg()
    sdfg.name = 'scatter_add_f64'
    progs.append((sdfg, {'kind': 'wcr_map', 'impl': 'wcr-omp-atomic'}))
    return progs


def _gpu_programs(dace):
    """The reference's own GPU lowerings of the path, on device-resident (GPU_Global) arrays, built
    for sm_100: the on-box GPU comparison SURVEY.md 8(d) asks for (C2 "vs the reference CUB
    expansion").  Expansions: 'CUDA (device)' = CUB (reference reduce.py:391-622), 'GPUAuto'
    (reduce.py:886-1343), and the generated one-atomic-per-key WCR kernel (cpu.py:965-1013)."""
    add, mx = 'lambda a, b: a + b', 'lambda a, b: max(a, b)'
    f32, i32 = dace.float32, dace.int32
    G = dace.StorageType.GPU_Global
    progs = []
    fmax = 3.4028234663852886e38
    for name, din, dout, nd, axes, wcr, ident, impl in (
        ('refgpu_full_sum_f32_cub', f32, f32, 1, None, add, 0, 'CUDA (device)'),
        ('refgpu_full_sum_f32_gpuauto', f32, f32, 1, None, add, 0, 'GPUAuto'),
        ('refgpu_row_sum_f32_cub', f32, f32, 2, [1], add, 0, 'CUDA (device)'),
        ('refgpu_row_sum_f32_gpuauto', f32, f32, 2, [1], add, 0, 'GPUAuto'),
        ('refgpu_row_max_f32_cub', f32, f32, 2, [1], mx, -fmax, 'CUDA (device)'),
        ('refgpu_col_sum_f32_gpuauto', f32, f32, 2, [0], add, 0, 'GPUAuto'),
        ('refgpu_row_sum_i32_cub', i32, i32, 2, [1], add, 0, 'CUDA (device)'),
        ('refgpu_row_sum_bf16_cub', dace.bfloat16, dace.bfloat16, 2, [1], add, 0, 'CUDA (device)'),
    ):
        sdfg = _reduce_sdfg(dace, name, din, dout, nd, axes, wcr, ident, impl)
        sdfg.arrays['A'].storage = G
        sdfg.arrays['B'].storage = G
        progs.append((sdfg, {'kind': 'reduce_gpu', 'impl': impl, 'axes': axes, 'wcr': wcr, 'identity': ident}))

    N, NB = dace.symbol('N', dace.int64), dace.symbol('NB', dace.int64)

    @dace.program
    def refgpu_hist_i32(keys: dace.int32[N], bins: dace.int32[NB]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = 1

    sdfg = refgpu_hist_i32.to_sdfg()
    sdfg.name = 'refgpu_hist_i32'  # the frontend prefixes the module name unless run as __main__
    sdfg.arrays['keys'].storage = G
    sdfg.arrays['bins'].storage = G
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, dace.nodes.MapEntry):
            n.map.schedule = dace.ScheduleType.GPU_Device
    progs.append((sdfg, {'kind': 'wcr_map_gpu', 'impl': 'wcr-gpu-atomic'}))

    # The reference's answer to map -> reduce chains: MapReduceFusion (reference
    # dace/transformation/dataflow/mapreduce.py:18-146) = a WCR map, atomics on the GPU.  Same program as
    # tests/dace_e2e b200_fused_softmax_den, device-resident; comparison baseline of dace_b200/fused_expansion.py.
    import numpy as np
    from dace.transformation.dataflow import MapReduceFusion
    M = dace.symbol('M', dace.int64)

    @dace.program
    def refgpu_softmax_den_mapreducefusion(A: dace.float32[N, M], mx: dace.float32[N], s: dace.float32[N]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = np.exp(A[i, j] - mx[i])
        dace.reduce(lambda a, b: a + b, tmp, s, axis=1, identity=0)   # writes s directly (the fusion cannot handle views)

    sdfg = refgpu_softmax_den_mapreducefusion.to_sdfg(simplify=True)
    sdfg.name = 'refgpu_softmax_den_mapreducefusion'
    for a in ('A', 'mx', 's'):
        sdfg.arrays[a].storage = G
    sdfg.apply_gpu_transformations()
    applied = sdfg.apply_transformations_repeated(MapReduceFusion)
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, dace.nodes.MapEntry) and n.map.schedule == dace.ScheduleType.GPU_Device:
            n.map.gpu_block_size = [256, 1, 1]
    progs.append((sdfg, {'kind': 'map_reduce_fusion_gpu', 'impl': 'MapReduceFusion', 'applied': applied}))
    return progs


_TODO = []


def _compile_one(i):
    """Compile _TODO[i] (inherited through fork); returns (name, manifest entry, failure text or None)."""
    sdfg, meta = _TODO[i]
    name = sdfg.name
    try:
        csdfg = sdfg.compile()
    except Exception as ex:   # only tolerated for the GPU comparison programs
        if not meta['kind'].endswith('_gpu'):
            raise
        return name, None, (str(ex).splitlines()[0][:300] if str(ex) else type(ex).__name__)
    lib = 'lib%s.so' % name
    shutil.copyfile(csdfg.filename, os.path.join(REF_DIR, lib))
    # Signatures are only final after library-node expansion: read them
    # from the SDFG the reference actually generated code for.
    sdfg = csdfg.sdfg
    args = []
    for aname in sdfg.signature_arglist(with_types=False, for_call=True):
        if aname in sdfg.arrays:
            args.append({'name': aname, 'kind': 'array', 'dtype': sdfg.arrays[aname].dtype.to_string()})
        else:
            args.append({'name': aname, 'kind': 'symbol', 'ctype': sdfg.symbols[aname].ctype})
    init_names = [n.strip() for n in sdfg.init_signature(for_call=True).split(',') if n.strip()]
    entry = dict(meta)
    entry.update({'lib': lib, 'args': args, 'init_symbols': [{'name': s, 'ctype': sdfg.symbols[s].ctype} for s in init_names]})
    del csdfg
    return name, entry, None


def build_all(verbose: bool = True, only=None) -> dict:
    tmp = tempfile.mkdtemp(prefix='dace_b200_refbuild_')
    cwd = os.getcwd()
    os.chdir(tmp)  # the reference writes .dacecache/ into the cwd
    try:
        dace = import_reference(os.path.join(tmp, '.dacecache'))
        if dace is None:
            raise RuntimeError('reference checkout not available')
        os.makedirs(REF_DIR, exist_ok=True)
        manifest = {'version': VERSION, 'reference_version': dace.__version__, 'programs': {}}
        manifest['failed'] = {}
        if only:   # rebuild a few programs, keep the rest of an existing manifest
            try:
                with open(MANIFEST) as f:
                    old = json.load(f)
                manifest['programs'] = {k: v for k, v in old['programs'].items()
                                        if os.path.exists(os.path.join(REF_DIR, v['lib']))}
                manifest['failed'] = old.get('failed', {})
            except Exception:
                pass
        todo = [(g, meta) for g, meta in _programs(dace) + _gpu_programs(dace) if not only or g.name in only]
        global _TODO
        _TODO = todo
        # compiles run in forked workers (each program has its own build folder)
        import concurrent.futures
        import multiprocessing
        jobs = max(1, min(len(todo), int(os.environ.get('DACE_B200_BUILD_JOBS', os.cpu_count() or 1))))
        if jobs > 1:
            with concurrent.futures.ProcessPoolExecutor(jobs, mp_context=multiprocessing.get_context('fork')) as pool:
                results = list(pool.map(_compile_one, range(len(todo))))
        else:
            results = [_compile_one(i) for i in range(len(todo))]
        for name, entry, failure in results:
            if failure is not None:
                manifest['failed'][name] = failure
                if verbose:
                    print('FAILED', name, failure, file=sys.stderr)
                continue
            manifest['programs'][name] = entry
            if verbose:
                print('built', name, [a['name'] for a in entry['args']], file=sys.stderr)
        with open(MANIFEST, 'w') as f:
            json.dump(manifest, f, indent=1)
        return manifest
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == '__main__':
    build_all(only=set(sys.argv[1:]) or None)
