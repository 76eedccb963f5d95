"""Build DaCe programs whose Reduce nodes use the 'B200' expansion.

Runs where /root/reference exists (this container): the UNMODIFIED reference
frontend + codegen + CMake build produce ordinary DaCe shared objects whose host
code calls ``dace_b200_reduce_nd`` (cross-compiled, no GPU needed).  Only the
built .so files + a manifest land in tests/dace_e2e/_built/ (git-ignored, shipped
to the GPU box), where tests/test_gpu_dace_e2e.py loads them with ctypes exactly
like ``CompiledSDFG`` does (reference dace/codegen/compiled_sdfg.py:233-240) and
checks the results against the oracle.  This is the drop-in proof: np.sum in a
@dace.program -> Reduce library node -> 'B200' expansion -> C ABI -> sm_100a kernel.

Usage:  PYTHONPATH=/root/repo python tests/dace_e2e/build_programs.py   (also run by __graft_entry__.build())
"""
import json
import os
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '_built')
VERSION = 23


def up_to_date() -> bool:
    try:
        m = json.load(open(os.path.join(OUT, 'manifest.json')))
        return m.get('version') == VERSION and all(os.path.exists(os.path.join(OUT, e['lib']))
                                                   for e in m['programs'].values())
    except Exception:
        return False


def _to_sdfg(prog, **kwargs):
    """The frontend prefixes the SDFG name with the defining module unless that is __main__; the manifest
    keys must not depend on how this file was loaded (script vs. __graft_entry__.build())."""
    sdfg = prog.to_sdfg(**kwargs)
    sdfg.name = prog.f.__name__
    return sdfg


def _programs(dace):
    import numpy as np
    from dace_b200 import expansion as E
    from oracle.build_ref import _reduce_sdfg
    add, mx = 'lambda a, b: a + b', 'lambda a, b: max(a, b)'
    progs = []
    N, M = dace.symbol('N', dace.int64), dace.symbol('M', dace.int64)
    K = dace.symbol('K', dace.int64)

    # (1) Unmodified @dace.program; the backend is chosen only through the registry default.
    E.Reduce.default_implementation = 'B200'
    try:
        @dace.program
        def b200_c1_npsum_axis1(A: dace.float32[N, M], B: dace.float32[N]):
            B[:] = np.sum(A, axis=1)

        @dace.program
        def b200_np_max_axis0_i64(A: dace.int64[N, M], B: dace.int64[M]):
            B[:] = np.max(A, axis=0)

        @dace.program
        def b200_softmax_stats(A: dace.float32[N, M], mx_out: dace.float32[N], sm_out: dace.float32[N]):
            mx_out[:] = np.max(A, axis=1)
            sm_out[:] = np.sum(A, axis=1)

        for p in (b200_c1_npsum_axis1, b200_np_max_axis0_i64, b200_softmax_stats):
            sdfg = _to_sdfg(p)
            sdfg.apply_gpu_transformations()
            E.use_b200_reductions(sdfg)
            progs.append(sdfg)
    finally:
        E.Reduce.default_implementation = 'pure'

    # (2) Hand-built Reduce nodes (reference tests/library/reduce_test.py:14-27), host arrays.
    f32, f64, i32, bf16 = dace.float32, dace.float64, dace.int32, dace.bfloat16
    fmax = float(np.finfo(np.float32).max)
    for name, din, dout, nd, axes, wcr, ident in (
        ('b200_full_sum_f32', f32, f32, 1, None, add, 0),
        ('b200_col_sum_f32', f32, f32, 2, [0], add, 0),
        ('b200_row_sum_i32', i32, i32, 2, [1], add, 0),
        ('b200_mid2_sum_f32', f32, f32, 4, [1, 2], add, 0),
        ('b200_outer_inner_sum_f64', f64, f64, 3, [0, 2], add, 0),
        ('b200_row_max_bf16_f32', bf16, f32, 2, [1], mx, -fmax),
    ):
        sdfg = _reduce_sdfg(dace, name, din, dout, nd, axes, wcr, ident, 'B200')
        sdfg.apply_gpu_transformations()
        E.use_b200_reductions(sdfg)
        progs.append(sdfg)

    # (3) Sub-range / offset memlets (reference tests/reduce_offsets_test.py:26-66).
    sdfg = dace.SDFG('b200_offsets')
    sdfg.add_array('A', [50, 50], f64)
    sdfg.add_array('B', [25], f64)
    st = sdfg.add_state()
    r = st.add_reduce(add, [0], 0)
    r.implementation = 'B200'
    st.add_nedge(st.add_read('A'), r, dace.Memlet('A[25:50, 13:20]'))
    st.add_nedge(r, st.add_write('B'), dace.Memlet('B[4:11]'))
    sdfg.apply_gpu_transformations()
    E.use_b200_reductions(sdfg)
    progs.append(sdfg)

    # (3b) Interiors of larger arrays (stencil domains): every element of A[1:N-1, 1:M-1] and per-plane sums of
    #      V[:, 1:H-1, 1:W-1] - all-reduced / innermost-reduced windows with a pitch on one and two levels, which take
    #      the inner-first plan and the row kernels on two levels of rows.
    @dace.program
    def b200_interior_sum(A: dace.float64[N, M], s: dace.float64[1]):
        s[0] = np.sum(A[1:N - 1, 1:M - 1])

    @dace.program
    def b200_interior_plane_sums(V: dace.float32[K, N, M], s: dace.float32[K], t: dace.float32[K, N - 2]):
        s[:] = np.sum(V[:, 1:N - 1, 1:M - 1], axis=(1, 2))
        t[:] = np.max(V[:, 1:N - 1, 1:M - 1], axis=2)

    for p in (b200_interior_sum, b200_interior_plane_sums):
        sdfg = _to_sdfg(p, simplify=True)
        sdfg.apply_gpu_transformations()
        E.use_b200_reductions(sdfg)
        progs.append(sdfg)

    # (4) Device-resident arrays (GPU_Global): no copies in the generated program.
    sdfg = dace.SDFG('b200_full_sum_f32_resident')
    D0 = dace.symbol('D0', dace.int64)
    sdfg.add_array('A', [D0], f32, storage=dace.StorageType.GPU_Global)
    sdfg.add_array('B', [1], f32, storage=dace.StorageType.GPU_Global)
    st = sdfg.add_state()
    r = st.add_reduce(add, None, 0)
    r.implementation = 'B200'
    st.add_nedge(st.add_read('A'), r, dace.Memlet.from_array('A', sdfg.arrays['A']))
    st.add_nedge(r, st.add_write('B'), dace.Memlet.from_array('B', sdfg.arrays['B']))
    progs.append(sdfg)

    # (4b) identity=None folds into the output already on the device.  Device-resident on purpose: with host
    #      arrays the reference's GPU transformation does not copy a write-only B in, so the fold would start
    #      from whatever the fresh cudaMalloc holds.
    sdfg = dace.SDFG('b200_noident_col_max_f64')
    D1 = dace.symbol('D1', dace.int64)
    sdfg.add_array('A', [D0, D1], f64, storage=dace.StorageType.GPU_Global)
    sdfg.add_array('B', [D1], f64, storage=dace.StorageType.GPU_Global)
    st = sdfg.add_state()
    r = st.add_reduce(mx, [0], None)
    r.implementation = 'B200'
    st.add_nedge(st.add_read('A'), r, dace.Memlet.from_array('A', sdfg.arrays['A']))
    st.add_nedge(r, st.add_write('B'), dace.Memlet.from_array('B', sdfg.arrays['B']))
    progs.append(sdfg)

    # (4c) Two independent large reductions in ONE state: DaCe's default stream assignment
    #      (compiler.cuda.max_concurrent_streams = 0) puts them on different streams, so they may overlap on
    #      the GPU - each expanded node owns its workspace (tickets + partials), nothing is shared.
    sdfg = dace.SDFG('b200_two_reductions_one_state')
    sdfg.add_array('A', [D0], f32, storage=dace.StorageType.GPU_Global)
    sdfg.add_array('C', [D1, 4096], f32, storage=dace.StorageType.GPU_Global)
    sdfg.add_array('s', [1], f32, storage=dace.StorageType.GPU_Global)
    sdfg.add_array('cs', [4096], f32, storage=dace.StorageType.GPU_Global)
    st = sdfg.add_state()
    r1 = st.add_reduce(add, None, 0)
    r2 = st.add_reduce(add, [0], 0)
    r1.implementation = r2.implementation = 'B200'
    st.add_nedge(st.add_read('A'), r1, dace.Memlet.from_array('A', sdfg.arrays['A']))
    st.add_nedge(r1, st.add_write('s'), dace.Memlet.from_array('s', sdfg.arrays['s']))
    st.add_nedge(st.add_read('C'), r2, dace.Memlet.from_array('C', sdfg.arrays['C']))
    st.add_nedge(r2, st.add_write('cs'), dace.Memlet.from_array('cs', sdfg.arrays['cs']))
    progs.append(sdfg)

    # (4d) Host arrays with persistent device copies: the reference's make_transients_persistent
    #      (reference dace/transformation/auto/auto_optimize.py:460-540) moves the cudaMalloc / cudaFree of gpu_A /
    #      gpu_B from every call into __dace_init / __dace_exit; together with the per-node workspace nothing is
    #      allocated on the call path.  This is the program bench.py times end to end.
    from dace.transformation.auto.auto_optimize import make_transients_persistent
    sdfg = _reduce_sdfg(dace, 'b200_full_sum_f32_persistent', f32, f32, 1, None, add, 0, 'B200')
    sdfg.apply_gpu_transformations()
    E.use_b200_reductions(sdfg)
    made = make_transients_persistent(sdfg, dace.DeviceType.GPU)
    if not any(made.values()):
        raise RuntimeError('b200_full_sum_f32_persistent: no transient became persistent')
    progs.append(sdfg)

    # (5) WCR maps (scatter-add / histograms): unmodified @dace.programs; the map is rewritten into a
    #     ScatterReduce library node expanded with 'B200' (dace_b200/scatter_node.py).
    from dace_b200 import scatter_node as SN

    @dace.program
    def b200_hist_augassign(keys: dace.int32[N], bins: dace.int32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    @dace.program
    def b200_hist_explicit_wcr(keys: dace.int32[N], bins: dace.uint32[M]):   # reference samples/explicit/histogram.py:13-26
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                out >> bins(1, lambda x, y: x + y)[:]
                out[k] = 1

    @dace.program
    def b200_hist_weighted_f32(keys: dace.int32[N], w: dace.float32[N], bins: dace.float32[M]):
        for i in dace.map[0:N]:
            bins[keys[i]] += w[i]

    @dace.program
    def b200_scatter_max_f64(keys: dace.int64[N], w: dace.float64[N], bins: dace.float64[M]):
        for i in dace.map[0:N]:
            with dace.tasklet:
                k << keys[i]
                v << w[i]
                out >> bins(1, lambda x, y: max(x, y))[:]
                out[k] = v

    for j, p in enumerate((b200_hist_augassign, b200_hist_explicit_wcr, b200_hist_weighted_f32, b200_scatter_max_f64)):
        sdfg = _to_sdfg(p, simplify=True)
        if j % 2 == 0:   # both orders of the two rewrites are supported
            sdfg.apply_gpu_transformations()
            applied = SN.use_b200_scatter(sdfg)
        else:
            applied = SN.use_b200_scatter(sdfg)
            sdfg.apply_gpu_transformations()
        if applied != 1:
            raise RuntimeError('%s: WCRMapToScatterReduce applied %d times' % (sdfg.name, applied))
        progs.append(sdfg)

    # (6) Whole-program route: the reference's auto_optimize(device=GPU) with the backend selected
    #     automatically (dace_b200/auto.py): two Reduce nodes and one WCR map in one program.
    from dace_b200 import auto as AUTO

    @dace.program
    def b200_auto_optimized(A: dace.float32[N, M], keys: dace.int32[N], bins: dace.int32[M], rs: dace.float32[N],
                            mx: dace.float32[1]):
        rs[:] = np.sum(A, axis=1)
        mx[0] = np.max(A)
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    sdfg = AUTO.auto_optimize(_to_sdfg(b200_auto_optimized, simplify=True), dace.DeviceType.GPU)
    progs.append(sdfg)

    # (7) In-kernel reductions: 'B200 (block)' on the SDFG of reference tests/blockreduce_cudatest.py:11-43
    #     (register items per thread -> one value per block), and 'B200 (block allreduce)' reached through
    #     the reference's ReduceExpansion transformation (reference
    #     tests/transformations/subgraph_fusion/block_allreduce_cudatest.py:17-45).
    from dace_b200 import block_expansion  # noqa: F401  (registers the expansions)
    progs.append(_blockreduce_sdfg(dace, 'b200_blockreduce_f32', add, f32, nblocks=2, nthreads=32, items=2))
    progs.append(_blockreduce_sdfg(dace, 'b200_blockreduce_i32_max', mx, i32, nblocks=5, nthreads=160, items=3))
    progs.append(_blockreduce_sdfg(dace, 'b200_blockreduce_f64_min', 'lambda a, b: min(a, b)', f64, nblocks=3,
                                   nthreads=1024, items=1))
    progs.append(_blockreduce_sdfg(dace, 'b200_blockreduce_custom_i64', 'lambda a, b: a + b + a * b', dace.int64,
                                   nblocks=4, nthreads=72, items=2))

    from dace.transformation.dataflow import ReduceExpansion

    @dace.program
    def b200_blockallreduce(A: dace.float32[N, 96], B: dace.float32[N]):
        B[:] = dace.reduce(lambda a, b: max(a, b), A, axis=1, identity=0)

    sdfg = _to_sdfg(b200_blockallreduce)
    sdfg.apply_gpu_transformations()
    found = [(n, st) for st in sdfg.nodes() for n in st.nodes() if isinstance(n, E.Reduce)]
    assert len(found) == 1
    node, st = found[0]
    node.implementation = 'B200 (block allreduce)'   # ReduceExpansion hands it on to the inner node
    xf = ReduceExpansion()
    xf.setup_match(sdfg, 0, sdfg.node_id(st), {ReduceExpansion.reduce: st.node_id(node)}, 0)
    xf.apply(st, sdfg)
    progs.append(sdfg)

    # (8) Multi-GPU: every rank runs this program on its own row / key shard; the sharding transformation
    #     (dace_b200/comm_node.py) inserts AllReduceCombine after the reductions over the sharded axis
    #     (column sum, full max with a non-neutral identity, histogram) and leaves the row sum local.
    from dace_b200 import comm_node as CN

    @dace.program
    def b200_sharded(A: dace.float32[N, M], keys: dace.int32[N], cs: dace.float32[M], tot: dace.float32[1],
                     rs: dace.float32[N], bins: dace.int32[M]):
        cs[:] = np.sum(A, axis=0)
        tot[0] = dace.reduce(lambda a, b: max(a, b), A, identity=0.25)
        rs[:] = np.sum(A, axis=1)
        for i in dace.map[0:N]:
            bins[keys[i]] += 1

    sdfg = _to_sdfg(b200_sharded, simplify=True)
    sdfg.apply_gpu_transformations()
    AUTO.apply_b200(sdfg)
    counts = CN.shard_reductions(sdfg, {'A': 0, 'keys': 0})
    if counts != {'reduce': 2, 'scatter': 1, 'local': 1}:
        raise RuntimeError('b200_sharded: unexpected sharding result %s' % counts)
    progs.append(sdfg)

    # (8b) Sharding + fusion in the right order: select (no fusing) -> shard -> fuse.  The local partial of the column
    #      sum of squares is computed by the generated fused kernel, then combined across the ranks.
    @dace.program
    def b200_sharded_fused(A: dace.float32[N, M], sq: dace.float32[M]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = A[i, j] * A[i, j]
        sq[:] = np.sum(tmp, axis=0)

    sdfg = _to_sdfg(b200_sharded_fused, simplify=True)
    sdfg.apply_gpu_transformations()
    AUTO.apply_b200(sdfg, fuse=False)
    if CN.shard_reductions(sdfg, {'A': 0})['reduce'] != 1 or AUTO.apply_b200(sdfg)['fused'] != 1:
        raise RuntimeError('b200_sharded_fused: expected one combined and fused reduction')
    progs.append(sdfg)

    # (9) np.argmax / np.argmin: the frontend's struct-WCR map -> ArgReduce library node -> dace_b200_reduce_arg
    #     (dace_b200/arg_node.py).  The flat form (axis=None) does not survive the reference's own
    #     apply_gpu_transformations (host interstate edge reading device data), so it is not built here.
    from dace_b200 import arg_node as AN
    K = dace.symbol('K', dace.int64)

    @dace.program
    def b200_argmax_f32_axis1(A: dace.float32[N, M], out: dace.int32[N]):
        out[:] = np.argmax(A, axis=1)

    @dace.program
    def b200_argmin_f64_axis0(A: dace.float64[N, M], out: dace.int32[M]):
        out[:] = np.argmin(A, axis=0)

    @dace.program
    def b200_argmin_i32_mid(A: dace.int32[N, M, K], out: dace.int32[N, K]):
        out[:] = np.argmin(A, axis=1)

    for j, p in enumerate((b200_argmax_f32_axis1, b200_argmin_f64_axis0, b200_argmin_i32_mid)):
        sdfg = _to_sdfg(p, simplify=True)
        sdfg.apply_gpu_transformations()
        applied = AN.use_b200_argreduce(sdfg) if j < 2 else AUTO.apply_b200(sdfg)['arg']
        if applied != 1:
            raise RuntimeError('%s: ArgReduceMapToLibrary applied %d times' % (sdfg.name, applied))
        progs.append(sdfg)

    # (10) Row max + row sum of one array (BASELINE C3): the two Reduce nodes are fused into one RowsMaxSum
    #      node = one pass over the input (dace_b200/fuse_node.py).  f32 from an unmodified @dace.program;
    #      bf16 -> fp32 outputs needs hand-built nodes (the frontend keeps the input dtype), device-resident.
    from dace_b200 import fuse_node as FN

    @dace.program
    def b200_softmax_stats_fused(A: dace.float32[N, M], mx_out: dace.float32[N], sm_out: dace.float32[N]):
        mx_out[:] = np.max(A, axis=1)
        sm_out[:] = np.sum(A, axis=1)

    sdfg = _to_sdfg(b200_softmax_stats_fused, simplify=True)
    sdfg.apply_gpu_transformations()
    AUTO.apply_b200(sdfg, fuse=False)
    if FN.fuse_row_max_sum(sdfg) != 1:
        raise RuntimeError('b200_softmax_stats_fused: the Reduce pair was not fused')
    progs.append(sdfg)

    sdfg = dace.SDFG('b200_c3_fused_bf16')
    G = dace.StorageType.GPU_Global
    sdfg.add_array('A', [N, M], bf16, storage=G)
    sdfg.add_array('mx', [N], f32, storage=G)
    sdfg.add_array('sm', [N], f32, storage=G)
    st = sdfg.add_state()
    a = st.add_read('A')
    r1 = st.add_reduce(mx, [1], -fmax)
    r2 = st.add_reduce(add, [1], 0)
    st.add_nedge(a, r1, dace.Memlet.from_array('A', sdfg.arrays['A']))
    st.add_nedge(a, r2, dace.Memlet.from_array('A', sdfg.arrays['A']))
    st.add_nedge(r1, st.add_write('mx'), dace.Memlet.from_array('mx', sdfg.arrays['mx']))
    st.add_nedge(r2, st.add_write('sm'), dace.Memlet.from_array('sm', sdfg.arrays['sm']))
    if FN.fuse_row_max_sum(sdfg) != 1:
        raise RuntimeError('b200_c3_fused_bf16: the Reduce pair was not fused')
    progs.append(sdfg)

    # (11) The other frontend producers of Reduce nodes (reference replacements/reduction.py:102-146):
    #      np.any / np.all (logical WCRs; np.all carries identity 0 in the reference, so it is all-false -
    #      a quirk parity has to keep), np.mean (sum + elementwise divide), np.min with `initial`.
    E.Reduce.default_implementation = 'B200'
    try:
        @dace.program
        def b200_np_any_all(A: dace.int32[N, M], any_out: dace.int32[N], all_out: dace.int32[N]):
            any_out[:] = np.any(A, axis=1)
            all_out[:] = np.all(A, axis=1)

        @dace.program
        def b200_np_mean_axis0(A: dace.float64[N, M], B: dace.float64[M]):
            B[:] = np.mean(A, axis=0)

        @dace.program
        def b200_np_min_initial(A: dace.int32[N, M], B: dace.int32[M]):
            B[:] = np.min(A, axis=0, initial=7)

        for p in (b200_np_any_all, b200_np_mean_axis0, b200_np_min_initial):
            sdfg = _to_sdfg(p)
            sdfg.apply_gpu_transformations()
            E.use_b200_reductions(sdfg)
            progs.append(sdfg)
    finally:
        E.Reduce.default_implementation = 'pure'

    # (12) Map -> transient -> Reduce fused into one generated kernel (dace_b200/fused_expansion.py): the map body
    #      becomes a load functor for dace_b200::fused_rows.  `_unfused` twins keep the transient (map kernel by the
    #      reference's code generator + 'B200' Reduce) for the A/B timing in tools/fused_probe.py.
    from dace_b200 import fused_expansion as FE

    @dace.program
    def b200_fused_softmax_den(A: dace.float32[N, M], mx: dace.float32[N], s: dace.float32[N]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = np.exp(A[i, j] - mx[i])
        s[:] = np.sum(tmp, axis=1)

    @dace.program
    def b200_unfused_softmax_den(A: dace.float32[N, M], mx: dace.float32[N], s: dace.float32[N]):
        tmp = np.ndarray([N, M], dace.float32)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = np.exp(A[i, j] - mx[i])
        s[:] = np.sum(tmp, axis=1)

    @dace.program
    def b200_fused_sumsq(A: dace.float64[N, M], s: dace.float64[1]):
        s[0] = np.sum(A * A)

    @dace.program
    def b200_fused_absdiff_max(A: dace.int32[N, M], B: dace.int32[N, M], out: dace.int32[N]):
        tmp = np.ndarray([N, M], dace.int32)
        for i, j in dace.map[0:N, 0:M]:
            with dace.tasklet:
                a << A[i, j]
                b << B[i, j]
                o >> tmp[i, j]
                o = abs(a - b)
        out[:] = np.max(tmp, axis=1)

    @dace.program
    def b200_fused_weighted_3d(A: dace.float64[N, M, K], w: dace.float64[K], out: dace.float64[N, M]):
        tmp = np.ndarray([N, M, K], dace.float64)
        for i, j, k in dace.map[0:N, 0:M, 0:K]:
            tmp[i, j, k] = A[i, j, k] * w[k] + i
        out[:] = np.sum(tmp, axis=2)

    @dace.program
    def b200_fused_colstats(A: dace.float32[N, M, K], mean: dace.float32[K], out: dace.float32[K]):
        tmp = np.ndarray([N, M, K], dace.float32)
        for i, j, k in dace.map[0:N, 0:M, 0:K]:
            tmp[i, j, k] = (A[i, j, k] - mean[k]) * (A[i, j, k] - mean[k])
        out[:] = np.sum(tmp, axis=(0, 1))

    for p in (b200_fused_softmax_den, b200_unfused_softmax_den, b200_fused_sumsq, b200_fused_absdiff_max,
              b200_fused_weighted_3d, b200_fused_colstats):
        sdfg = _to_sdfg(p, simplify=True)
        sdfg.apply_gpu_transformations()
        # device-resident arguments: the probe times kernels, not PCIe copies (tests pass device pointers)
        if 'softmax_den' in sdfg.name:
            sdfg = _to_sdfg(p, simplify=True)
            for a in ('A', 'mx', 's'):
                sdfg.arrays[a].storage = dace.StorageType.GPU_Global
            sdfg.apply_gpu_transformations()
        fused = FE.fuse_maps_into_reductions(sdfg) if '_fused_' in sdfg.name else 0
        if fused != (1 if '_fused_' in sdfg.name else 0):
            raise RuntimeError('%s: FuseMapIntoReduce applied %d times' % (sdfg.name, fused))
        AUTO.apply_b200(sdfg, fuse=False)
        for n, _ in sdfg.all_nodes_recursive():       # a sensible block size for the unfused twin's map kernel
            if isinstance(n, dace.nodes.MapEntry) and n.map.schedule == dace.ScheduleType.GPU_Device:
                n.map.gpu_block_size = [256, 1, 1]
        progs.append(sdfg)

    # (13) Reductions written as WCR maps (no indirection): `s[i] += f(A[i, j])` through the frontend's AugAssignToWCR,
    #      an explicit WCR tasklet output folding everything into one element, and a column fold.  The reference emits
    #      one atomic per element for these (reference codegen/targets/cpu.py:965-1013); WCRMapToFusedReduce
    #      (dace_b200/fused_expansion.py) turns each into one tree-reduction kernel that folds into the output.
    @dace.program
    def b200_wcr_rows(A: dace.float32[N, M], mx: dace.float32[N], s: dace.float32[N]):
        for i, j in dace.map[0:N, 0:M]:
            s[i] += np.exp(A[i, j] - mx[i])

    @dace.program
    def b200_wcr_all_f64(A: dace.float64[N, M], s: dace.float64[1]):
        for i, j in dace.map[0:N, 0:M]:
            with dace.tasklet:
                xin << A[i, j]                      # (names that do not shadow locals of this function)
                res >> s(1, lambda x, y: x + y)[0]
                res = xin * xin

    @dace.program
    def b200_wcr_cols_max_i32(A: dace.int32[N, M], B: dace.int32[N, M], out: dace.int32[M]):
        for i, j in dace.map[0:N, 0:M]:
            with dace.tasklet:
                xin << A[i, j]
                yin << B[i, j]
                res >> out(1, lambda x, y: max(x, y))[j]
                res = xin - yin

    @dace.program
    def b200_fused_f16_rows(A: dace.float16[N, M], s: dace.float16[N]):
        tmp = np.ndarray([N, M], dace.float16)
        for i, j in dace.map[0:N, 0:M]:
            tmp[i, j] = A[i, j] * A[i, j]
        s[:] = np.sum(tmp, axis=1)

    sdfg = _to_sdfg(b200_fused_f16_rows, simplify=True)
    sdfg.apply_gpu_transformations()
    if FE.fuse_maps_into_reductions(sdfg) != 1:
        raise RuntimeError('b200_fused_f16_rows: not fused')
    progs.append(sdfg)

    for p in (b200_wcr_rows, b200_wcr_all_f64, b200_wcr_cols_max_i32):
        sdfg = _to_sdfg(p, simplify=True)
        sdfg.apply_gpu_transformations()
        if FE.lower_wcr_maps(sdfg) != 1:
            raise RuntimeError('%s: WCRMapToFusedReduce did not apply' % sdfg.name)
        progs.append(sdfg)
    # (12b) np.argmax(A) / np.argmin(A) over ALL elements: the frontend reads the (value, index) pair on an interstate
    #       edge, which the reference's GPU transformation rejects; arg_node.lower_flat_arg_results turns the read into
    #       a memlet first, then the 1-D arg-reduce map becomes an ArgReduce node like the axis forms.
    from dace_b200 import arg_node as AN

    @dace.program
    def b200_flat_argmax_f32(A: dace.float32[N, M], out: dace.int64[1]):
        out[0] = np.argmax(A)

    @dace.program
    def b200_flat_argmin_i32(A: dace.int32[N, M, K], out: dace.int64[1]):
        out[0] = np.argmin(A)

    for p in (b200_flat_argmax_f32, b200_flat_argmin_i32):
        sdfg = _to_sdfg(p, simplify=True)
        if AN.lower_flat_arg_results(sdfg) != 1:
            raise RuntimeError('%s: flat arg result not lowered' % sdfg.name)
        sdfg.apply_gpu_transformations()
        if AN.use_b200_argreduce(sdfg) != 1:
            raise RuntimeError('%s: ArgReduceMapToLibrary did not apply' % sdfg.name)
        progs.append(sdfg)

    # (13b) ONE map with several WCR outputs computed by different tasklets (reference tests/multiple_cr_test.py: a
    #       sum and a product from the same map): split into one map per output (dace_b200/split_wcr.py), each
    #       piece lowered to a fused tree reduction - through apply_b200, the entry point auto_optimize uses.
    from dace_b200 import auto as AUTO
    sdfg = dace.SDFG('b200_multiple_cr')
    sdfg.add_array('A', [N], dace.int64)
    sdfg.add_array('s', [1], dace.int64)
    sdfg.add_array('p', [1], dace.int64)
    st = sdfg.add_state()
    me, mx = st.add_map('mymap', dict(i='0:N'))
    rd = st.add_read('A')
    for tname, arr, wcr in (('task1', 's', 'lambda a, b: a + b'), ('task2', 'p', 'lambda a, b: a * b')):
        t = st.add_tasklet(tname, {'a'}, {'b'}, 'b = a')
        st.add_memlet_path(rd, me, t, dst_conn='a', memlet=dace.Memlet('A[i]'))
        st.add_memlet_path(t, mx, st.add_write(arr), src_conn='b', memlet=dace.Memlet('%s[0]' % arr, wcr=wcr))
    sdfg.apply_gpu_transformations()
    got = AUTO.apply_b200(sdfg)
    if got['split'] != 1 or got['wcr_fused'] != 2:
        raise RuntimeError('b200_multiple_cr: %s' % got)
    progs.append(sdfg)

    @dace.program
    def b200_two_wcr_rows(A: dace.float32[N, M], s: dace.float32[N], big: dace.float32[N]):
        for i, j in dace.map[0:N, 0:M]:
            with dace.tasklet:
                xin << A[i, j]
                res >> s(1, lambda x, y: x + y)[i]
                res = xin
            with dace.tasklet:
                xin2 << A[i, j]
                res2 >> big(1, lambda x, y: max(x, y))[i]
                res2 = xin2 * xin2

    sdfg = _to_sdfg(b200_two_wcr_rows, simplify=True)
    sdfg.apply_gpu_transformations()
    got = AUTO.apply_b200(sdfg)
    if got['split'] != 1 or got['wcr_fused'] != 2:
        raise RuntimeError('b200_two_wcr_rows: %s' % got)
    progs.append(sdfg)

    # (14) Custom WCR lambdas on the device-level Reduce node (ReductionType.Custom): the lambda becomes a generated
    #      functor instantiated with the header kernels (dace_b200/expansion.py:_expand_custom), as the reference's
    #      CUB expansion does (reference reduce.py:453-468).  Same SDFGs as the golden vectors the reference's `pure`
    #      expansion produced (oracle/make_golden.py CUSTOM_CASES), device-resident arrays.
    from oracle.make_golden import CUSTOM_CASES
    for tag, dtype, shape, axes, wcr, identity in CUSTOM_CASES:
        dt = {'int64': dace.int64, 'float32': f32}[dtype]
        sdfg = _reduce_sdfg(dace, 'b200_custom_' + tag, dt, dt, len(shape), axes, wcr, identity, 'B200')
        for a in ('A', 'B'):
            sdfg.arrays[a].storage = dace.StorageType.GPU_Global
        progs.append(sdfg)

    # (15) Privatised shared-memory WCR tiles: the reference's own AccumulateTransient between a
    #      thread-block map and its kernel map, tile in shared memory - generated call sites
    #      dace::ResetShared<..>::Reset / dace::SharedToGlobal1D<..>::Accum<Sum> are served by
    #      include/dace_b200_shared_accum.cuh (dace_b200/shared_tile.py).
    from dace_b200 import shared_tile as ST

    @dace.program
    def b200_shared_tile_hist(keys: dace.int32[N], bins: dace.int32[256]):
        for b in dace.map[0:N // 1024]:
            for t in dace.map[0:256]:
                for j in range(4):
                    with dace.tasklet:
                        k << keys[b * 1024 + j * 256 + t]
                        out >> bins(1, lambda x, y: x + y)[:]
                        out[k] = 1

    @dace.program
    def b200_shared_tile_weighted_f32(keys: dace.int32[N], w: dace.float32[N], bins: dace.float32[512]):
        for b in dace.map[0:N // 512]:
            for t in dace.map[0:128]:
                for j in range(4):
                    with dace.tasklet:
                        k << keys[b * 512 + j * 128 + t]
                        v << w[b * 512 + j * 128 + t]
                        out >> bins(1, lambda x, y: x + y)[:]
                        out[k] = v

    for p in (b200_shared_tile_hist, b200_shared_tile_weighted_f32):
        sdfg = _to_sdfg(p, simplify=True)
        sdfg.apply_gpu_transformations()
        st = next(s_ for s_ in sdfg.states() if sum(isinstance(n, dace.nodes.MapExit) for n in s_.nodes()) >= 2)
        exits = [n for n in st.nodes() if isinstance(n, dace.nodes.MapExit)]
        inner = next(e for e in exits if e.map.params == ['t'])
        outer = next(e for e in exits if e.map.params == ['b'])
        st.entry_node(inner).map.schedule = dace.ScheduleType.GPU_ThreadBlock
        ST.privatise_wcr_in_shared(sdfg, st, inner, outer)
        progs.append(sdfg)

    return progs


def _blockreduce_sdfg(dace, name, wcr, dt, nblocks, nthreads, items):
    """nblocks x nthreads x items values -> one value per block through a Reduce node inside the kernel."""
    from dace.memlet import Memlet
    from dace.transformation.interstate import GPUTransformSDFG
    per_block = nthreads * items
    sdfg = dace.SDFG(name)
    sdfg.add_array('A', (nblocks * per_block, ), dt)
    sdfg.add_array('B', (nblocks, ), dt)
    sdfg.add_transient('tA', (items, ), dt)
    sdfg.add_transient('tB', (1, ), dt)
    st = sdfg.add_state('a')
    A, B, tA, tB = st.add_access('A'), st.add_access('B'), st.add_access('tA'), st.add_access('tB')
    me, mx = st.add_map('blocks', dict(bi='0:%d' % nblocks))
    mei, mxi = st.add_map('threads', dict(i='0:%d' % nthreads))
    red = st.add_reduce(wcr, None, 0)
    red.implementation = 'B200 (block)'
    wt = st.add_tasklet('writeout', {'inp'}, {'out'}, 'if i == 0: out = inp')
    st.add_edge(A, None, me, None, Memlet.simple(A, '0:%d' % (nblocks * per_block)))
    st.add_edge(me, None, mei, None, Memlet.simple(A, '(%d*bi):(%d*bi+%d)' % (per_block, per_block, per_block)))
    st.add_edge(mei, None, tA, None,
                Memlet.simple('A', '(%d*bi+%d*i):(%d*bi+%d*i+%d)' % (per_block, items, per_block, items, items)))
    st.add_edge(tA, None, red, None, Memlet.simple(tA, '0:%d' % items))
    st.add_edge(red, None, tB, None, Memlet.simple(tB, '0'))
    st.add_edge(tB, None, wt, 'inp', Memlet.simple(tB, '0'))
    st.add_edge(wt, 'out', mxi, None, Memlet.simple('B', 'bi', num_accesses=-1))
    st.add_edge(mxi, None, mx, None, Memlet.simple(B, 'bi'))
    st.add_edge(mx, None, B, None, Memlet.simple(B, '0:%d' % nblocks))
    sdfg.fill_scope_connectors()
    sdfg.apply_transformations(GPUTransformSDFG, options={'sequential_innermaps': False})
    return sdfg


_TODO = []


def _compile_one(i):
    """Compile _TODO[i] (inherited through fork) and describe it for the manifest."""
    sdfg = _TODO[i]
    name = sdfg.name
    csdfg = sdfg.compile()
    lib = 'lib%s.so' % name
    shutil.copyfile(csdfg.filename, os.path.join(OUT, lib))
    g = csdfg.sdfg
    args = []
    for aname in g.signature_arglist(with_types=False, for_call=True):
        if aname in g.arrays:
            args.append({'name': aname, 'kind': 'array', 'dtype': g.arrays[aname].dtype.to_string(),
                         'storage': str(g.arrays[aname].storage)})
        else:
            args.append({'name': aname, 'kind': 'symbol', 'ctype': g.symbols[aname].ctype})
    init_names = [n.strip() for n in g.init_signature(for_call=True).split(',') if n.strip()]
    srcdir = os.path.join(os.path.dirname(os.path.dirname(csdfg.filename)), 'src')
    src = open(os.path.join(srcdir, 'cpu', name + '.cpp')).read()
    cu = os.path.join(srcdir, 'cuda', name + '_cuda.cu')
    cu = open(cu).read() if os.path.exists(cu) else ''
    entry = {
        'lib': lib, 'args': args,
        'init_symbols': [{'name': s, 'ctype': g.symbols[s].ctype} for s in init_names],
        'b200_calls': src.count('dace_b200_reduce_nd(') + src.count('dace_b200_wcr_scatter(') +
        src.count('dace_b200_reduce_arg(') + src.count('dace_b200_reduce_rows_max_sum(') + src.count('__b200_fused_') // 2 + src.count('DACE_GPU_CHECK(__b200_custom_') + cu.count('dace_b200 shared tiles') +
        cu.count('dace_b200::block_reduce<') + cu.count('dace_b200::block_allreduce<'),
        'comm_calls': src.count('dace_b200_comm_allreduce('),
    }
    del csdfg
    return name, entry


def build_all(verbose=True, only=None):
    from oracle.refenv import import_reference
    tmp = tempfile.mkdtemp(prefix='dace_b200_e2e_')
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        dace = import_reference(os.path.join(tmp, '.dacecache'))
        if dace is None:
            raise RuntimeError('reference checkout not available')
        os.makedirs(OUT, exist_ok=True)
        manifest = {'version': VERSION, 'reference_version': dace.__version__, 'programs': {}}
        if only:
            try:
                old = json.load(open(os.path.join(OUT, 'manifest.json')))
                manifest['programs'] = {k: v for k, v in old['programs'].items()
                                        if os.path.exists(os.path.join(OUT, v['lib']))}
            except Exception:
                pass
        todo = [g for g in _programs(dace) if not only or g.name in only]
        global _TODO
        _TODO = todo
        # The SDFGs are generated here; the compiles (CMake + nvcc + g++, ~10 s each) run in forked workers,
        # every program in its own build folder.
        import concurrent.futures
        import multiprocessing
        jobs = max(1, min(len(todo), int(os.environ.get('DACE_B200_BUILD_JOBS', os.cpu_count() or 1))))
        if jobs > 1 and sys.modules.get(__name__) is not None and getattr(sys.modules[__name__], '_compile_one', None) is _compile_one:
            with concurrent.futures.ProcessPoolExecutor(jobs, mp_context=multiprocessing.get_context('fork')) as pool:
                results = list(pool.map(_compile_one, range(len(todo))))
        else:   # loaded under a name the workers could not import: compile one after the other
            results = [_compile_one(i) for i in range(len(todo))]
        for name, entry in results:
            manifest['programs'][name] = entry
            if verbose:
                print('built', name, [a['name'] for a in entry['args']], 'b200 calls:', entry['b200_calls'], file=sys.stderr)
        json.dump(manifest, open(os.path.join(OUT, 'manifest.json'), 'w'), indent=1)
        return manifest
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == '__main__':
    build_all(only=set(sys.argv[1:]) or None)
