"""DaCe plug-in: the ``'B200'`` expansion of ``dace.libraries.standard.Reduce``.

Importing this module (DaCe must be importable) registers

* ``B200Reduce`` - a ``@dace.library.environment`` that adds ``include/`` to the
  include path, links ``libdace_b200_reduce.so`` (with an rpath) and includes the C
  header into the generated frame code (reference environment mechanics:
  dace/library.py:149-164, dace/codegen/compiler.py:782-860,
  dace/codegen/targets/framecode.py:322-327,386-391);
* ``ExpandReduceB200`` - ``Reduce.implementations['B200']`` via
  ``dace.library.register_expansion`` (reference dace/library.py:137-145).  Like
  the reference's ``CUDA (device)`` expansion it returns a host C++ tasklet with
  pointer connectors ``_in`` / ``_out`` (reference
  dace/libraries/standard/nodes/reduce.py:599-620); unlike it, the tasklet does not
  call a per-node generated CUB wrapper but ONE C-ABI entry point,
  ``dace_b200_reduce_nd``, passing the memlet shape/strides as run-time arrays,
  so every axis combination, offset, stride and step is served by the
  hand-written sm_100a kernels.

Select it with ``node.implementation = 'B200'``, or for whole programs with
``Reduce.default_implementation = 'B200'`` (the pattern of reference
tests/numpy/reductions_test.py:263-264), or call :func:`use_b200_reductions`.
Custom WCR lambdas become a generated functor instantiated with the header kernels of
``include/dace_b200_fused_reduce.cuh`` inside the program's .cu (the reference's CUB expansion does
the same with cub::DeviceReduce).  There is no CPU fallback: data that is not in ``GPU_Global`` /
``CPU_Pinned`` storage and argmin/argmax raise at expansion time.
"""
import copy
import itertools
import os

import dace
import dace.library
from dace import dtypes, nodes
from dace.frontend.operations import detect_reduction_type
from dace.libraries.standard.nodes.reduce import Reduce
from dace.symbolic import symstr
from dace.transformation.transformation import ExpandTransformation

from . import _lib
from .redtypes import ReductionType, lower_custom

_DTYPE_CODES = {
    dace.float32: 0, dace.float64: 1, dace.int32: 2, dace.int64: 3, dace.uint32: 4, dace.uint64: 5,
    dace.float16: 7, dace.int8: 8, dace.uint8: 9, dace.int16: 10, dace.uint16: 11, dace.bool_: 9,
}
if hasattr(dace, 'bfloat16'):
    _DTYPE_CODES[dace.bfloat16] = 6

_SUPPORTED = {ReductionType.Min, ReductionType.Max, ReductionType.Sum, ReductionType.Product,
              ReductionType.Logical_And, ReductionType.Bitwise_And, ReductionType.Logical_Or,
              ReductionType.Bitwise_Or, ReductionType.Logical_Xor, ReductionType.Bitwise_Xor}


@dace.library.environment
class B200Reduce:
    """Links the prebuilt sm_100a kernel library into the generated program."""
    cmake_minimum_version = None
    cmake_packages = ["CUDA"]
    cmake_variables = {}
    cmake_includes = [_lib.INCLUDE_DIR]
    cmake_libraries = [_lib.LIB_PATH]
    cmake_compile_flags = []
    cmake_link_flags = ['-Wl,-rpath,' + os.path.dirname(_lib.LIB_PATH)]
    cmake_files = []

    headers = {'frame': [os.path.join(_lib.INCLUDE_DIR, 'dace_b200_reduce.h')],
               'cuda': [os.path.join(_lib.INCLUDE_DIR, 'dace_b200_reduce.h')]}
    # No shared state: every expanded node owns its workspace (node_workspace() below).
    state_fields = []
    init_code = ""
    finalize_code = ""
    dependencies = []


@dace.library.environment
class B200FusedReduce:
    """Header-only: puts ``dace_b200_fused_reduce.cuh`` (kernels instantiated with GENERATED functors: fused
    map -> reduce producers, custom WCR lambdas) into the program's .cu file."""
    cmake_minimum_version = None
    cmake_packages = ["CUDA"]
    cmake_variables = {}
    cmake_includes = []
    cmake_libraries = []
    cmake_compile_flags = []
    cmake_link_flags = []
    cmake_files = []
    headers = {'frame': [], 'cuda': ['dace_b200_fused_reduce.cuh']}
    state_fields = []
    init_code = ""
    finalize_code = ""
    dependencies = [B200Reduce]


_ws_counter = itertools.count()


def node_workspace(sdfg, state, node, bytes_code: str = None, symbols_used=(), zero_init: bool = True):
    """A device workspace owned by ONE expanded node: a pointer + size in the generated frame code,
    ``cudaMalloc``'ed in ``__dace_init_<sdfg>`` and freed in ``__dace_exit_<sdfg>`` - the protocol of the
    reference's CUB temp storage (reference dace/libraries/standard/nodes/reduce.py:514-570), so nothing is
    allocated on the call path and independent nodes that DaCe schedules on different streams
    (``compiler.cuda.max_concurrent_streams``) never share tickets or partials.

    ``bytes_code`` is a C++ block that assigns the wanted size to ``__b200_need`` (e.g. from
    ``dace_b200_reduce_nd_workspace_bytes``); it runs inside ``__dace_init`` where only the top-level SDFG's
    free symbols exist, so it is used only when every symbol it mentions is one of them - otherwise, and by
    default, the size is the base ``dace_b200_workspace_bytes()``.  Returns ``(pointer, bytes)`` C++ expressions."""
    sid = state.parent_graph.node_id(state) if hasattr(state, 'parent_graph') else sdfg.node_id(state)
    name = '__b200_ws_%s_%d_%d_%d' % (sdfg.name, sid, state.node_id(node), next(_ws_counter))
    init_syms = set(str(x) for x in sdfg.free_symbols) if sdfg.parent_sdfg is None else set()
    sized = bytes_code is not None and sdfg.parent_sdfg is None and all(str(x) in init_syms for x in symbols_used)
    sdfg.append_global_code('void *%s = nullptr; size_t %s_bytes = 0;\n' % (name, name), 'frame')
    sdfg.append_init_code("""
    {{
        size_t __b200_need = dace_b200_workspace_bytes();
        {sized}
        {name}_bytes = __b200_need > 0 ? __b200_need : 256;
        if (cudaMalloc(&{name}, {name}_bytes) != cudaSuccess{init}) {{
            __result = 1;
        }}
    }}
""".format(name=name, sized=bytes_code if sized else '',
           init=(' ||\n            dace_b200_workspace_init(%s, %s_bytes, nullptr) != 0 ||\n'
                 '            cudaStreamSynchronize(nullptr) != cudaSuccess' % (name, name)) if zero_init else ''), 'frame')
    sdfg.append_exit_code('    if (%s) { cudaFree(%s); %s = nullptr; }\n' % (name, name, name), 'frame')
    return name, name + '_bytes'


def _resolve_redtype(wcr: str, allow_custom: bool = False) -> ReductionType:
    rt = detect_reduction_type(wcr)
    rt = ReductionType[str(rt).split('.')[-1]]
    if rt == ReductionType.Custom:
        rt = lower_custom(wcr)
    if rt == ReductionType.Custom and allow_custom:
        # `a - b` / `a / b` are "reductions" only for the reference's OpenMP expansion (reduce.py:274-286); they are
        # neither associative nor commutative, so a tree over them is not the reference's result: not taken
        if str(detect_reduction_type(wcr, openmp=True)).split('.')[-1] in ('Sub', 'Div'):
            raise NotImplementedError('Reduce expansion "B200": %s is not associative; use the "pure" expansion' % wcr)
        return rt
    if rt not in _SUPPORTED:
        raise NotImplementedError('Reduce expansion "B200": reduction type %s has no precompiled sm_100a kernel '
                                  '(custom lambdas, argmin/argmax and exchange are not supported)' % rt.name)
    return rt


def _same(a, b) -> bool:
    """Symbolic equality of two size / stride expressions."""
    try:
        d = dace.symbolic.pystr_to_symbolic(a) - dace.symbolic.pystr_to_symbolic(b)
        return bool(dace.symbolic.simplify(d) == 0)
    except Exception:
        return False


def _merge(sizes, strides):
    """(extent, stride) of dims that merge into ONE linear index (stride[i] == stride[i+1] * size[i+1]), else None."""
    if not sizes:
        return 1, 0
    for i in range(len(sizes) - 1):
        if not _same(strides[i], strides[i + 1] * sizes[i + 1]):
            return None
    total = sizes[0]
    for s_ in sizes[1:]:
        total = total * s_
    return total, strides[-1]


def _expand_custom(node, state, sdfg, in_desc, out_desc, shape, istr, reduced, ostr):
    """ReductionType.Custom on the device: the WCR lambda becomes a generated functor (unparsed with the
    reference's own ``unparse_cr_split`` exactly as its CUB expansion does, reference
    dace/libraries/standard/nodes/reduce.py:453-468) and the streaming kernels of
    ``include/dace_b200_fused_reduce.cuh`` are instantiated with it inside the program's .cu file - a lambda
    cannot sit behind a precompiled C ABI.  Served layouts: all axes, trailing axes (rows) or leading axes
    (columns) reduced, each group of axes dense enough to merge; anything else is refused (use 'pure')."""
    from dace.codegen.targets.cpp import unparse_cr_split
    cols, nk, sk, nr, sr, so = custom_layout(shape, istr, reduced, ostr)
    body, (arg1, arg2) = unparse_cr_split(sdfg, node.wcr)
    ident = '%s_%d_%d_%d' % (sdfg.name, state.parent_graph.node_id(state), state.node_id(node), next(_ws_counter))
    tin, tout = in_desc.dtype.ctype, out_desc.dtype.ctype
    if node.identity is None:
        ident_decl, ident_arg = '', '(const %s*)nullptr' % tout
    else:
        ident_decl = 'const %s __identity = %s;' % (tout, symstr(node.identity, cpp_mode=True))
        ident_arg = '&__identity'
    # vector interface of the load functor (include/dace_b200_fused_reduce.cuh, VECTOR LOADS): 32 bytes per access
    kvec = min(8, 32 // in_desc.dtype.bytes) if in_desc.dtype.bytes in (1, 2, 4, 8) else 0
    vec_code = '' if kvec < 2 else """
    static constexpr int kVec = %d;
    using value_type = %s;
    template <int __V>
    __device__ __forceinline__ void vec(size_t __row, size_t __col, %s (&__v)[__V]) const {
        %s __t[__V];
        dace_b200::stream_load<%s, __V>(in + (long long)__row * s_row + (long long)__col, __t);
        #pragma unroll
        for (int __k = 0; __k < __V; ++__k) __v[__k] = (%s)__t[__k];
    }""" % (kvec, tout, tout, tin, tin, tout)
    sig = ('const %s* __in, %s* __out, size_t __kept, size_t __red, long long __s_kept, long long __s_red, '
           'long long __s_out, void* __workspace, size_t __workspace_bytes, cudaStream_t __stream' % (tin, tout))
    sdfg.append_global_code("""
struct __b200_custom_op_{id} {{
    template <typename T>
    __device__ __forceinline__ T operator()(const T &{a}, const T &{b}) const {{
        {body}
    }}
}};
struct __b200_custom_load_{id} {{
    const {tin}* __restrict__ in; long long s_row, s_col;
    __device__ __forceinline__ {tout} operator()(size_t __row, size_t __col) const {{
        return ({tout})in[(long long)__row * s_row + (long long)__col * s_col];
    }}{vec_code}
}};
DACE_EXPORTED cudaError_t __b200_custom_{id}({sig});
cudaError_t __b200_custom_{id}({sig}) {{
    {ident_decl}
    // widest vector access that is legal for this call: unit column stride, rows and base aligned to it
    const long long __s_col = {s_col}, __s_row = {s_row};
    const size_t __n = {n};
    int __max_vec = 1;
    for (int __v = {kvec}; __v >= 2 && __s_col == 1; __v /= 2)
        if (__n % __v == 0 && __s_row % __v == 0 && ((size_t)__in % (__v * sizeof({tin}))) == 0) {{ __max_vec = __v; break; }}
    return {call};
}}
""".format(id=ident, a=arg1, b=arg2, body=body, tin=tin, tout=tout, sig=sig, ident_decl=ident_decl,
           kvec=kvec, vec_code=vec_code, s_col='__s_kept' if cols else '__s_red', s_row='__s_red' if cols else '__s_kept',
           n='__kept' if cols else '__red',
           call=('dace_b200::fused_cols_custom<__b200_custom_op_{id}, {t}, {t}>(__b200_custom_load_{id}{{__in, __s_red, __s_kept}}, '
                 '__out, (ptrdiff_t)__s_out, __red, __kept, {ia}, __workspace, __workspace_bytes, __stream, __max_vec)' if cols else
                 'dace_b200::fused_rows_custom<__b200_custom_op_{id}, {t}, {t}>(__b200_custom_load_{id}{{__in, __s_kept, __s_red}}, '
                 '__out, (ptrdiff_t)__s_out, __kept, __red, {ia}, __workspace, __workspace_bytes, __stream, __max_vec)'
                 ).format(id=ident, t=tout, ia=ident_arg)), 'cuda')
    sdfg.append_global_code('DACE_EXPORTED cudaError_t __b200_custom_%s(%s);\n' % (ident, sig), 'frame')
    ws, ws_bytes = node_workspace(sdfg, state, node)
    code = ('DACE_GPU_CHECK(__b200_custom_%s(_in, _out, (size_t)(%s), (size_t)(%s), (long long)(%s), (long long)(%s), '
            '(long long)(%s), %s, %s, __dace_current_stream));' %
            (ident, symstr(nk, cpp_mode=True), symstr(nr, cpp_mode=True), symstr(sk, cpp_mode=True),
             symstr(sr, cpp_mode=True), symstr(so, cpp_mode=True), ws, ws_bytes))
    return nodes.Tasklet('reduce_b200_custom', {'_in': dace.pointer(in_desc.dtype)}, {'_out': dace.pointer(out_desc.dtype)},
                         code, language=dace.Language.CPP)


def _is_one(expr) -> bool:
    try:
        return bool(expr == 1)
    except TypeError:
        return False


def _memlet_dims(node, in_edge, out_edge, in_desc, out_desc):
    """Input memlet as (extents, element strides, reduced flags) + output strides of the kept axes; compile-time
    unit dims are squeezed like the reference does (reference reduce.py:40-54)."""
    in_subset = copy.deepcopy(in_edge.data.subset)
    in_sizes = in_subset.size()
    input_dims = len(in_sizes)
    axes = list(node.axes) if node.axes is not None else list(range(input_dims))
    axes = [a % input_dims for a in axes]
    shape, istr, reduced = [], [], []
    for i, (rng, size) in enumerate(zip(in_subset.ranges, in_sizes)):
        if _is_one(size):
            continue
        shape.append(size)
        istr.append(in_desc.strides[i] * rng[2])
        reduced.append(i in axes)
    out_subset = copy.deepcopy(out_edge.data.subset)
    ostr = [out_desc.strides[j] * rng[2]
            for j, (rng, size) in enumerate(zip(out_subset.ranges, out_subset.size())) if not _is_one(size)]
    return shape, istr, reduced, ostr


def custom_layout(shape, istr, reduced, ostr):
    """(cols, n_kept, s_kept, n_red, s_red, s_out) of a layout the custom-lambda kernels serve, else
    NotImplementedError: all / trailing (rows) / leading (columns) axes reduced, each group merging into one index."""
    kept = [i for i, r in enumerate(reduced) if not r]
    red = [i for i, r in enumerate(reduced) if r]
    if not red:
        raise NotImplementedError('Reduce expansion "B200": custom WCR without a reduced axis')
    if kept and not (max(kept) < min(red) or max(red) < min(kept)):
        raise NotImplementedError('Reduce expansion "B200": custom WCR lambdas are served for all / trailing / leading '
                                  'reduced axes only (got interleaved axes); use the "pure" expansion')
    cols = bool(kept) and max(red) < min(kept)
    mk = _merge([shape[i] for i in kept], [istr[i] for i in kept])
    mr = _merge([shape[i] for i in red], [istr[i] for i in red])
    mo = _merge([shape[i] for i in kept], list(ostr))
    if mk is None or mr is None or mo is None:
        raise NotImplementedError('Reduce expansion "B200": custom WCR needs kept and reduced axes that each merge into '
                                  'one index (dense sub-ranges); use the "pure" expansion')
    (nk, sk), (nr, sr), (_, so) = mk, mr, mo
    return cols, nk, sk, nr, sr, (so if kept else 1)


@dace.library.register_expansion(Reduce, 'B200')
class ExpandReduceB200(ExpandTransformation):
    """Reduce -> one call into libdace_b200_reduce.so (hand-written sm_100a kernels)."""
    environments = [B200Reduce, B200FusedReduce]

    @staticmethod
    def expansion(node: Reduce, state, sdfg):
        node.validate(sdfg, state)
        in_edge = state.in_edges(node)[0]
        out_edge = state.out_edges(node)[0]
        in_desc = sdfg.arrays[in_edge.data.data]
        out_desc = sdfg.arrays[out_edge.data.data]

        ok_storage = (dtypes.StorageType.GPU_Global, dtypes.StorageType.CPU_Pinned)
        if in_desc.storage not in ok_storage or out_desc.storage not in ok_storage:
            raise ValueError('Reduce expansion "B200" needs input and output in GPU_Global or CPU_Pinned storage '
                             '(got %s -> %s); apply sdfg.apply_gpu_transformations() first. There is no CPU '
                             'fallback.' % (in_desc.storage, out_desc.storage))
        if in_desc.dtype not in _DTYPE_CODES or out_desc.dtype not in _DTYPE_CODES:
            raise NotImplementedError('Reduce expansion "B200": unsupported dtype %s -> %s' %
                                      (in_desc.dtype, out_desc.dtype))
        redtype = _resolve_redtype(node.wcr, allow_custom=True)

        # Input memlet: per-dimension extent and element stride (array stride x range step)
        shape, istr, reduced, ostr = _memlet_dims(node, in_edge, out_edge, in_desc, out_desc)
        n_kept = sum(1 for r in reduced if not r)
        if len(ostr) != n_kept:
            raise ValueError('Reduce expansion "B200": output memlet %s has %d non-unit dimensions, expected %d' %
                             (out_edge.data, len(ostr), n_kept))
        if redtype == ReductionType.Custom:
            tnode = _expand_custom(node, state, sdfg, in_desc, out_desc, shape, istr, reduced, ostr)
            in_edge._dst_conn = '_in'
            out_edge._src_conn = '_out'
            node.add_in_connector('_in')
            node.add_out_connector('_out')
            return tnode
        ndim = len(shape)
        if ndim > 12:
            raise NotImplementedError('Reduce expansion "B200": more than 12 non-unit dimensions')
        mask = sum(1 << i for i, r in enumerate(reduced) if r)

        def carr(name, vals):
            vals = [symstr(v, cpp_mode=True) for v in vals] or ['0']
            return 'const int64_t %s[] = {%s};' % (name, ', '.join('(int64_t)(%s)' % v for v in vals))

        if node.identity is None:
            ident_decl, ident_arg = '', 'nullptr'
        else:
            ident_decl = '%s __b200_identity = %s;' % (out_desc.dtype.ctype, symstr(node.identity, cpp_mode=True))
            ident_arg = '&__b200_identity'
        used = set()
        for v in list(shape) + list(istr):
            used |= set(str(x) for x in getattr(v, 'free_symbols', ()))
        ws, ws_bytes = node_workspace(
            sdfg, state, node, symbols_used=used,
            bytes_code='%s %s __b200_need = dace_b200_reduce_nd_workspace_bytes(%d, %d, %d, __b200_shape, __b200_istr, %du);'
            % (carr('__b200_shape', shape), carr('__b200_istr', istr), _DTYPE_CODES[in_desc.dtype],
               _DTYPE_CODES[out_desc.dtype], ndim, mask))
        code = """
{shape}
{istr}
{ostr}
{ident_decl}
DACE_GPU_CHECK((cudaError_t)dace_b200_reduce_nd({red}, {dtin}, {dtout}, _in, _out, {ndim}, __b200_shape, __b200_istr,
    __b200_ostr, {mask}u, {ident_arg}, {ws}, {ws_bytes},
    (dace_b200_stream_t)__dace_current_stream));
""".format(shape=carr('__b200_shape', shape), istr=carr('__b200_istr', istr), ostr=carr('__b200_ostr', ostr),
           ident_decl=ident_decl, red=int(redtype), dtin=_DTYPE_CODES[in_desc.dtype],
           dtout=_DTYPE_CODES[out_desc.dtype], ndim=ndim, mask=mask, ident_arg=ident_arg, ws=ws, ws_bytes=ws_bytes)

        tnode = nodes.Tasklet('reduce_b200', {'_in': dace.pointer(in_desc.dtype)}, {'_out': dace.pointer(out_desc.dtype)},
                              code, language=dace.Language.CPP)
        # Rename the outer connectors like the reference expansions do (reduce.py:617-620).
        in_edge._dst_conn = '_in'
        out_edge._src_conn = '_out'
        node.add_in_connector('_in')
        node.add_out_connector('_out')
        return tnode


def use_b200_reductions(sdfg=None):
    """Make ``'B200'`` the implementation of every Reduce node: of ``sdfg`` if
    given, else globally via ``Reduce.default_implementation``."""
    if sdfg is None:
        Reduce.default_implementation = 'B200'
        return
    for n, _ in sdfg.all_nodes_recursive():
        if isinstance(n, Reduce):
            n.implementation = 'B200'
