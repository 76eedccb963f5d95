"""DaCe plug-in: the ``'B200'`` expansion of ``dace.libraries.standard.Reduce``.

Importing this module (DaCe must be importable) registers

* ``B200Reduce`` - a ``@dace.library.environment`` that adds ``include/`` to the
  include path, links ``libdace_b200_reduce.so`` (with an rpath), includes the C
  header into the generated frame code and keeps one device workspace per
  compiled SDFG in the state struct (reference environment mechanics:
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
There is no CPU fallback: data that is not in ``GPU_Global`` / ``CPU_Pinned``
storage, custom lambdas and argmin/argmax raise at expansion time.
"""
import copy
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
    # Workspace protocol of the CUB expansion (allocate once at init, free at
    # exit; reference reduce.py:550-570), but per compiled SDFG instead of file-scope globals.
    state_fields = ["void *dace_b200_workspace = nullptr;"]
    init_code = """
    if (cudaMalloc(&__state->dace_b200_workspace, dace_b200_workspace_bytes()) != cudaSuccess ||
        dace_b200_workspace_init(__state->dace_b200_workspace, dace_b200_workspace_bytes(), nullptr) != 0 ||
        cudaStreamSynchronize(nullptr) != cudaSuccess) {
        __result = 1;
    }
"""
    finalize_code = """
    if (__state->dace_b200_workspace) cudaFree(__state->dace_b200_workspace);
"""
    dependencies = []


def _resolve_redtype(wcr: str) -> ReductionType:
    rt = detect_reduction_type(wcr)
    rt = ReductionType[str(rt).split('.')[-1]]
    if rt == ReductionType.Custom:
        rt = lower_custom(wcr)
    if rt not in _SUPPORTED:
        raise NotImplementedError('Reduce expansion "B200": reduction type %s has no precompiled sm_100a kernel '
                                  '(custom lambdas, argmin/argmax and exchange are not supported)' % rt.name)
    return rt


def _is_one(expr) -> bool:
    try:
        return bool(expr == 1)
    except TypeError:
        return False


@dace.library.register_expansion(Reduce, 'B200')
class ExpandReduceB200(ExpandTransformation):
    """Reduce -> one call into libdace_b200_reduce.so (hand-written sm_100a kernels)."""
    environments = [B200Reduce]

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
        redtype = _resolve_redtype(node.wcr)

        # Input memlet: per-dimension extent and element stride (array stride x range step);
        # compile-time unit dims are squeezed like the reference does (reduce.py:40-54).
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
        n_kept = sum(1 for r in reduced if not r)
        if len(ostr) != n_kept:
            raise ValueError('Reduce expansion "B200": output memlet %s has %d non-unit dimensions, expected %d' %
                             (out_edge.data, len(ostr), n_kept))
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
        code = """
{shape}
{istr}
{ostr}
{ident_decl}
DACE_GPU_CHECK((cudaError_t)dace_b200_reduce_nd({red}, {dtin}, {dtout}, _in, _out, {ndim}, __b200_shape, __b200_istr,
    __b200_ostr, {mask}u, {ident_arg}, __state->dace_b200_workspace, dace_b200_workspace_bytes(),
    (dace_b200_stream_t)__dace_current_stream));
""".format(shape=carr('__b200_shape', shape), istr=carr('__b200_istr', istr), ostr=carr('__b200_ostr', ostr),
           ident_decl=ident_decl, red=int(redtype), dtin=_DTYPE_CODES[in_desc.dtype],
           dtout=_DTYPE_CODES[out_desc.dtype], ndim=ndim, mask=mask, ident_arg=ident_arg)

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
