"""DaCe plug-in: in-kernel Reduce expansions ``'B200 (block)'`` and ``'B200 (block allreduce)'``.

They take the place of the reference's ``CUDA (block)`` / ``CUDA (block allreduce)`` expansions
(reference dace/libraries/standard/nodes/reduce.py:625-747 and :750-883), which lower a Reduce node
that sits INSIDE a GPU kernel to ``cub::BlockReduce<T, NTHREADS>(temp).Reduce(in, op)``.  Here the
generated tasklet calls ``dace_b200::block_reduce`` / ``block_allreduce`` from
``include/dace_b200_block_reduce.cuh`` (``redux.sync`` for 32-bit integers, shuffle trees otherwise,
one shared-memory slot per warp).  In-kernel code cannot cross a C ABI, so - like CUB in the
reference - the boundary of this part is a device header that DaCe's own nvcc pass compiles.

Same preconditions and errors as the reference expansion (reduce.py:665-716): the node must be
inside a GPU kernel with a constant block size, reduce all axes of a Register input into a Register
output, and have an identity.  Custom WCR lambdas are supported (a functor is generated like
reduce.py:676-689); argmin/argmax are rejected like the reference (reduce.py:715-716).

* ``'B200 (block)'``            result in thread 0 of the block (CUB semantics);
* ``'B200 (block allreduce)'``  the Reduce node reads a block-shared / global array along ``axes``:
  the node is wrapped in a map over the reduced axes (which becomes the thread-block map), each
  thread takes one element into a register, and the block-wide result is stored once by thread 0
  - the structure the reference builds in reduce.py:783-883, re-done with plain SDFG API calls.
"""
import copy
import functools

import dace
import dace.library
from dace import dtypes, nodes, subsets
from dace.frontend.operations import detect_reduction_type
from dace.libraries.standard.nodes.reduce import Reduce
from dace.sdfg import devicelevel_block_size
from dace.symbolic import issymbolic
from dace.transformation.transformation import ExpandTransformation

from . import _lib
from .redtypes import ReductionType, lower_custom

_BUILTIN = {ReductionType.Min, ReductionType.Max, ReductionType.Sum, ReductionType.Product,
            ReductionType.Logical_And, ReductionType.Bitwise_And, ReductionType.Logical_Or,
            ReductionType.Bitwise_Or, ReductionType.Logical_Xor, ReductionType.Bitwise_Xor}


@dace.library.environment
class B200BlockReduce:
    """Header-only: puts ``dace_b200_block_reduce.cuh`` into the generated .cu file."""
    cmake_minimum_version = None
    cmake_packages = ["CUDA"]
    cmake_variables = {}
    cmake_includes = [_lib.INCLUDE_DIR]
    cmake_libraries = []
    cmake_compile_flags = []
    cmake_link_flags = []
    cmake_files = []
    headers = {'frame': [], 'cuda': ['dace_b200_block_reduce.cuh']}
    state_fields = []
    init_code = ""
    finalize_code = ""
    dependencies = []


def _block_threads(sdfg, state, node) -> int:
    bs = devicelevel_block_size(sdfg, state, node)
    if bs is None:
        raise ValueError('Block-wide GPU reduction must occur within a GPU kernel')
    n = functools.reduce(lambda a, b: a * b, bs, 1)
    if issymbolic(n, sdfg.constants):
        raise ValueError('Block size has to be constant for block-wide reduction (got %s)' % str(n))
    return int(n)


def _emit(node: Reduce, state, sdfg, fn: str):
    """Common part of both expansions: checks + the tasklet that calls ``dace_b200::<fn>``."""
    from dace.codegen.targets.cpp import unparse_cr_split

    node.validate(sdfg, state)
    in_edge = state.in_edges(node)[0]
    out_edge = state.out_edges(node)[0]
    in_desc = sdfg.arrays[in_edge.data.data]
    out_desc = sdfg.arrays[out_edge.data.data]
    if node.identity is None:
        raise ValueError('For device reduce nodes, initial value must be specified')
    nthreads = _block_threads(sdfg, state, node)
    if nthreads > 1024:
        raise ValueError('Block-wide reduction: %d threads per block' % nthreads)
    if node.axes is not None and len(node.axes) < len(in_edge.data.subset):
        raise ValueError('Only full reduction is supported for block-wide reduce, please use the pure expansion')
    if in_desc.storage != dtypes.StorageType.Register or out_desc.storage != dtypes.StorageType.Register:
        raise ValueError('Block-wise reduction only supports GPU register inputs and outputs')

    rt = ReductionType[str(detect_reduction_type(node.wcr)).split('.')[-1]]
    if rt == ReductionType.Custom:
        rt = lower_custom(node.wcr)   # hand-written min/max conditionals; else stays Custom
    if rt in (ReductionType.Min_Location, ReductionType.Max_Location):
        raise ValueError('%s block reduction not supported' % rt)

    ctype = out_desc.dtype.ctype
    state_id = state.parent_graph.node_id(state)
    node_id = state.node_id(node)
    idstr = '%s_%d_%d' % (sdfg.name, state_id, node_id)
    if rt in _BUILTIN:
        functor = 'dace_b200::op<%d, %s>()' % (int(rt), ctype)
    else:
        # Any other lambda: a generated functor, as the reference does for CUB.
        body, (arg1, arg2) = unparse_cr_split(sdfg, node.wcr)
        sdfg.append_global_code("""
struct __b200_reduce_{id} {{
    template <typename T>
    __device__ __forceinline__ T operator()(const T &{a}, const T &{b}) const {{
        {body}
    }}
}};""".format(id=idstr, a=arg1, b=arg2, body=body), 'cuda')
        functor = '__b200_reduce_%s()' % idstr

    total = in_edge.data.subset.num_elements()
    if issymbolic(total, sdfg.constants):
        raise ValueError('Block-wide reduction needs a constant number of register items per thread (got %s)' % total)
    uid = (sdfg.cfg_id * 1000 + state_id) * 10000 + node_id
    # `_in` / `_out` are the connector pointers, already offset to the memlet start (register scalars
    # and register arrays alike).
    if int(total) == 1:
        src = '*_in'
    else:
        src = 'reinterpret_cast<const %s (&)[%d]>(*_in)' % (in_desc.dtype.ctype, int(total))
    code = '*_out = dace_b200::%s<%s, %d, %d>(%s, %s);' % (fn, ctype, nthreads, uid, src, functor)
    tnode = nodes.Tasklet('reduce_b200_block', {'_in': dace.pointer(in_desc.dtype)},
                          {'_out': dace.pointer(out_desc.dtype)}, code, language=dace.Language.CPP)
    in_edge._dst_conn = '_in'
    out_edge._src_conn = '_out'
    node.add_in_connector('_in')
    node.add_out_connector('_out')
    return tnode


@dace.library.register_expansion(Reduce, 'B200 (block)')
class ExpandReduceB200Block(ExpandTransformation):
    """Reduce inside a GPU kernel -> ``dace_b200::block_reduce`` (result in thread 0)."""
    environments = [B200BlockReduce]

    @staticmethod
    def expansion(node: Reduce, state, sdfg):
        return _emit(node, state, sdfg, 'block_reduce')


@dace.library.register_expansion(Reduce, 'B200 (block allreduce)')
class ExpandReduceB200BlockAll(ExpandTransformation):
    """Reduce over an array inside a GPU kernel: wrap in a thread map + registers, then block-reduce."""
    environments = [B200BlockReduce]

    @staticmethod
    def expansion(node: Reduce, state, sdfg):
        in_edge = state.in_edges(node)[0]
        out_edge = state.out_edges(node)[0]
        ndim = len(in_edge.data.subset)
        axes = sorted(a % ndim for a in node.axes) if node.axes is not None else list(range(ndim))
        in_desc = sdfg.arrays[in_edge.data.data]
        out_desc = sdfg.arrays[out_edge.data.data]

        # Thread map over the reduced axes; every thread holds one input element in a register.
        params = ['__b200_t%d' % i for i in range(len(axes))]
        ranges = {p: '%s:%s:%s' % (in_edge.data.subset[a][0], in_edge.data.subset[a][1] + 1, in_edge.data.subset[a][2])
                  for p, a in zip(params, axes)}
        entry, exit_ = state.add_map('inner_reduce_block', ranges, schedule=dtypes.ScheduleType.GPU_ThreadBlock)
        reg_in, _ = sdfg.add_scalar('__b200_rin', in_desc.dtype, storage=dtypes.StorageType.Register, transient=True,
                                    find_new_name=True)
        reg_out, _ = sdfg.add_scalar('__b200_rout', out_desc.dtype, storage=dtypes.StorageType.Register,
                                     transient=True, find_new_name=True)
        rin, rout = state.add_access(reg_in), state.add_access(reg_out)

        elem = subsets.Range([(params[axes.index(i)], params[axes.index(i)], 1) if i in axes else in_edge.data.subset[i]
                              for i in range(ndim)])
        outer_in = copy.deepcopy(in_edge.data)
        outer_out = copy.deepcopy(out_edge.data)
        src, src_conn = in_edge.src, in_edge.src_conn
        dst, dst_conn = out_edge.dst, out_edge.dst_conn
        state.remove_edge(in_edge)
        state.remove_edge(out_edge)
        # Edges are added by hand (no propagation: the scope is inconsistent until all of them exist).
        cin, cout = 'IN_' + outer_in.data, 'OUT_' + outer_in.data
        entry.add_in_connector(cin)
        entry.add_out_connector(cout)
        state.add_edge(src, src_conn, entry, cin, outer_in)
        state.add_edge(entry, cout, rin, None, dace.Memlet(data=outer_in.data, subset=elem))
        state.add_nedge(rin, node, dace.Memlet(data=reg_in, subset='0'))
        state.add_nedge(node, rout, dace.Memlet(data=reg_out, subset='0'))
        cond = ' and '.join('%s == %s' % (p, in_edge.data.subset[a][0]) for p, a in zip(params, axes))
        writer = state.add_tasklet('block_reduce_write', {'inp'}, {'out'}, 'if %s:\n    out = inp' % cond)
        state.add_edge(rout, None, writer, 'inp', dace.Memlet(data=reg_out, subset='0'))
        wm = copy.deepcopy(outer_out)
        wm.volume, wm.dynamic = 0, True
        xin, xout = 'IN_' + outer_out.data, 'OUT_' + outer_out.data
        exit_.add_in_connector(xin)
        exit_.add_out_connector(xout)
        state.add_edge(writer, 'out', exit_, xin, wm)
        state.add_edge(exit_, xout, dst, dst_conn, outer_out)

        node.axes = None
        node.implementation = 'B200 (block)'
        # The barrier inside block_reduce needs every thread of the block: the new thread map has
        # to BE the block (constant size, no larger sibling thread-block map in the kernel).
        mine = functools.reduce(lambda a, b: a * b, entry.map.range.size(), 1)
        if issymbolic(mine, sdfg.constants):
            raise ValueError('Block-wide reduction needs constant extents along the reduced axes (got %s)' % mine)
        if _block_threads(sdfg, state, node) != int(mine):
            raise ValueError('Block-wide reduction over %s elements inside a kernel with %d threads per block' %
                             (mine, _block_threads(sdfg, state, node)))
        return _emit(node, state, sdfg, 'block_reduce')
