"""DaCe plug-in: ``np.argmax`` / ``np.argmin`` in ``@dace.program``s -> ``dace_b200_reduce_arg``.

The reference frontend lowers argmin/argmax to a map with a struct-valued WCR
(reference dace/frontend/python/replacements/reduction.py:276-348):

    init state:   reduced[o] = _val_and_idx(val=min_value / max_value, idx=-1)
    reduce map:   reduced[o] (WCR) = _val_and_idx(idx=__i<axis>, val=A[...])
                  wcr: lambda x, y: _val_and_idx(val=max(x.val, y.val), idx=y.idx if x.val < y.val else x.idx)

which on a GPU becomes one compare-and-swap loop per element
(``dace::wcr_custom<T>::reduce_atomic``, reference dace/runtime/include/dace/reduction.h:56-88) - the
slowest WCR case of the reference.  ``ArgReduceMapToLibrary`` recognises that map
and puts an ``ArgReduce`` library node in its place; the node's ``'B200'`` expansion calls
``dace_b200_reduce_arg`` with the two outputs pointing at the ``idx`` / ``val`` fields of the struct
array (byte strides = sizeof(struct)) and ``init == NULL``, i.e. it folds into the pairs the init state
wrote - the program around the map (init state, extraction of ``.idx``) stays exactly as generated.
"""
import ctypes
import re

import dace
import dace.library
from dace import dtypes, nodes, properties
from dace.sdfg import utils as sdutil
from dace.symbolic import symstr
from dace.transformation import transformation as xf
from dace.transformation.transformation import ExpandTransformation

from .expansion import B200Reduce, _DTYPE_CODES, node_workspace
from .redtypes import ReductionType

_CODE_RE = re.compile(r'^\s*__out\s*=\s*_val_and_idx\(\s*idx\s*=\s*(\w+)\s*,\s*val\s*=\s*__in\s*\)\s*;?\s*$')
_WCR_RE = re.compile(r'_val_and_idx\(\s*val\s*=\s*(max|min)\(x\.val,\s*y\.val\)')
_IDX_CODES = {dace.int32: 2, dace.int64: 3}


def _struct_layout(stype):
    """(offset of idx, offset of val, sizeof) of the C struct the code generator emits for ``stype``
    (natural alignment, fields in declaration order - what ctypes.Structure computes too)."""
    ct = stype.as_ctypes()
    return getattr(ct, 'idx').offset, getattr(ct, 'val').offset, ctypes.sizeof(ct)


def _packed_full(memlet, desc) -> bool:
    from .comm_node import _full_contiguous
    return _full_contiguous(memlet, desc)


class ExpandArgReduceB200(ExpandTransformation):
    """ArgReduce -> one ``dace_b200_reduce_arg`` call on the program's stream."""
    environments = [B200Reduce]

    @staticmethod
    def expansion(node, state, sdfg):
        node.validate(sdfg, state)
        ie, oe = state.in_edges(node)[0], state.out_edges(node)[0]
        in_desc, out_desc = sdfg.arrays[ie.data.data], sdfg.arrays[oe.data.data]
        ok_storage = (dtypes.StorageType.GPU_Global, dtypes.StorageType.CPU_Pinned)
        if in_desc.storage not in ok_storage or out_desc.storage not in ok_storage:
            raise ValueError('ArgReduce expansion "B200" needs GPU_Global or CPU_Pinned storage (got %s -> %s). There is '
                             'no CPU fallback.' % (in_desc.storage, out_desc.storage))
        stype = out_desc.dtype
        if not isinstance(stype, dtypes.struct) or list(stype.fields.keys()) != ['idx', 'val']:
            raise TypeError('ArgReduce: output must be an array of struct {idx; val} (got %s)' % stype)
        idx_t, val_t = stype.fields['idx'], stype.fields['val']
        if val_t != in_desc.dtype or in_desc.dtype not in _DTYPE_CODES or idx_t not in _IDX_CODES:
            raise NotImplementedError('ArgReduce expansion "B200": %s -> {%s, %s}' % (in_desc.dtype, idx_t, val_t))
        if not _packed_full(ie.data, in_desc) or not _packed_full(oe.data, out_desc):
            raise NotImplementedError('ArgReduce expansion "B200": input and output must be whole contiguous arrays')
        shape = list(in_desc.shape)
        axis = node.axis % len(shape)
        outer, inner = 1, 1
        for s in shape[:axis]:
            outer = outer * s
        for s in shape[axis + 1:]:
            inner = inner * s
        reduced = shape[axis]
        off_idx, off_val, size = _struct_layout(stype)
        red = int(ReductionType.Max_Location if node.func == 'max' else ReductionType.Min_Location)
        ws, ws_bytes = node_workspace(sdfg, state, node)
        code = """
DACE_GPU_CHECK((cudaError_t)dace_b200_reduce_arg({red}, {dtin}, {dtidx}, _in, (char *)_out + {off_idx}, (char *)_out + {off_val},
    (size_t)({outer}), (size_t)({reduced}), (size_t)({inner}), (int64_t)(({reduced}) * ({inner})), (int64_t)({inner}), (int64_t)1,
    (size_t){size}, (size_t){size}, nullptr /* fold into the pairs written by the init state */,
    {ws}, {ws_bytes}, (dace_b200_stream_t)__dace_current_stream));
""".format(ws=ws, ws_bytes=ws_bytes, red=red, dtin=_DTYPE_CODES[in_desc.dtype], dtidx=_IDX_CODES[idx_t], off_idx=off_idx, off_val=off_val,
           outer=symstr(outer, cpp_mode=True), reduced=symstr(reduced, cpp_mode=True),
           inner=symstr(inner, cpp_mode=True), size=size)
        return nodes.Tasklet('arg_reduce_b200', {'_in': dace.pointer(in_desc.dtype)}, {'_out': dace.pointer(stype)}, code,
                             language=dace.Language.CPP)


@dace.library.node
class ArgReduce(nodes.LibraryNode):
    """``_out[o, i] = fold of (val, idx) pairs of _in[o, :, i]`` along ``axis`` with the frontend's argmax /
    argmin rule (strictly better replaces, first occurrence wins), starting from the pairs in ``_out``."""

    implementations = {'B200': ExpandArgReduceB200}
    default_implementation = 'B200'

    func = properties.Property(dtype=str, default='max', choices=['max', 'min'])
    axis = properties.Property(dtype=int, default=0)

    def __init__(self, name='arg_reduce', func='max', axis=0, **kwargs):
        super().__init__(name, inputs={'_in'}, outputs={'_out'}, **kwargs)
        self.func = func
        self.axis = axis

    @staticmethod
    def from_json(json_obj, context=None):
        ret = ArgReduce('arg_reduce')
        dace.serialize.set_properties_from_json(ret, json_obj, context=context)
        return ret

    def __str__(self):
        return 'ArgReduce (arg%s, axis %d)' % (self.func, self.axis)

    def validate(self, sdfg, state):
        ins, outs = state.in_edges(self), state.out_edges(self)
        if len(ins) != 1 or len(outs) != 1 or ins[0].dst_conn != '_in' or outs[0].src_conn != '_out':
            raise ValueError('ArgReduce needs exactly one _in input and one _out output')
        nd = len(sdfg.arrays[ins[0].data.data].shape)
        if not -nd <= self.axis < nd:
            raise ValueError('ArgReduce: axis %d out of range for %d dimensions' % (self.axis, nd))


def _match(state, sdfg, entry):
    """(func, axis, in_edge, out_edge, tasklet) when ``entry`` is the frontend's arg-reduce map, else None."""
    scope = state.scope_subgraph(entry, include_entry=False, include_exit=False)
    tasklets = [n for n in scope.nodes()]
    if len(tasklets) != 1 or not isinstance(tasklets[0], nodes.Tasklet):
        return None
    t = tasklets[0]
    if t.language != dace.Language.Python or set(t.in_connectors) != {'__in'} or set(t.out_connectors) != {'__out'}:
        return None
    m = _CODE_RE.match(t.code.as_string.strip())
    if m is None:
        return None
    exit_ = state.exit_node(entry)
    inner_out = state.out_edges(t)[0]
    w = _WCR_RE.search(inner_out.data.wcr or '')
    if w is None or 'y.idx if x.val %s y.val else x.idx' % ('<' if w.group(1) == 'max' else '>') not in inner_out.data.wcr:
        return None
    inner_in = state.in_edges(t)[0]
    idx_param = m.group(1)
    params = list(entry.map.params)
    if idx_param not in params or len(inner_in.data.subset) != len(params):
        return None
    # the tasklet reads A[__i0, ..., __i(n-1)] with every map parameter in its own position, over the whole array
    desc = sdfg.arrays[inner_in.data.data]
    for d, (p, (b, e, s)) in enumerate(zip(params, inner_in.data.subset.ranges)):
        if str(b) != p or str(e) != p:
            return None
        rb, re_, rs = entry.map.range.ranges[d]
        if (rb != 0) is True or (rs != 1) is True or (re_ + 1 != desc.shape[d]) is True:
            return None
    outer_in = [e for e in state.in_edges(entry) if e.data.data == inner_in.data.data]
    outer_out = [e for e in state.out_edges(exit_) if e.data.data == inner_out.data.data]
    if len(outer_in) != 1 or len(outer_out) != 1 or len(state.in_edges(entry)) != 1 or len(state.out_edges(exit_)) != 1:
        return None
    if not isinstance(sdfg.arrays[inner_out.data.data].dtype, dtypes.struct):
        return None
    return w.group(1), params.index(idx_param), outer_in[0], outer_out[0], t


@properties.make_properties
class ArgReduceMapToLibrary(xf.SingleStateTransformation):
    """The frontend's ``_arg{max,min}_reduce_`` struct-WCR map -> ``ArgReduce`` library node."""

    map_entry = xf.PatternNode(nodes.MapEntry)

    @classmethod
    def expressions(cls):
        return [sdutil.node_path_graph(cls.map_entry)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        return _match(graph, sdfg, self.map_entry) is not None

    def apply(self, graph, sdfg):
        func, axis, outer_in, outer_out, tasklet = _match(graph, sdfg, self.map_entry)
        entry = self.map_entry
        exit_ = graph.exit_node(entry)
        node = ArgReduce('arg%s_%s' % (func, outer_in.data.data), func=func, axis=axis)
        node.implementation = 'B200'
        graph.add_node(node)
        in_desc, out_desc = sdfg.arrays[outer_in.data.data], sdfg.arrays[outer_out.data.data]
        graph.add_edge(outer_in.src, outer_in.src_conn, node, '_in', dace.Memlet.from_array(outer_in.data.data, in_desc))
        # plain output: the library call folds into the existing pairs itself (a WCR memlet on a pointer
        # connector would make the code generator emit its own atomics)
        graph.add_edge(node, '_out', outer_out.dst, outer_out.dst_conn,
                       dace.Memlet.from_array(outer_out.data.data, out_desc))
        graph.remove_nodes_from([entry, tasklet, exit_])
        return node


_IDX_OF_PAIR = re.compile(r'\s*(\w+)\[0\]\.idx\s*')


def lower_flat_arg_results(sdfg) -> int:
    """``np.argmax(A)`` / ``np.argmin(A)`` over ALL elements: keep the result on the data-flow side.

    The frontend reduces a flat view of ``A`` into a one-element (value, index) transient and then reads
    ``pair[0].idx`` on an INTERSTATE edge into a symbol that the next state stores (reference
    dace/frontend/python/replacements/reduction.py:276-348).  On a GPU that edge is host code reading a device
    container, and the reference's own ``apply_gpu_transformations`` rejects the program
    (``InvalidSDFGInterstateEdgeError``).  Here the symbol is replaced by a memlet: the storing tasklet of the next
    state reads ``pair[0]`` itself and takes ``.idx`` - after that the GPU transformation goes through, and the
    reduction map is the ordinary 1-D arg-reduce map that :class:`ArgReduceMapToLibrary` takes.

    Call BEFORE ``apply_gpu_transformations`` (``dace_b200.auto.auto_optimize`` does).  Only the frontend's exact
    shape is rewritten: the symbol has one use, a tasklet ``__out = <symbol>`` without inputs.  Returns how many."""
    done = 0
    for ise in list(sdfg.all_interstate_edges()):
        for sym, expr in list(ise.data.assignments.items()):
            m = _IDX_OF_PAIR.fullmatch(expr)
            if m is None or m.group(1) not in sdfg.arrays:
                continue
            arr = m.group(1)
            desc = sdfg.arrays[arr]
            if not desc.transient or not isinstance(desc.dtype, dtypes.struct):
                continue
            dst = ise.dst
            if not isinstance(dst, dace.SDFGState):
                continue
            users = [n for n in dst.nodes() if isinstance(n, nodes.Tasklet)]
            if len(users) != 1 or users[0].in_connectors or users[0].code.as_string.strip() != '__out = %s' % sym:
                continue
            if any(sym in other.data.free_symbols for other in sdfg.all_interstate_edges() if other is not ise):
                continue
            if any(sym in map(str, st.free_symbols) for st in sdfg.states() if st is not dst):
                continue
            t = users[0]
            t.add_in_connector('__pair')
            t.code = dace.properties.CodeBlock('__out = __pair.idx')
            dst.add_edge(dst.add_read(arr), None, t, '__pair', dace.Memlet('%s[0]' % arr))
            del ise.data.assignments[sym]
            if sym in sdfg.symbols:
                sdfg.remove_symbol(sym)
            done += 1
    return done


def use_b200_argreduce(sdfg) -> int:
    """Rewrite every argmin/argmax map the frontend generated; returns how many."""
    return sdfg.apply_transformations_repeated(ArgReduceMapToLibrary, validate=True)
