"""DaCe plug-in: the ``ScatterReduce`` library node and the transformation that
creates it from WCR maps (scatter-add / histogram).

The reference has no library node for ``bins[idx[i]] = wcr(bins[idx[i]], v[i])``:
the frontend emits a map whose tasklet writes through a dynamic WCR memlet and
the code generator turns every element into one ``reduce_atomic`` (reference
dace/codegen/targets/cpu.py:965-1013, dace/codegen/targets/cpp.py:1133-1208,
dace/runtime/include/dace/reduction.h:167-340).  SURVEY.md 8(b) names the boundary
to create: a ``@dace.library.node`` plus a ``PatternTransformation`` that replaces
``MapEntry -> Tasklet -> (WCR memlet) -> MapExit`` with it.  This module is that:

* :class:`ScatterReduce` - library node with connectors ``_idx`` (1-D indices),
  optional ``_val`` (1-D values, same length) and ``_bins`` (the array updated in
  place); properties ``wcr`` and ``const_value`` (used when there is no ``_val``).
  Expansions: ``'pure'`` rebuilds the reference's own WCR map (any storage, any
  lambda - it is the reference code path again), ``'B200'`` is one call to
  ``dace_b200_wcr_scatter`` (include/dace_b200_reduce.h) and, like the Reduce
  expansion, has no CPU fallback.
* :class:`WCRMapToScatterReduce` - matches both shapes the Python frontend
  produces: the explicit tasklet ``out[k] = v`` with ``out >> bins(1, wcr)[:]``
  (reference samples/explicit/histogram.py:13-26) and the nested SDFG that
  ``bins[idx[i]] += v[i]`` lowers to (reference tests/npbench/misc/azimint_hist_test.py:39-59).

Registered as DaCe library ``dace_b200`` (reference dace/library.py:84-101).
"""
import ast
import copy

import dace
import dace.library
from dace import dtypes, nodes, subsets, symbolic
from dace.frontend.operations import detect_reduction_type
from dace.memlet import Memlet
from dace.properties import LambdaProperty, Property
from dace.sdfg import utils as sdutil
from dace.symbolic import symstr
from dace.transformation import transformation as xf
from dace.transformation.transformation import ExpandTransformation

from .expansion import B200Reduce, _DTYPE_CODES, _resolve_redtype, node_workspace
from .redtypes import ReductionType

_SCATTER_OPS = {ReductionType.Sum, ReductionType.Min, ReductionType.Max, ReductionType.Product,
                ReductionType.Bitwise_And, ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor}

_IDX_DTYPES = (dace.int32, dace.uint32, dace.int64, dace.uint8)
_VAL_DTYPES = (dace.int32, dace.uint32, dace.int64, dace.uint64, dace.float32, dace.float64)


def _edges_of(node, state):
    ins = {e.dst_conn: e for e in state.in_edges(node)}
    outs = {e.src_conn: e for e in state.out_edges(node)}
    return ins, outs


def _one_dim(memlet: Memlet, what: str):
    """(extent, step) of the only non-unit dimension of a memlet subset."""
    sizes = memlet.subset.size()
    dims = [i for i, s in enumerate(sizes) if not (s == 1) is True]
    if len(dims) > 1:
        raise ValueError('ScatterReduce: %s memlet %s must be one-dimensional' % (what, memlet))
    if not dims:
        return 1, 1, 0
    d = dims[0]
    return sizes[d], memlet.subset.ranges[d][2], d


@dace.library.expansion
class ExpandScatterPure(ExpandTransformation):
    """Back to the reference's own lowering: a map with a WCR tasklet."""
    environments = []

    @staticmethod
    def expansion(node, state, sdfg):
        node.validate(sdfg, state)
        ins, outs = _edges_of(node, state)
        idx_desc = sdfg.arrays[ins['_idx'].data.data]
        bins_desc = sdfg.arrays[outs['_bins'].data.data]
        n, _, _ = _one_dim(ins['_idx'].data, 'index')
        nbins, _, _ = _one_dim(outs['_bins'].data, 'bins')

        nsdfg = dace.SDFG('scatter_reduce_pure')
        nsdfg.add_array('_idx', [n], idx_desc.dtype, storage=idx_desc.storage, strides=[_stride_of(ins['_idx'], sdfg)])
        nsdfg.add_array('_bins', [nbins], bins_desc.dtype, storage=bins_desc.storage,
                        strides=[_stride_of(outs['_bins'], sdfg)])
        has_val = '_val' in ins
        if has_val:
            val_desc = sdfg.arrays[ins['_val'].data.data]
            nsdfg.add_array('_val', [n], val_desc.dtype, storage=val_desc.storage, strides=[_stride_of(ins['_val'], sdfg)])
        nstate = nsdfg.add_state('scatter')
        schedule = node.schedule if node.schedule != dtypes.ScheduleType.Default else dtypes.ScheduleType.Default
        if idx_desc.storage == dtypes.StorageType.GPU_Global:
            schedule = dtypes.ScheduleType.GPU_Device
        me, mx = nstate.add_map('scatter_map', {'__i': '0:%s' % symstr(n)}, schedule=schedule)
        value = '__v' if has_val else str(node.const_value if node.const_value is not None else 1)
        t = nstate.add_tasklet('scatter', {'__k': None, '__v': None} if has_val else {'__k': None}, {'__out': None},
                               '__out[__k] = %s' % value)
        nstate.add_memlet_path(nstate.add_read('_idx'), me, t, dst_conn='__k', memlet=Memlet('_idx[__i]'))
        if has_val:
            nstate.add_memlet_path(nstate.add_read('_val'), me, t, dst_conn='__v', memlet=Memlet('_val[__i]'))
        out_memlet = Memlet('_bins[0:%s]' % symstr(nbins), wcr=node.wcr, volume=1)
        nstate.add_memlet_path(t, mx, nstate.add_write('_bins'), src_conn='__out', memlet=out_memlet)
        outs['_bins'].data.wcr = None   # the update happens inside, like the frontend's own nested form
        return nsdfg


def _stride_of(edge, sdfg):
    """Element stride of the 1-D view a memlet describes (array stride x range step)."""
    desc = sdfg.arrays[edge.data.data]
    _, step, d = _one_dim(edge.data, 'ScatterReduce')
    return desc.strides[d] * step


@dace.library.expansion
class ExpandScatterB200(ExpandTransformation):
    """ScatterReduce -> one call into libdace_b200_reduce.so (shared-memory privatised /
    radix-partitioned / global-atomic sm_100a kernels, chosen inside the library)."""
    environments = [B200Reduce]

    @staticmethod
    def expansion(node, state, sdfg):
        node.validate(sdfg, state)
        ins, outs = _edges_of(node, state)
        idx_edge, bins_edge = ins['_idx'], outs['_bins']
        val_edge = ins.get('_val')
        idx_desc = sdfg.arrays[idx_edge.data.data]
        bins_desc = sdfg.arrays[bins_edge.data.data]
        ok_storage = (dtypes.StorageType.GPU_Global, dtypes.StorageType.CPU_Pinned)
        for e in (idx_edge, bins_edge) + ((val_edge, ) if val_edge else ()):
            if sdfg.arrays[e.data.data].storage not in ok_storage:
                raise ValueError('ScatterReduce expansion "B200" needs %s in GPU_Global or CPU_Pinned storage (got %s); '
                                 'apply sdfg.apply_gpu_transformations() first. There is no CPU fallback.' %
                                 (e.data.data, sdfg.arrays[e.data.data].storage))
            if not (_stride_of(e, sdfg) == 1) is True:
                raise NotImplementedError('ScatterReduce expansion "B200": %s must be contiguous' % e.data)
        if idx_desc.dtype not in _IDX_DTYPES:
            raise NotImplementedError('ScatterReduce expansion "B200": index dtype %s' % idx_desc.dtype)
        if bins_desc.dtype not in _VAL_DTYPES:
            raise NotImplementedError('ScatterReduce expansion "B200": bins dtype %s' % bins_desc.dtype)
        if val_edge is not None and sdfg.arrays[val_edge.data.data].dtype != bins_desc.dtype:
            raise NotImplementedError('ScatterReduce expansion "B200": values (%s) and bins (%s) must have one dtype' %
                                      (sdfg.arrays[val_edge.data.data].dtype, bins_desc.dtype))
        redtype = _resolve_redtype(node.wcr)
        if redtype not in _SCATTER_OPS:
            # dace_b200_wcr_scatter has no logical operators: refuse here, not with cudaErrorNotSupported at run time
            raise NotImplementedError('ScatterReduce expansion "B200": reduction type %s has no scatter kernel' % redtype.name)
        n, _, _ = _one_dim(idx_edge.data, 'index')
        nbins, _, _ = _one_dim(bins_edge.data, 'bins')

        if val_edge is None:
            const = node.const_value if node.const_value is not None else 1
            const_decl = '%s __b200_const = %s;' % (bins_desc.dtype.ctype, symstr(const, cpp_mode=True))
            val_arg, const_arg = 'nullptr', '&__b200_const'
        else:
            const_decl, val_arg, const_arg = '', '_val', 'nullptr'
        # Scratch of the radix-partition path (large bin arrays): sized in __dace_init from the SDFG's
        # symbols, owned by this node; 0 bytes wanted = the path does not apply (a 256-byte stub is kept).
        used = set(str(x) for v in (n, nbins) for x in getattr(v, 'free_symbols', ()))
        ws, ws_bytes = node_workspace(
            sdfg, state, node, symbols_used=used, zero_init=False,
            bytes_code='__b200_need = dace_b200_wcr_scatter_workspace_bytes(%d, %d, %d, %d, (size_t)(%s), (size_t)(%s));'
            % (int(redtype), _DTYPE_CODES[idx_desc.dtype], _DTYPE_CODES[bins_desc.dtype], 0 if val_edge is None else 1,
               symstr(n, cpp_mode=True), symstr(nbins, cpp_mode=True)))
        code = """
{const_decl}
DACE_GPU_CHECK((cudaError_t)dace_b200_wcr_scatter({red}, {dtidx}, {dtval}, _idx, {val_arg}, {const_arg}, (size_t)({n}),
    _bins, (size_t)({nbins}), {ws}, {ws_bytes}, (dace_b200_stream_t)__dace_current_stream));
""".format(const_decl=const_decl, red=int(redtype), dtidx=_DTYPE_CODES[idx_desc.dtype],
           dtval=_DTYPE_CODES[bins_desc.dtype], val_arg=val_arg, const_arg=const_arg,
           n=symstr(n, cpp_mode=True), nbins=symstr(nbins, cpp_mode=True), ws=ws, ws_bytes=ws_bytes)
        in_conns = {'_idx': dace.pointer(idx_desc.dtype)}
        if val_edge is not None:
            in_conns['_val'] = dace.pointer(bins_desc.dtype)
        tnode = nodes.Tasklet('scatter_reduce_b200', in_conns, {'_bins': dace.pointer(bins_desc.dtype)}, code,
                              language=dace.Language.CPP)
        # The library call updates the bins in place; for the code generator this is a plain
        # pointer output (a WCR memlet on a pointer connector would make it emit its own atomic).
        bins_edge.data.wcr = None
        return tnode


@dace.library.node
class ScatterReduce(nodes.LibraryNode):
    """``_bins[_idx[i]] = wcr(_bins[_idx[i]], _val[i] or const_value)`` for every i."""

    implementations = {'pure': ExpandScatterPure, 'B200': ExpandScatterB200}
    default_implementation = 'pure'

    wcr = LambdaProperty(default='lambda a, b: a + b')
    const_value = Property(allow_none=True, default=None, to_json=lambda x: None if x is None else str(x),
                           desc='value applied for every index when there is no _val input (default 1)')

    def __init__(self, name='scatter_reduce', wcr='lambda a, b: a + b', const_value=None, has_values=False,
                 schedule=dtypes.ScheduleType.Default, **kwargs):
        super().__init__(name, inputs={'_idx', '_val'} if has_values else {'_idx'}, outputs={'_bins'}, **kwargs)
        self.wcr = wcr
        self.const_value = const_value
        self.schedule = schedule

    @staticmethod
    def from_json(json_obj, context=None):
        ret = ScatterReduce('scatter_reduce')
        dace.serialize.set_properties_from_json(ret, json_obj, context=context)
        return ret

    def __str__(self):
        red = str(detect_reduction_type(self.wcr)).split('.')[-1]
        return 'ScatterReduce (%s)' % red

    def validate(self, sdfg, state):
        ins, outs = _edges_of(self, state)
        if '_idx' not in ins or set(ins) - {'_idx', '_val'}:
            raise ValueError('ScatterReduce needs an _idx input and at most a _val input')
        if set(outs) != {'_bins'}:
            raise ValueError('ScatterReduce needs exactly one output, _bins')
        n, _, _ = _one_dim(ins['_idx'].data, 'index')
        _one_dim(outs['_bins'].data, 'bins')
        if '_val' in ins:
            nv, _, _ = _one_dim(ins['_val'].data, 'value')
            if (symbolic.pystr_to_symbolic(str(nv)) - symbolic.pystr_to_symbolic(str(n))) != 0:
                raise ValueError('ScatterReduce: %s indices but %s values' % (n, nv))
        if not sdfg.arrays[ins['_idx'].data.data].dtype in dtypes.INTEGER_TYPES:
            raise TypeError('ScatterReduce: indices must be integers')


# ----------------------------------------------------------------------------------------------
# WCR map -> ScatterReduce
# ----------------------------------------------------------------------------------------------

def _const_of(node):
    """Python constant of an AST expression like ``1``, ``-2.5`` or ``True``; None if it is not one."""
    if isinstance(node, ast.Constant) and isinstance(node.value, (int, float, bool)):
        return node.value
    if isinstance(node, ast.UnaryOp) and isinstance(node.op, (ast.USub, ast.UAdd)):
        inner = _const_of(node.operand)
        if inner is not None:
            return -inner if isinstance(node.op, ast.USub) else inner
    return None


def _single_assign(tasklet):
    if not isinstance(tasklet, nodes.Tasklet) or tasklet.language != dtypes.Language.Python:
        return None
    try:
        body = ast.parse(tasklet.code.as_string).body
    except SyntaxError:
        return None
    if len(body) != 1 or not isinstance(body[0], ast.Assign) or len(body[0].targets) != 1:
        return None
    return body[0]


def _parse_indexed_store(tasklet):
    """``out[k] = v`` / ``out[k] = <const>`` -> (out_conn, idx_conn, val_conn | None, const | None)."""
    stmt = _single_assign(tasklet)
    if stmt is None or not isinstance(stmt.targets[0], ast.Subscript):
        return None
    tgt = stmt.targets[0]
    index = tgt.slice.value if isinstance(tgt.slice, getattr(ast, 'Index', ())) else tgt.slice
    if not isinstance(tgt.value, ast.Name) or not isinstance(index, ast.Name):
        return None
    out_conn, idx_conn = tgt.value.id, index.id
    if out_conn not in tasklet.out_connectors or idx_conn not in tasklet.in_connectors or len(tasklet.out_connectors) != 1:
        return None
    const = _const_of(stmt.value)
    if const is not None:
        return (out_conn, idx_conn, None, const) if set(tasklet.in_connectors) == {idx_conn} else None
    if isinstance(stmt.value, ast.Name) and stmt.value.id in tasklet.in_connectors and stmt.value.id != idx_conn:
        if set(tasklet.in_connectors) == {idx_conn, stmt.value.id}:
            return out_conn, idx_conn, stmt.value.id, None
    return None


def _parse_plain_store(tasklet):
    """``__out = __inp`` / ``__out = <const>`` -> (out_conn, val_conn | None, const | None)."""
    stmt = _single_assign(tasklet)
    if stmt is None or not isinstance(stmt.targets[0], ast.Name) or len(tasklet.out_connectors) != 1:
        return None
    out_conn = stmt.targets[0].id
    if out_conn not in tasklet.out_connectors:
        return None
    const = _const_of(stmt.value)
    if const is not None:
        return (out_conn, None, const) if not tasklet.in_connectors else None
    if isinstance(stmt.value, ast.Name) and set(tasklet.in_connectors) == {stmt.value.id}:
        return out_conn, stmt.value.id, None
    return None


def _affine_in_param(memlet, desc, param):
    """For a one-element memlet ``arr[param + c]`` on a 1-D array return c, else None."""
    if memlet is None or memlet.is_empty() or len(desc.shape) != 1 or not isinstance(memlet.subset, subsets.Range):
        return None
    (b, e, s), = memlet.subset.ranges
    if not (symbolic.simplify(e - b) == 0) is True:
        return None
    psym = symbolic.pystr_to_symbolic(param)
    c = symbolic.simplify(symbolic.pystr_to_symbolic(str(b)) - psym)
    if psym in getattr(c, 'free_symbols', ()) or any(str(fs) == param for fs in getattr(c, 'free_symbols', ())):
        return None
    return c


def _match_wcr_map(state, sdfg, map_entry):
    """Describe the scatter a map scope performs, or None if it is not one."""
    m = map_entry.map
    if len(m.params) != 1 or not (m.range.ranges[0][2] == 1) is True:
        return None
    map_exit = state.exit_node(map_entry)
    body = [n for n in state.scope_subgraph(map_entry, include_entry=False, include_exit=False).nodes()]
    if len(body) != 1:
        return None
    body = body[0]
    exit_in, exit_out = state.in_edges(map_exit), state.out_edges(map_exit)
    if len(exit_in) != 1 or len(exit_out) != 1 or not isinstance(exit_out[0].dst, nodes.AccessNode):
        return None
    if any(e.dst_conn is None or not e.dst_conn.startswith('IN_') or not isinstance(e.src, nodes.AccessNode)
           for e in state.in_edges(map_entry)):
        return None   # dynamic map ranges, empty memlets or non-array producers
    inner_out, outer_out = exit_in[0], exit_out[0]
    bins_desc = sdfg.arrays[outer_out.data.data]
    if len(bins_desc.shape) != 1 or inner_out.src is not body:
        return None
    if not (outer_out.data.subset.ranges[0][2] == 1) is True:
        return None
    param = m.params[0]
    body_in = {e.dst_conn: e for e in state.in_edges(body)}
    if any(e.src is not map_entry for e in body_in.values()):
        return None

    def outer_of(inner_edge):
        conn = 'IN_' + inner_edge.src_conn[4:]
        return next(e for e in state.in_edges(map_entry) if e.dst_conn == conn)

    wcr, idx_conn, val_conn, const = None, None, None, None
    if isinstance(body, nodes.Tasklet):
        parsed = _parse_indexed_store(body)
        if parsed is None or inner_out.data.wcr is None:
            return None
        _, idx_conn, val_conn, const = parsed
        wcr = inner_out.data.wcr
    elif isinstance(body, nodes.NestedSDFG):
        # what ``bins[idx[i]] += v`` lowers to: init --(sym = idx connector)--> one tasklet writing w[sym] with WCR
        ns = body.sdfg
        states, isedges = list(ns.states()), list(ns.edges())
        if len(states) != 2 or len(isedges) != 1 or len(body.out_connectors) != 1:
            return None
        ise = isedges[0]
        if ise.src is not ns.start_state or ise.src.number_of_nodes() != 0 or len(ise.data.assignments) != 1:
            return None
        if not ise.data.is_unconditional():
            return None
        (sym, src), = ise.data.assignments.items()
        if src not in body.in_connectors:
            return None
        idx_conn = src
        inner = ise.dst
        tasklets = [n for n in inner.nodes() if isinstance(n, nodes.Tasklet)]
        if len(tasklets) != 1 or any(not isinstance(n, (nodes.Tasklet, nodes.AccessNode)) for n in inner.nodes()):
            return None
        parsed = _parse_plain_store(tasklets[0])
        if parsed is None:
            return None
        _, inner_val, const = parsed
        t_out = inner.out_edges(tasklets[0])
        if len(t_out) != 1 or t_out[0].data.wcr is None or not isinstance(t_out[0].dst, nodes.AccessNode):
            return None
        if t_out[0].dst.data != inner_out.src_conn or inner.out_degree(t_out[0].dst) != 0:
            return None
        (wb, we, _), = t_out[0].data.subset.ranges if len(t_out[0].data.subset.ranges) == 1 else ((None, None, None), )
        if wb is None or str(wb) != sym or str(we) != sym:
            return None
        if inner_val is not None:
            t_in = inner.in_edges(tasklets[0])
            if len(t_in) != 1 or not isinstance(t_in[0].src, nodes.AccessNode) or t_in[0].src.data not in body.in_connectors:
                return None
            val_conn = t_in[0].src.data
            if val_conn == idx_conn or inner.in_degree(t_in[0].src) != 0:
                return None
        if set(body.in_connectors) != ({idx_conn, val_conn} - {None}):
            return None
        # the index symbol must not be used anywhere else
        if sym in body.symbol_mapping:
            return None
        wcr = t_out[0].data.wcr
    else:
        return None

    idx_inner = body_in[idx_conn]
    idx_c = _affine_in_param(idx_inner.data, sdfg.arrays[idx_inner.data.data], param)
    if idx_c is None or sdfg.arrays[idx_inner.data.data].dtype not in dtypes.INTEGER_TYPES:
        return None
    val_inner, val_c = None, None
    if val_conn is not None:
        val_inner = body_in[val_conn]
        val_c = _affine_in_param(val_inner.data, sdfg.arrays[val_inner.data.data], param)
        if val_c is None:
            return None
    return dict(map_exit=map_exit, body=body, wcr=wcr, const=const, param=param,
                idx=(outer_of(idx_inner), idx_c), val=(outer_of(val_inner), val_c) if val_inner is not None else None,
                bins=outer_out)


@dace.properties.make_properties
class WCRMapToScatterReduce(xf.SingleStateTransformation):
    """Replace a map that scatters through a WCR memlet (``bins[idx[i]] += v[i]``,
    histograms) by a :class:`ScatterReduce` library node."""

    map_entry = xf.PatternNode(nodes.MapEntry)

    implementation = Property(dtype=str, default='B200', allow_none=True,
                              desc='implementation set on the new node (None: the library default, "pure")')

    @classmethod
    def expressions(cls):
        return [sdutil.node_path_graph(cls.map_entry)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        return _match_wcr_map(graph, sdfg, self.map_entry) is not None

    def apply(self, graph, sdfg):
        state, map_entry = graph, self.map_entry
        m = _match_wcr_map(state, sdfg, map_entry)
        lo, hi, _ = map_entry.map.range.ranges[0]
        node = ScatterReduce('scatter_' + map_entry.map.label, wcr=m['wcr'], const_value=m['const'],
                             has_values=m['val'] is not None)
        node.implementation = self.implementation
        node.debuginfo = map_entry.map.debuginfo
        state.add_node(node)

        def connect(which, conn):
            outer, c = m[which]
            memlet = Memlet(data=outer.data.data, subset=subsets.Range([(lo + c, hi + c, 1)]))
            state.add_edge(outer.src, outer.src_conn, node, conn, memlet)

        connect('idx', '_idx')
        if m['val'] is not None:
            connect('val', '_val')
        bins = m['bins']
        out_memlet = Memlet(data=bins.data.data, subset=copy.deepcopy(bins.data.subset), wcr=m['wcr'])
        state.add_edge(node, '_bins', bins.dst, bins.dst_conn, out_memlet)
        state.remove_node(m['body'])
        state.remove_node(m['map_exit'])
        state.remove_node(map_entry)
        return node


def use_b200_scatter(sdfg) -> int:
    """Rewrite every WCR scatter map of ``sdfg`` into a ScatterReduce node expanded by the sm_100a
    library; returns how many were rewritten."""
    return sdfg.apply_transformations_repeated(WCRMapToScatterReduce, options={'implementation': 'B200'})


dace.library.register_library(__name__, 'dace_b200')
