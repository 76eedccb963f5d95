"""DaCe plug-in: Map -> transient -> Reduce fused into ONE generated kernel (``FuseMapIntoReduce``).

``s[:] = np.sum(np.exp(A - mx[:, None]) ...)``-style programs are, in the SDFG, an elementwise Map that
writes a transient array followed by a ``Reduce`` library node that reads it back.  The reference's
tool for this is ``MapReduceFusion`` (reference dace/transformation/dataflow/mapreduce.py:18-146): the
Reduce node disappears into a WCR memlet on the map, which the GPU code generator resolves with one
atomic per element (reference dace/codegen/targets/cpu.py:965-1013).  This rewrite keeps the reduction a
tree instead: the tasklets of the map body become a *load functor* (C++ generated with the reference's own
``cppunparse``), and ``dace_b200::fused_rows`` from ``include/dace_b200_fused_reduce.cuh`` is instantiated
with it inside the program's .cu file - the elementwise work happens in registers on the way into the
reduction, the transient is never written or read, one pass over HBM.

The emitted pieces mirror the reference's CUB expansion (reference
dace/libraries/standard/nodes/reduce.py:574-614): a ``DACE_EXPORTED cudaError_t __b200_fused_<id>(...)``
host function in the CUDA global code, its declaration in the frame code, and a host C++ tasklet that
calls it under ``DACE_GPU_CHECK`` on ``__dace_current_stream``.

Matches (anything else is left untouched):
* a top-level map whose ranges are exactly the transient's shape, writing ``T[p0, ..., pk]``, nothing else;
* a body of Python tasklets chained through register scalars, reading single elements of device arrays;
* ``T`` used nowhere else; the Reduce folds the last axis, all axes, or all leading axes (column fold:
  ``dace_b200::fused_cols``) with a built-in operator;
* accumulator types f32 / f64 / i32 / i64 / u32 / u64; a dense output.
"""
import ast
import copy

import dace
import dace.library
from dace import data, dtypes, nodes, properties
from dace.codegen import cppunparse
from dace.libraries.standard.nodes.reduce import Reduce
from dace.sdfg import utils as sdutil
from dace.symbolic import symstr
from dace.transformation import transformation as xf

from .expansion import B200FusedReduce, B200Reduce, node_workspace
from .redtypes import ReductionType, detect_reduction_type, lower_custom

_ACC_TYPES = (dace.float32, dace.float64, dace.int32, dace.int64, dace.uint32, dace.uint64)
# 16-bit floats are produced / stored as they are but ACCUMULATED in fp32 (north_star: "bf16 accumulated in fp32")
_HALF_TYPES = tuple(t for t in (dace.float16, getattr(dace, 'bfloat16', None)) if t is not None)


def _acc_of(dtype):
    return dace.float32 if dtype in _HALF_TYPES else dtype
_OPS = {ReductionType.Min, ReductionType.Max, ReductionType.Sum, ReductionType.Product,
        ReductionType.Logical_And, ReductionType.Bitwise_And, ReductionType.Logical_Or,
        ReductionType.Bitwise_Or, ReductionType.Logical_Xor, ReductionType.Bitwise_Xor}
_INT_ONLY = {ReductionType.Bitwise_And, ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor}
_GPU = (dtypes.StorageType.GPU_Global, dtypes.StorageType.CPU_Pinned)


def _redtype(node: Reduce):
    rt = detect_reduction_type(node.wcr)
    if rt == ReductionType.Custom:
        rt = lower_custom(node.wcr)
    return rt


def _is_true(x) -> bool:
    return (x is True) or (x == True) is True   # noqa: E712  (sympy relational -> only literal True counts)


class _Match:
    """Everything ``apply`` needs, collected (and validated) once."""

    def __init__(self):
        self.entry = self.exit = self.trans = self.reduce = None
        self.tasklets = []          # topological order
        self.scalars = []           # access nodes between tasklets
        self.all_axes = False
        self.cols = False           # fold the leading axes, keep the last one
        self.redtype = None
        # what the code generator needs, independent of where the pattern came from
        self.shape = None           # iteration space = extents of the map
        self.acc_dtype = None       # accumulator type
        self.value_dtype = None     # type of the value the map body produces (16-bit floats widen to the accumulator)
        self.out_edge = None        # edge whose destination / memlet receives the result
        self.identity = None        # None: fold into the existing output
        self.uid = ''


def _collect_body(state, sdfg, entry, mexit):
    """(tasklets in topological order, scalar access nodes) of a map body made of Python tasklets chained
    through register scalars and reading single elements of device arrays; None if it is anything else."""
    scope = state.scope_subgraph(entry, include_entry=False, include_exit=False)
    tasklets, scalars = [], []
    for n in sdutil.dfs_topological_sort(scope, [x for x in scope.nodes() if scope.in_degree(x) == 0]):
        if isinstance(n, nodes.Tasklet):
            if n.language != dtypes.Language.Python or n.side_effects or len(n.out_connectors) != 1:
                return None
            tasklets.append(n)
        elif isinstance(n, nodes.AccessNode):
            d = sdfg.arrays[n.data]
            if not d.transient or not _is_true(d.total_size == 1) or state.in_degree(n) != 1:
                return None
            scalars.append(n)
        else:
            return None
    if not tasklets:
        return None
    for t in tasklets:
        for e in state.in_edges(t):
            if e.data.is_empty():
                continue
            if e.src is entry:
                d = sdfg.arrays[e.data.data]
                if (d.storage not in _GPU or e.data.wcr is not None or e.data.dynamic
                        or not _is_true(e.data.subset.num_elements() == 1) or isinstance(d.dtype, dtypes.struct)):
                    return None
            elif not isinstance(e.src, nodes.AccessNode):
                return None
        outs = state.out_edges(t)
        if len(outs) != 1 or not (outs[0].dst is mexit or isinstance(outs[0].dst, nodes.AccessNode)):
            return None
    for e in state.in_edges(entry):
        if e.data.is_empty() or not isinstance(e.src, nodes.AccessNode):
            return None
    return tasklets, scalars


def _match(state, sdfg, red: Reduce):
    rt = _redtype(red)
    if rt not in _OPS:
        return None
    ie, oe = state.in_edges(red)[0], state.out_edges(red)[0]
    tnode = ie.src
    if not isinstance(tnode, nodes.AccessNode):
        return None
    tdesc = sdfg.arrays[tnode.data]
    if (not tdesc.transient or isinstance(tdesc, data.View) or not isinstance(tdesc, data.Array)
            or tdesc.storage not in _GPU or tdesc.dtype not in _ACC_TYPES + _HALF_TYPES):
        return None
    if rt in _INT_ONLY and tdesc.dtype in (dace.float32, dace.float64) + _HALF_TYPES:
        return None
    if state.in_degree(tnode) != 1 or state.out_degree(tnode) != 1:
        return None
    uses = sum(1 for st in sdfg.states() for n in st.data_nodes() if n.data == tnode.data)
    if uses != 1 or any(tnode.data in str(e.data.free_symbols) for e in sdfg.all_interstate_edges()):
        return None
    mexit = state.in_edges(tnode)[0].src
    if not isinstance(mexit, nodes.MapExit):
        return None
    entry = state.entry_node(mexit)
    if state.entry_node(entry) is not None:
        return None
    nd = len(tdesc.shape)
    params = list(entry.map.params)
    if len(params) != nd or state.in_degree(mexit) != 1 or state.out_degree(mexit) != 1:
        return None
    for (b, e, s), sh in zip(entry.map.range.ranges, tdesc.shape):
        if not (_is_true(b == 0) and _is_true(s == 1) and _is_true(e + 1 == sh)):
            return None
    # packed C strides of the transient (the functor never touches it, but the row / column split relies on its shape)
    w = state.in_edges(mexit)[0]
    if w.data.wcr is not None or w.data.data != tnode.data:
        return None
    for p, (b, e, s) in zip(params, w.data.subset.ranges):
        if str(b) != p or str(e) != p:
            return None
    # the Reduce reads all of T and folds the last axis or everything
    if not all(_is_true(sz == sh) for sz, sh in zip(ie.data.subset.size(), tdesc.shape)):
        return None
    axes = sorted(a % nd for a in red.axes) if red.axes is not None else list(range(nd))
    cols = False
    if axes == list(range(nd)):
        all_axes = True
    elif axes == [nd - 1]:
        all_axes = False
    elif nd >= 2 and axes == list(range(nd - 1)):
        all_axes, cols = False, True       # every leading axis folded, the last (contiguous) one kept
    else:
        return None
    odesc = sdfg.arrays[oe.data.data]
    if odesc.storage not in _GPU or odesc.dtype not in _ACC_TYPES + _HALF_TYPES or oe.data.wcr is not None:
        return None
    from .comm_node import _full_contiguous
    if not all_axes and not _full_contiguous(oe.data, odesc):
        return None
    if all_axes and not _is_true(oe.data.subset.num_elements() == 1):
        return None
    body = _collect_body(state, sdfg, entry, mexit)
    if body is None:
        return None
    tasklets, scalars = body
    m = _Match()
    m.entry, m.exit, m.trans, m.reduce = entry, mexit, tnode, red
    m.tasklets, m.scalars, m.all_axes, m.redtype = tasklets, scalars, all_axes, rt
    m.cols = cols
    m.shape, m.acc_dtype, m.out_edge, m.identity = list(tdesc.shape), _acc_of(tdesc.dtype), oe, red.identity
    m.value_dtype = tdesc.dtype
    m.uid = '%s_%d_%d' % (sdfg.name, sdfg.node_id(state), state.node_id(red))
    return m


def _unparse(code: str, defined: dict) -> str:
    # connectors are declared by the functor itself: tell the unparser, or it emits `auto __out = ...`
    return cppunparse.cppunparse(ast.parse(code), expr_semicolon=True, defined_symbols=dict(defined))


def _generate(state, sdfg, m: _Match):
    """(cuda global code, frame declaration, tasklet call code, [input array names])."""
    entry = m.entry
    oe = m.out_edge
    odesc = sdfg.arrays[oe.data.data]
    acc, out_t = m.acc_dtype.ctype, odesc.dtype.ctype
    shape = list(m.shape)
    nd = len(shape)
    params = list(entry.map.params)
    ident = m.uid

    packed = [1] * nd
    for d in range(nd - 2, -1, -1):
        packed[d] = packed[d + 1] * shape[d + 1]

    arrays, symbols = [], set()
    body, need_params = [], False
    vbody, vec_arrays = [], []          # the same tasklets with flat reads taken from vector registers

    def note_symbols(expr):
        for s in getattr(expr, 'free_symbols', []):
            if str(s) not in params:
                symbols.add(str(s))

    for t in m.tasklets:
        out_edge = state.out_edges(t)[0]
        out_conn = next(iter(t.out_connectors))
        if out_edge.dst is m.exit:
            out_name, out_ctype, final = '__b200_value', m.value_dtype.ctype, True
        else:
            out_name, out_ctype, final = out_edge.dst.data, sdfg.arrays[out_edge.dst.data].dtype.ctype, False
        # map parameters / program symbols used inside the tasklet code itself (`tmp[i, j] = A[i, j] * i / N`)
        for fs in map(str, t.free_symbols):
            if fs in params:
                need_params = True
            elif fs in sdfg.symbols:
                symbols.add(fs)
        lines, vlines = [], []
        defined = {out_conn: m.value_dtype if final else sdfg.arrays[out_edge.dst.data].dtype}
        for e in state.in_edges(t):
            if e.data.is_empty():
                continue
            defined[e.dst_conn] = sdfg.arrays[e.data.data].dtype if e.src is entry else sdfg.arrays[e.src.data].dtype
            if e.src is entry:
                name = e.data.data
                d = sdfg.arrays[name]
                if name not in arrays:
                    arrays.append(name)
                idx = e.data.subset.min_element()
                flat = (len(idx) == nd and all(str(i) == p for i, p in zip(idx, params))
                        and all(_is_true(a == b) for a, b in zip(d.strides, packed)))
                if flat:
                    off = '__flat'
                else:
                    expr = sum(i * s for i, s in zip(idx, d.strides))
                    note_symbols(expr)
                    off = symstr(expr, cpp_mode=True)
                    need_params = True
                lines.append('const %s %s = %s[%s];' % (d.dtype.ctype, e.dst_conn, name, off))
                if flat and d.dtype.bytes in (1, 2, 4, 8):
                    if name not in vec_arrays:
                        vec_arrays.append(name)
                    vlines.append('const %s %s = __b200_v_%s[__k];' % (d.dtype.ctype, e.dst_conn, name))
                else:
                    vlines.append(lines[-1])
            else:
                lines.append('const %s %s = %s;' % (sdfg.arrays[e.src.data].dtype.ctype, e.dst_conn, e.src.data))
                vlines.append(lines[-1])
        tail = ['%s %s;' % (out_ctype, out_conn), _unparse(t.code.as_string, defined), '%s = %s;' % (out_name, out_conn)]
        decl = '' if final else '%s %s;\n        ' % (out_ctype, out_name)
        body.append(decl + '{ ' + '\n          '.join(lines + tail) + ' }')
        vbody.append(decl + '{ ' + '\n          '.join(vlines + tail) + ' }')

    # (row, col) -> map parameters
    rows_expr = 1
    if m.all_axes:
        n_expr = 1
        for s in shape:
            n_expr = n_expr * s
        decomp, src, flat_expr = list(range(nd)), '__col', '__col'
    else:
        n_expr = shape[-1]
        for s in shape[:-1]:
            rows_expr = rows_expr * s
        decomp, src = list(range(nd - 1)), '__row'
        flat_expr = '__row * (size_t)(%s) + __col' % symstr(n_expr, cpp_mode=True)
    note_symbols(n_expr)
    note_symbols(rows_expr)
    pcode = []
    if need_params:
        if not m.all_axes:
            pcode.append('const long long %s = (long long)__col;' % params[-1])
        if len(decomp) == 1:
            pcode.append('const long long %s = (long long)%s;' % (params[decomp[0]], src))
        elif decomp:
            pcode.append('size_t __rem = %s;' % src)
            for d in reversed(decomp[1:]):
                note_symbols(shape[d])
                ext = '(size_t)(%s)' % symstr(shape[d], cpp_mode=True)
                pcode.append('const long long %s = (long long)(__rem %% %s); __rem /= %s;' % (params[d], ext, ext))
            pcode.append('const long long %s = (long long)__rem;' % params[decomp[0]])

    syms = sorted(symbols)
    sym_types = {s: (sdfg.symbols[s].ctype if s in sdfg.symbols else 'long long') for s in syms}
    members = ['const %s* __restrict__ %s;' % (sdfg.arrays[a].dtype.ctype, a) for a in arrays] + \
              ['%s %s;' % (sym_types[s], s) for s in syms]
    sig = ['const %s* %s' % (sdfg.arrays[a].dtype.ctype, a) for a in arrays] + ['%s* __out' % out_t] + \
          ['%s %s' % (sym_types[s], s) for s in syms] + ['void* __workspace', 'size_t __workspace_bytes', 'cudaStream_t __stream']
    if m.identity is None:
        ident_decl, ident_arg = '', '(const %s*)nullptr' % acc
    else:
        ident_decl = 'const %s __identity = %s;' % (acc, symstr(m.identity, cpp_mode=True))
        ident_arg = '&__identity'
    # Vector interface (include/dace_b200_fused_reduce.cuh, VECTOR LOADS): operands read at the map's own flat index
    # are fetched V elements at a time (32 bytes of the widest operand, at most 8), everything else stays as it is.
    vec_code, vec_arg = '', ''
    if vec_arrays:
        kvec = min(8, 32 // max(sdfg.arrays[a].dtype.bytes for a in vec_arrays))
        loads = ['%s __b200_v_%s[__V]; dace_b200::stream_load<%s, __V>(%s + __flat0, __b200_v_%s);' %
                 (sdfg.arrays[a].dtype.ctype, a, sdfg.arrays[a].dtype.ctype, a, a) for a in vec_arrays]
        vec_code = """
    static constexpr int kVec = {kvec};
    using value_type = {acc};
    template <int __V>
    __device__ __forceinline__ void vec(size_t __row, size_t __col0, {acc} (&__b200_out)[__V]) const {{
        const size_t __flat0 = {flat0};
        {loads}
        #pragma unroll
        for (int __k = 0; __k < __V; ++__k) {{
            const size_t __col = __col0 + __k; (void)__col;
            const size_t __flat = __flat0 + __k; (void)__flat;
            {pcode}
            {val} __b200_value;
            {body}
            __b200_out[__k] = ({acc})__b200_value;
        }}
    }}""".format(kvec=kvec, acc=acc, flat0=flat_expr.replace('__col', '__col0'), loads='\n        '.join(loads),
                 pcode='\n            '.join(pcode), val=m.value_dtype.ctype, body='\n            '.join(vbody))
        aligned = ' && '.join('((size_t)%s %% (__v * sizeof(%s))) == 0' % (a, sdfg.arrays[a].dtype.ctype) for a in vec_arrays)
        vec_arg = ', __b200_max_vec'
        vec_pick = ('int __b200_max_vec = 1;\n    for (int __v = %d; __v >= 2; __v /= 2)\n'
                    '        if (((size_t)(%s)) %% __v == 0 && %s) { __b200_max_vec = __v; break; }' %
                    (kvec, symstr(n_expr, cpp_mode=True), aligned))
    else:
        vec_pick = ''
    cuda_code = """
struct __b200_load_{id} {{
    {members}
    __device__ __forceinline__ {acc} operator()(size_t __row, size_t __col) const {{
        const size_t __flat = {flat}; (void)__flat;
        {pcode}
        {val} __b200_value;
        {body}
        return ({acc})__b200_value;
    }}{vec_code}
}};
DACE_EXPORTED cudaError_t __b200_fused_{id}({sig});
cudaError_t __b200_fused_{id}({sig}) {{
    __b200_load_{id} __load{{{init}}};
    {ident_decl}
    {vec_pick}
    return dace_b200::{fn}<{red}, {acc}, {out_t}>(__load, __out, (ptrdiff_t)1, (size_t)({rows}), (size_t)({n}), {ident_arg},
                                                      __workspace, __workspace_bytes, __stream{vec_arg});
}}
""".format(id=ident, members='\n    '.join(members), acc=acc, out_t=out_t, flat=flat_expr, pcode='\n        '.join(pcode),
           val=m.value_dtype.ctype, body='\n        '.join(body), sig=', '.join(sig), init=', '.join(arrays + syms), ident_decl=ident_decl,
           fn='fused_cols' if m.cols else 'fused_rows', red=int(m.redtype), rows=symstr(rows_expr, cpp_mode=True), n=symstr(n_expr, cpp_mode=True), ident_arg=ident_arg,
           vec_code=vec_code, vec_pick=vec_pick, vec_arg=vec_arg)
    frame_decl = 'DACE_EXPORTED cudaError_t __b200_fused_%s(%s);\n' % (ident, ', '.join(sig))
    call = ('DACE_GPU_CHECK(__b200_fused_%s(%s));' %
            (ident, ', '.join(['_in_' + a for a in arrays] + ['_out'] + syms +
                              ['@B200_WS@', '@B200_WS_BYTES@', '__dace_current_stream'])))
    return cuda_code, frame_decl, call, arrays


def _apply(graph, sdfg, m: _Match):
    """Replace the matched subgraph by one host tasklet that launches the generated kernel."""
    cuda_code, frame_decl, call, arrays = _generate(graph, sdfg, m)
    ws, ws_bytes = node_workspace(sdfg, graph, m.entry)      # tickets + split partials of this kernel only
    call = call.replace('@B200_WS@', ws).replace('@B200_WS_BYTES@', ws_bytes)
    entry, oe = m.entry, m.out_edge
    out_desc = sdfg.arrays[oe.data.data]
    in_conns = {'_in_' + a: dace.pointer(sdfg.arrays[a].dtype) for a in arrays}
    tasklet = nodes.Tasklet('fused_map_reduce_b200', in_conns, {'_out': dace.pointer(out_desc.dtype)}, call,
                            language=dace.Language.CPP)
    tasklet.environments = {B200FusedReduce.full_class_path(), B200Reduce.full_class_path()}
    graph.add_node(tasklet)
    seen = set()
    for e in graph.in_edges(entry):
        a = e.data.data
        if a in seen or a not in arrays:
            continue
        seen.add(a)
        graph.add_edge(e.src, e.src_conn, tasklet, '_in_' + a, dace.Memlet.from_array(a, sdfg.arrays[a]))
    out_memlet = copy.deepcopy(oe.data)
    out_memlet.wcr = None          # the kernel folds into the output itself
    graph.add_edge(tasklet, '_out', oe.dst, oe.dst_conn, out_memlet)
    inner = [n.data for n in m.scalars]
    graph.remove_nodes_from(list(graph.scope_subgraph(entry).nodes()) + [n for n in (m.trans, m.reduce) if n is not None])
    for name in inner + ([m.trans.data] if m.trans is not None else []):
        if not any(n.data == name for st in sdfg.states() for n in st.data_nodes()):
            sdfg.remove_data(name, validate=False)
    sdfg.append_global_code(cuda_code, 'cuda')
    sdfg.append_global_code(frame_decl)
    return tasklet


def _match_wcr(state, sdfg, entry: nodes.MapEntry):
    """A reduction written as a WCR map - ``for i, j in dace.map[...]: s[i] += f(A[i, j])`` after the frontend's
    AugAssignToWCR, an explicit ``o >> s(1, lambda x, y: x + y)[i]`` tasklet output, or what the reference's
    MapReduceFusion leaves behind: the output index uses only some of the map parameters, the others are folded
    with the WCR operator.  Supported index forms: the leading parameters in order (row fold), only the last
    parameter (column fold), a constant (everything folded)."""
    if state.entry_node(entry) is not None:
        return None
    mexit = state.exit_node(entry)
    if state.in_degree(mexit) != 1 or state.out_degree(mexit) != 1:
        return None
    w, outer = state.in_edges(mexit)[0], state.out_edges(mexit)[0]
    if w.data.wcr is None or outer.data.wcr is None or w.data.dynamic or not isinstance(outer.dst, nodes.AccessNode):
        return None
    rt = detect_reduction_type(w.data.wcr)
    if rt == ReductionType.Custom:
        rt = lower_custom(w.data.wcr)
    odesc = sdfg.arrays[w.data.data]
    if (rt not in _OPS or odesc.storage not in _GPU or odesc.dtype not in _ACC_TYPES + _HALF_TYPES
            or isinstance(odesc, data.View) or (rt in _INT_ONLY and odesc.dtype in (dace.float32, dace.float64) + _HALF_TYPES)):
        return None
    params = list(entry.map.params)
    nd = len(params)
    shape = []
    for b, e, st in entry.map.range.ranges:
        if not (_is_true(b == 0) and _is_true(st == 1)):
            return None
        shape.append(e + 1)
    idx = []
    for b, e, _ in w.data.subset.ranges:
        if str(b) != str(e):
            return None
        idx.append(str(b))
    from .comm_node import _full_contiguous
    all_axes = cols = False
    uses_params = any(set(map(str, getattr(b, 'free_symbols', []))) & set(params) for b, _, _ in w.data.subset.ranges)
    if not uses_params:
        all_axes = True                                       # s[const] (+)= ...
        if not _is_true(outer.data.subset.num_elements() == 1):
            return None
    elif nd >= 2 and idx == params[:-1]:
        pass                                                  # rows: s[i...] (+)= f(..., j)
    elif nd >= 2 and idx == [params[-1]]:
        cols = True                                           # cols: s[j] (+)= f(i..., j)
    else:
        return None
    if not all_axes:
        kept = shape[:-1] if not cols else shape[-1:]
        if len(odesc.shape) != len(kept) or not all(_is_true(a == b) for a, b in zip(odesc.shape, kept)):
            return None
        if not _full_contiguous(outer.data, odesc):
            return None
    body = _collect_body(state, sdfg, entry, mexit)
    if body is None:
        return None
    tasklets, scalars = body
    if any(e.data.data == w.data.data for t in tasklets for e in state.in_edges(t) if e.src is entry):
        return None                                           # the map also reads what it accumulates into
    m = _Match()
    m.entry, m.exit, m.trans, m.reduce = entry, mexit, None, None
    m.tasklets, m.scalars, m.all_axes, m.cols, m.redtype = tasklets, scalars, all_axes, cols, rt
    m.shape, m.acc_dtype, m.out_edge, m.identity = shape, _acc_of(odesc.dtype), outer, None   # WCR: fold into the output
    m.value_dtype = odesc.dtype
    m.uid = '%s_%d_%d' % (sdfg.name, sdfg.node_id(state), state.node_id(entry))
    return m


@properties.make_properties
class FuseMapIntoReduce(xf.SingleStateTransformation):
    """Map -> transient -> Reduce  =>  one host tasklet that launches ``dace_b200::fused_rows`` /
    ``fused_cols`` with a load functor generated from the map body."""

    reduce = xf.PatternNode(Reduce)

    @classmethod
    def expressions(cls):
        return [sdutil.node_path_graph(cls.reduce)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        return _match(graph, sdfg, self.reduce) is not None

    def apply(self, graph, sdfg):
        return _apply(graph, sdfg, _match(graph, sdfg, self.reduce))


@properties.make_properties
class WCRMapToFusedReduce(xf.SingleStateTransformation):
    """A reduction written as a WCR map  =>  one ``dace_b200::fused_rows`` / ``fused_cols`` launch (a tree
    reduction folded into the existing output) instead of one atomic per element - the WCR lowering hook
    for maps without indirection (indirect WCR maps go to ``ScatterReduce``, dace_b200/scatter_node.py)."""

    map_entry = xf.PatternNode(nodes.MapEntry)

    @classmethod
    def expressions(cls):
        return [sdutil.node_path_graph(cls.map_entry)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        return _match_wcr(graph, sdfg, self.map_entry) is not None

    def apply(self, graph, sdfg):
        return _apply(graph, sdfg, _match_wcr(graph, sdfg, self.map_entry))


def fuse_maps_into_reductions(sdfg) -> int:
    """Apply :class:`FuseMapIntoReduce` wherever it matches (call after ``apply_gpu_transformations``);
    returns the number of fused pairs."""
    return sdfg.apply_transformations_repeated(FuseMapIntoReduce, validate=True)


def lower_wcr_maps(sdfg) -> int:
    """Apply :class:`WCRMapToFusedReduce` wherever it matches; returns the number of rewritten maps."""
    return sdfg.apply_transformations_repeated(WCRMapToFusedReduce, validate=True)
