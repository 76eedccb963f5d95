"""DaCe plug-in: two sibling Reduce nodes (row max + row sum of the same array) -> one pass.

BASELINE config C3 ("row-wise max + sum (softmax-style) reductions over [65536, 8192] bf16 with fp32
accumulation") is, in a DaCe program, two ``Reduce`` library nodes that read the same array
(``mx[:] = np.max(A, axis=1); sm[:] = np.sum(A, axis=1)``, cf. reference tests/softmax_test.py:8-31).
Expanded one by one they read the input twice.  ``FuseRowMaxSum`` replaces the pair with a
``RowsMaxSum`` library node whose ``'B200'`` expansion calls ``dace_b200_reduce_rows_max_sum``:
(max, sum) pairs per row in a single pass over HBM.

Conditions (else the nodes stay as they are): both read the same access node with the same memlet,
both reduce exactly the last (contiguous) axis, input f32 / bf16 / f16 on the device, fp32 outputs
that are whole contiguous arrays, Sum identity 0, Max identity <= -3e38 (the frontend's
``min_value``).  The kernel seeds the maximum with -inf: the result differs from the reference's
only for a row made of -inf alone (reference: the identity; here: -inf).
"""
import dace
import dace.library
from dace import dtypes, nodes, properties
from dace.libraries.standard.nodes.reduce import Reduce
from dace.sdfg import utils as sdutil
from dace.symbolic import symstr
from dace.transformation import transformation as xf
from dace.transformation.transformation import ExpandTransformation

from .expansion import B200Reduce, _DTYPE_CODES
from .redtypes import ReductionType, detect_reduction_type, lower_custom

_IN_TYPES = [dace.float32, dace.float16] + ([dace.bfloat16] if hasattr(dace, 'bfloat16') else [])
_GPU = (dtypes.StorageType.GPU_Global, dtypes.StorageType.CPU_Pinned)


def _rows_view(memlet, desc):
    """(rows, n, row_stride) when the memlet is [rows..., n] with a contiguous last axis and leading axes
    that merge into one row index; else None."""
    sizes = memlet.subset.size()
    steps = [r[2] for r in memlet.subset.ranges]
    strides = [st * sp for st, sp in zip(desc.strides, steps)]
    dims = [(sz, st) for sz, st in zip(sizes, strides) if (sz == 1) is not True]
    if not dims or (dims[-1][1] == 1) is not True:
        return None
    n = dims[-1][0]
    lead = dims[:-1]
    if not lead:
        return 1, n, n
    for (s0, t0), (s1, t1) in zip(lead[:-1], lead[1:]):
        if (t0 == t1 * s1) is not True:
            return None
    rows = 1
    for s, _ in lead:
        rows = rows * s
    return rows, n, lead[-1][1]


class ExpandRowsMaxSumB200(ExpandTransformation):
    environments = [B200Reduce]

    @staticmethod
    def expansion(node, state, sdfg):
        node.validate(sdfg, state)
        ie = state.in_edges(node)[0]
        outs = {e.src_conn: e for e in state.out_edges(node)}
        in_desc = sdfg.arrays[ie.data.data]
        view = _rows_view(ie.data, in_desc)
        if view is None or in_desc.dtype not in _IN_TYPES:
            raise NotImplementedError('RowsMaxSum expansion "B200": %s of %s' % (ie.data, in_desc.dtype))
        for e in [ie] + list(outs.values()):
            if sdfg.arrays[e.data.data].storage not in _GPU:
                raise ValueError('RowsMaxSum expansion "B200" needs %s in GPU_Global or CPU_Pinned storage. There is no '
                                 'CPU fallback.' % e.data.data)
        rows, n, stride = view
        code = """
DACE_GPU_CHECK((cudaError_t)dace_b200_reduce_rows_max_sum({dt}, _in, _max, _sum, (size_t)({rows}), (size_t)({n}),
    (size_t)({stride}), (dace_b200_stream_t)__dace_current_stream));
""".format(dt=_DTYPE_CODES[in_desc.dtype], rows=symstr(rows, cpp_mode=True), n=symstr(n, cpp_mode=True),
           stride=symstr(stride, cpp_mode=True))
        return nodes.Tasklet('rows_max_sum_b200', {'_in': dace.pointer(in_desc.dtype)},
                             {'_max': dace.pointer(dace.float32), '_sum': dace.pointer(dace.float32)}, code,
                             language=dace.Language.CPP)


@dace.library.node
class RowsMaxSum(nodes.LibraryNode):
    """``_max[r] = max(_in[r, :])`` and ``_sum[r] = sum(_in[r, :])`` (fp32) in one pass."""

    implementations = {'B200': ExpandRowsMaxSumB200}
    default_implementation = 'B200'

    def __init__(self, name='rows_max_sum', **kwargs):
        super().__init__(name, inputs={'_in'}, outputs={'_max', '_sum'}, **kwargs)

    @staticmethod
    def from_json(json_obj, context=None):
        ret = RowsMaxSum('rows_max_sum')
        dace.serialize.set_properties_from_json(ret, json_obj, context=context)
        return ret

    def validate(self, sdfg, state):
        if len(state.in_edges(self)) != 1 or {e.src_conn for e in state.out_edges(self)} != {'_max', '_sum'}:
            raise ValueError('RowsMaxSum needs one _in input and the outputs _max and _sum')


def _kind(node: Reduce):
    rt = detect_reduction_type(node.wcr)
    if rt == ReductionType.Custom:
        rt = lower_custom(node.wcr)
    return rt


def _dense_f32_out(state, sdfg, node):
    from .comm_node import _full_contiguous
    oe = state.out_edges(node)[0]
    d = sdfg.arrays[oe.data.data]
    return d.dtype == dace.float32 and d.storage in _GPU and _full_contiguous(oe.data, d) and not oe.data.wcr


def _partner(state, sdfg, mx: Reduce):
    """The Sum Reduce that can share a pass with the Max Reduce ``mx``, or None."""
    if _kind(mx) != ReductionType.Max or mx.identity is None:
        return None
    try:
        if not float(mx.identity) <= -3e38:
            return None
    except (TypeError, ValueError):
        return None
    ie = state.in_edges(mx)[0]
    src = ie.src
    desc = sdfg.arrays[ie.data.data]
    if not isinstance(src, nodes.AccessNode) or desc.dtype not in _IN_TYPES or desc.storage not in _GPU:
        return None
    nd = len(ie.data.subset)
    if mx.axes is None or [a % nd for a in mx.axes] != [nd - 1] or _rows_view(ie.data, desc) is None:
        return None
    if not _dense_f32_out(state, sdfg, mx):
        return None
    for e in state.out_edges(src):
        other = e.dst
        if other is mx or not isinstance(other, Reduce) or e.data != ie.data:
            continue
        if _kind(other) != ReductionType.Sum or other.identity is None or other.axes is None:
            continue
        try:
            zero = float(other.identity) == 0.0
        except (TypeError, ValueError):
            zero = False
        if zero and [a % nd for a in other.axes] == [nd - 1] and _dense_f32_out(state, sdfg, other):
            return other
    return None


@properties.make_properties
class FuseRowMaxSum(xf.SingleStateTransformation):
    """Reduce(max, last axis) + Reduce(sum, last axis) over the same array -> RowsMaxSum."""

    max_node = xf.PatternNode(Reduce)

    @classmethod
    def expressions(cls):
        return [sdutil.node_path_graph(cls.max_node)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        return _partner(graph, sdfg, self.max_node) is not None

    def apply(self, graph, sdfg):
        mx = self.max_node
        sm = _partner(graph, sdfg, mx)
        ie = graph.in_edges(mx)[0]
        node = RowsMaxSum('rows_max_sum_' + ie.data.data)
        node.implementation = 'B200'
        graph.add_node(node)
        graph.add_edge(ie.src, ie.src_conn, node, '_in', dace.Memlet(data=ie.data.data, subset=ie.data.subset))
        for red, conn in ((mx, '_max'), (sm, '_sum')):
            oe = graph.out_edges(red)[0]
            graph.add_edge(node, conn, oe.dst, oe.dst_conn, dace.Memlet(data=oe.data.data, subset=oe.data.subset))
        graph.remove_nodes_from([mx, sm])
        return node


def fuse_row_max_sum(sdfg) -> int:
    """Apply :class:`FuseRowMaxSum` wherever it matches; returns how many pairs were fused."""
    return sdfg.apply_transformations_repeated(FuseRowMaxSum, validate=True)
