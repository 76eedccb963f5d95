"""DaCe plug-in, multi-GPU part: the ``AllReduceCombine`` library node and the sharding
transformation that inserts it after reductions over a sharded axis.

Process model is DaCe's: one process per GPU, every process runs the same SDFG on its own shard
(reference dace/library.py:206-219; the reference's distributed programs are written the same way
with the MPI library, reference dace/libraries/mpi/nodes/allreduce.py:9-60).  What is added:

* ``B200Comm`` - environment: a communicator handle in the SDFG state struct, taken at init from
  ``dace_b200_comm_default()`` (NCCL bound at run time + the one-shot NVLink peer-memory kernel;
  include/dace_b200_reduce.h, dace_b200/csrc/comm.cu).  Mirrors the MPI environment
  (reference dace/libraries/mpi/environments/mpi.py:6-22).
* ``AllReduceCombine`` - library node, ``_inbuf -> _outbuf`` (normally the same array: in place)
  with a ``wcr`` lambda; expansion ``'B200'`` emits one ``dace_b200_comm_allreduce`` on the
  program's stream; expansion ``'pure'`` is the single-process meaning (a copy).
* :func:`shard_reductions` / ``ShardedReduceCombine`` - the transformation north_star asks for:
  given which program arrays are per-rank shards and along which axis, every ``Reduce`` node
  that folds over the sharded axis gets its partial combined (all-reduce right after the local
  kernel, same stream); ``ScatterReduce`` nodes over sharded keys scatter into a neutral-initialised
  private copy that is combined and folded into the bins once.  Reductions that keep the sharded
  axis produce disjoint slices and are left alone (no collective).
"""
import copy
from typing import Dict

import dace
import dace.library
from dace import dtypes, nodes, properties
from dace.frontend.operations import detect_reduction_type
from dace.libraries.standard.nodes.reduce import Reduce
from dace.properties import LambdaProperty
from dace.symbolic import symstr
from dace.transformation import transformation as xf
from dace.transformation.transformation import ExpandTransformation

from .expansion import B200Reduce, _DTYPE_CODES
from .redtypes import ReductionType, lower_custom

_COMBINABLE = {ReductionType.Min, ReductionType.Max, ReductionType.Sum, ReductionType.Product,
               ReductionType.Logical_And, ReductionType.Bitwise_And, ReductionType.Logical_Or,
               ReductionType.Bitwise_Or, ReductionType.Logical_Xor, ReductionType.Bitwise_Xor}


@dace.library.environment
class B200Comm:
    """Communicator of the process group in the state struct (collective init: every rank's
    ``__dace_init_*`` runs it)."""
    cmake_minimum_version = None
    cmake_packages = ["CUDA"]
    cmake_variables = {}
    cmake_includes = []
    cmake_libraries = []
    cmake_compile_flags = []
    cmake_link_flags = []
    cmake_files = []
    headers = {'frame': [], 'cuda': []}
    state_fields = ["void *dace_b200_comm = nullptr;"]
    init_code = """
    if (dace_b200_comm_default(&__state->dace_b200_comm) != 0) {
        __result = 1;
    }
"""
    finalize_code = ""      # the process-wide communicator outlives a single program
    dependencies = [B200Reduce]


def _redtype_of(wcr: str) -> ReductionType:
    rt = ReductionType[str(detect_reduction_type(wcr)).split('.')[-1]]
    if rt == ReductionType.Custom:
        rt = lower_custom(wcr)
    if rt not in _COMBINABLE:
        raise NotImplementedError('AllReduceCombine: reduction type %s has no collective' % rt.name)
    return rt


def neutral_element(rt: ReductionType, dtype: dace.typeclass):
    """The value ``e`` with ``op(e, x) == x`` (what every rank's partial is seeded with)."""
    if rt in (ReductionType.Sum, ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor, ReductionType.Logical_Or,
              ReductionType.Logical_Xor):
        return 0
    if rt in (ReductionType.Product, ReductionType.Logical_And):
        return 1
    if rt == ReductionType.Min:
        return dtypes.max_value(dtype)
    if rt == ReductionType.Max:
        return dtypes.min_value(dtype)
    if rt == ReductionType.Bitwise_And:
        return -1 if dtype in (dace.int8, dace.int16, dace.int32, dace.int64) else dtypes.max_value(dtype)
    raise NotImplementedError(rt)


def _full_contiguous(memlet, desc) -> bool:
    """True when the memlet covers a whole C-contiguous array (what one collective call can move)."""
    if any((b != 0) is True or (s != 1) is True for b, _, s in memlet.subset.ranges):
        return False
    if any((sz != sh) is True or (sz != sh) is not False for sz, sh in zip(memlet.subset.size(), desc.shape)):
        return False
    expect = 1
    for sh, st in zip(reversed(desc.shape), reversed(desc.strides)):
        if (sh != 1) is not False and ((st != expect) is True or (st != expect) is not False):
            return False
        expect = expect * sh
    return True


class ExpandAllReducePure(ExpandTransformation):
    """Single-process meaning of the node: the partial IS the result."""
    environments = []

    @staticmethod
    def expansion(node, state, sdfg):
        ie, oe = state.in_edges(node)[0], state.out_edges(node)[0]
        desc = sdfg.arrays[ie.data.data]
        n = symstr(ie.data.subset.num_elements(), cpp_mode=True)
        code = 'if (_inbuf != _outbuf) { for (size_t __i = 0; __i < (size_t)(%s); ++__i) _outbuf[__i] = _inbuf[__i]; }' % n
        return nodes.Tasklet('allreduce_combine_pure', {'_inbuf': dace.pointer(desc.dtype)},
                             {'_outbuf': dace.pointer(desc.dtype)}, code, language=dace.Language.CPP)


class ExpandAllReduceB200(ExpandTransformation):
    """AllReduceCombine -> ``dace_b200_comm_allreduce`` on the program's stream."""
    environments = [B200Comm]

    @staticmethod
    def expansion(node, state, sdfg):
        node.validate(sdfg, state)
        ie, oe = state.in_edges(node)[0], state.out_edges(node)[0]
        in_desc, out_desc = sdfg.arrays[ie.data.data], sdfg.arrays[oe.data.data]
        for d, e in ((in_desc, ie), (out_desc, oe)):
            if d.storage != dtypes.StorageType.GPU_Global:
                raise ValueError('AllReduceCombine expansion "B200" needs %s in GPU_Global storage (got %s). There is '
                                 'no CPU fallback.' % (e.data.data, d.storage))
            if not _full_contiguous(e.data, d):
                raise NotImplementedError('AllReduceCombine expansion "B200": %s must cover a whole contiguous array' %
                                          e.data)
        if in_desc.dtype != out_desc.dtype or in_desc.dtype not in _DTYPE_CODES:
            raise NotImplementedError('AllReduceCombine expansion "B200": dtype %s -> %s' % (in_desc.dtype, out_desc.dtype))
        rt = _redtype_of(node.wcr)
        n = symstr(ie.data.subset.num_elements(), cpp_mode=True)
        code = """
if (_inbuf != _outbuf) {{
    DACE_GPU_CHECK(cudaMemcpyAsync(_outbuf, _inbuf, (size_t)({n}) * sizeof({ctype}), cudaMemcpyDeviceToDevice,
                                   __dace_current_stream));
}}
DACE_GPU_CHECK((cudaError_t)dace_b200_comm_allreduce(__state->dace_b200_comm, {red}, {dt}, _outbuf, (size_t)({n}),
    (dace_b200_stream_t)__dace_current_stream));
""".format(n=n, ctype=in_desc.dtype.ctype, red=int(rt), dt=_DTYPE_CODES[in_desc.dtype])
        return nodes.Tasklet('allreduce_combine_b200', {'_inbuf': dace.pointer(in_desc.dtype)},
                             {'_outbuf': dace.pointer(out_desc.dtype)}, code, language=dace.Language.CPP)


@dace.library.node
class AllReduceCombine(nodes.LibraryNode):
    """``_outbuf = all-reduce over the ranks of _inbuf`` with the operator ``wcr``."""

    implementations = {'pure': ExpandAllReducePure, 'B200': ExpandAllReduceB200}
    default_implementation = 'B200'

    wcr = LambdaProperty(default='lambda a, b: a + b')

    def __init__(self, name='allreduce_combine', wcr='lambda a, b: a + b', **kwargs):
        super().__init__(name, inputs={'_inbuf'}, outputs={'_outbuf'}, **kwargs)
        self.wcr = wcr

    @staticmethod
    def from_json(json_obj, context=None):
        ret = AllReduceCombine('allreduce_combine')
        dace.serialize.set_properties_from_json(ret, json_obj, context=context)
        return ret

    def __str__(self):
        return 'AllReduceCombine (%s)' % str(detect_reduction_type(self.wcr)).split('.')[-1]

    def validate(self, sdfg, state):
        ins, outs = state.in_edges(self), state.out_edges(self)
        if len(ins) != 1 or len(outs) != 1 or ins[0].dst_conn != '_inbuf' or outs[0].src_conn != '_outbuf':
            raise ValueError('AllReduceCombine needs exactly one _inbuf input and one _outbuf output')
        if (ins[0].data.subset.num_elements() != outs[0].data.subset.num_elements()) is True:
            raise ValueError('AllReduceCombine: input and output sizes differ')


# ----------------------------------------------------------------------------------------------
# Sharding transformation
# ----------------------------------------------------------------------------------------------

def _map_schedule(desc):
    return dtypes.ScheduleType.GPU_Device if desc.storage == dtypes.StorageType.GPU_Global else dtypes.ScheduleType.Default


def _block256(map_entry, desc):
    if desc.storage == dtypes.StorageType.GPU_Global:
        map_entry.map.gpu_block_size = [256, 1, 1]


def _wcr_cpp(rt: ReductionType, ctype: str) -> str:
    return 'dace::_wcr_fixed<dace::ReductionType::%s, %s>()' % (rt.name, ctype)


def _through_views(sdfg, state, acc):
    """The frontend often lets a Reduce write through a View of the real array (``B[:] = np.sum(..)``):
    follow it to the access node of the array itself, or None when the view is a proper sub-range."""
    while isinstance(sdfg.arrays[acc.data], dace.data.View):
        outs = state.out_edges(acc)
        if len(outs) != 1 or not isinstance(outs[0].dst, nodes.AccessNode):
            return None
        nxt = outs[0].dst
        m = outs[0].data
        other = m.subset if m.data == nxt.data else m.other_subset
        if other is None:
            return None
        probe = dace.Memlet(data=nxt.data, subset=other)
        if not _full_contiguous(probe, sdfg.arrays[nxt.data]):
            return None
        acc = nxt
    return acc


def _append_combine(sdfg, state, acc, wcr: str):
    """``acc`` (access node just written by a local reduction) -> AllReduceCombine -> new access
    node of the same array, which takes over ``acc``'s readers.  Returns the new access node."""
    desc = sdfg.arrays[acc.data]
    comb = AllReduceCombine('combine_' + acc.data, wcr=wcr)
    comb.implementation = 'B200'
    state.add_node(comb)
    after = state.add_access(acc.data)
    for e in list(state.out_edges(acc)):
        state.add_edge(after, e.src_conn, e.dst, e.dst_conn, e.data)
        state.remove_edge(e)
    full = dace.Memlet.from_array(acc.data, desc)
    state.add_edge(acc, None, comb, '_inbuf', copy.deepcopy(full))
    state.add_edge(comb, '_outbuf', after, None, copy.deepcopy(full))
    return after


def _fold_scalar_after(sdfg, state, acc, rt: ReductionType, value):
    """acc[i] = op(value, acc[i]) as an ordinary GPU map (only for identities that are not neutral)."""
    desc = sdfg.arrays[acc.data]
    rng = {'__i%d' % i: '0:%s' % s for i, s in enumerate(desc.shape)}
    idx = ', '.join(rng.keys())
    after = state.add_access(acc.data)
    for e in list(state.out_edges(acc)):
        state.add_edge(after, e.src_conn, e.dst, e.dst_conn, e.data)
        state.remove_edge(e)
    code = '__o = %s(%s(%s), __v);' % (_wcr_cpp(rt, desc.dtype.ctype), desc.dtype.ctype, symstr(value, cpp_mode=True))
    _, me, _ = state.add_mapped_tasklet('fold_identity_' + acc.data, rng, {'__v': dace.Memlet('%s[%s]' % (acc.data, idx))}, code,
                             {'__o': dace.Memlet('%s[%s]' % (acc.data, idx))}, schedule=_map_schedule(desc),
                             language=dace.Language.CPP, input_nodes={acc.data: acc}, output_nodes={acc.data: after},
                             external_edges=True)
    _block256(me, desc)
    return after


def propagate_sharding(sdfg, sharded: Dict[str, int]) -> Dict[str, int]:
    """Which other arrays are shards, and along which axis:
    * whole-array copies of a sharded array (the host -> device copies of ``apply_gpu_transformations``);
    * outputs of elementwise maps: a map that reads ``X[..., p, ...]`` with parameter ``p`` on X's sharded
      axis and writes ``T[..., p, ...]`` makes T a shard along the axis indexed by ``p``."""
    out = dict(sharded)
    changed = True
    while changed:
        changed = False
        for state in sdfg.states():
            for e in state.edges():
                if isinstance(e.src, nodes.AccessNode) and isinstance(e.dst, nodes.AccessNode):
                    a, b = e.src.data, e.dst.data
                    if a in out and b not in out and len(sdfg.arrays[a].shape) == len(sdfg.arrays[b].shape):
                        out[b] = out[a]
                        changed = True
            for entry in state.nodes():
                if not isinstance(entry, nodes.MapEntry) or state.entry_node(entry) is not None:
                    continue
                params = set(entry.map.params)
                shard_params = set()
                for e in state.out_edges(entry):
                    name = e.data.data
                    if name in out and e.data.subset is not None and out[name] < len(e.data.subset):
                        b, en, _ = e.data.subset.ranges[out[name]]
                        if str(b) in params and str(b) == str(en):
                            shard_params.add(str(b))
                if not shard_params:
                    continue
                for e in state.in_edges(state.exit_node(entry)):
                    name = e.data.data
                    if name is None or name in out or e.data.subset is None:
                        continue
                    for d, (b, en, _) in enumerate(e.data.subset.ranges):
                        if str(b) in shard_params and str(b) == str(en):
                            out[name] = d
                            changed = True
                            break
    return out


def _reduce_needs_combine(node: Reduce, state, sharded: Dict[str, int]) -> bool:
    ie = state.in_edges(node)[0]
    if ie.data.data not in sharded:
        return False
    ndim = len(ie.data.subset)
    axes = list(range(ndim)) if node.axes is None else [a % ndim for a in node.axes]
    return sharded[ie.data.data] % ndim in axes


@properties.make_properties
class ShardedReduceCombine(xf.SingleStateTransformation):
    """Reduce over a sharded axis -> Reduce (partial, neutral identity) + AllReduceCombine
    [+ fold of a non-neutral identity]."""

    reduce = xf.PatternNode(Reduce)
    out = xf.PatternNode(nodes.AccessNode)

    sharded = properties.DictProperty(key_type=str, value_type=int, default={},
                                      desc='array name -> axis along which every rank holds one block')

    @classmethod
    def expressions(cls):
        return [dace.sdfg.utils.node_path_graph(cls.reduce, cls.out)]

    def can_be_applied(self, graph, expr_index, sdfg, permissive=False):
        node = self.reduce
        if getattr(node, '_b200_combined', False) or not _reduce_needs_combine(node, graph, self.sharded):
            return False
        if node.identity is None:
            return False          # folding into existing output would count it once per rank
        try:
            _redtype_of(node.wcr)
        except NotImplementedError:
            return False
        oe = graph.out_edges(node)[0]
        return (_full_contiguous(oe.data, sdfg.arrays[oe.data.data])
                and _through_views(sdfg, graph, self.out) is not None)

    def apply(self, graph, sdfg):
        node, acc = self.reduce, _through_views(sdfg, graph, self.out)
        rt = _redtype_of(node.wcr)
        desc = sdfg.arrays[acc.data]
        neutral = neutral_element(rt, desc.dtype)
        identity = node.identity
        node.identity = neutral
        node._b200_combined = True
        after = _append_combine(sdfg, graph, acc, node.wcr)
        try:
            same = bool(dace.symbolic.pystr_to_symbolic(str(identity)) == dace.symbolic.pystr_to_symbolic(str(neutral)))
        except Exception:
            same = False
        if not same:
            _fold_scalar_after(sdfg, graph, after, rt, identity)


def _shard_scatter(sdfg, state, node, wcr: str):
    """ScatterReduce over sharded keys: scatter into a neutral private copy (filled in a state
    before), combine it, fold it into the bins with an ordinary GPU map."""
    oe = [e for e in state.out_edges(node) if e.src_conn == '_bins'][0]
    bins_acc = oe.dst
    bins = bins_acc.data
    desc = sdfg.arrays[bins]
    if not _full_contiguous(oe.data, desc) or len(desc.shape) != 1:
        raise NotImplementedError('sharded ScatterReduce: %s must cover a whole 1-D array' % oe.data)
    rt = _redtype_of(wcr)
    pname, pdesc = sdfg.add_transient('__b200_partial_' + bins, desc.shape, desc.dtype, storage=desc.storage,
                                      find_new_name=True)
    init = sdfg.add_state_before(state, 'b200_neutral_' + bins)
    _, me, _ = init.add_mapped_tasklet('neutral_' + pname, {'__i': '0:%s' % desc.shape[0]}, {},
                            '__o = %s(%s);' % (desc.dtype.ctype, symstr(neutral_element(rt, desc.dtype), cpp_mode=True)),
                            {'__o': dace.Memlet('%s[__i]' % pname)}, schedule=_map_schedule(desc),
                            language=dace.Language.CPP, external_edges=True)
    _block256(me, desc)
    part = state.add_access(pname)
    wcr_on_edge = oe.data.wcr
    state.remove_edge(oe)
    m = dace.Memlet.from_array(pname, pdesc)
    m.wcr = wcr_on_edge
    state.add_edge(node, '_bins', part, None, m)
    combined = _append_combine(sdfg, state, part, wcr)
    old = state.add_read(bins)
    code = '__o = %s(__a, __p);' % _wcr_cpp(rt, desc.dtype.ctype)
    _, me, _ = state.add_mapped_tasklet('fold_partial_' + bins, {'__i': '0:%s' % desc.shape[0]},
                             {'__a': dace.Memlet('%s[__i]' % bins), '__p': dace.Memlet('%s[__i]' % pname)}, code,
                             {'__o': dace.Memlet('%s[__i]' % bins)}, schedule=_map_schedule(desc),
                             language=dace.Language.CPP, input_nodes={bins: old, pname: combined},
                             output_nodes={bins: bins_acc}, external_edges=True)
    _block256(me, desc)


def shard_reductions(sdfg, sharded: Dict[str, int]) -> Dict[str, int]:
    """Insert the multi-GPU combine wherever a reduction folds over a sharded axis.

    ``sharded`` maps program arrays (as every rank sees them: its own block) to the axis that is
    split over the ranks, e.g. ``{'A': 0}`` for row shards, ``{'keys': 0}`` for key shards.  Call
    it after ``apply_gpu_transformations()`` / :func:`dace_b200.auto.apply_b200`.  Returns counts
    ``{'reduce': n, 'scatter': m, 'local': k}`` (``local`` = reductions that keep the sharded axis
    and need no collective)."""
    from .scatter_node import ScatterReduce
    sh = propagate_sharding(sdfg, sharded)
    counts = {'reduce': 0, 'scatter': 0, 'local': 0}
    # Reductions that were already fused into generated kernels / single-pass nodes no longer show their axes:
    # refuse instead of silently leaving a partial result uncombined.
    for state in sdfg.states():
        for node in state.nodes():
            fused = (isinstance(node, nodes.Tasklet) and node.label == 'fused_map_reduce_b200') or \
                type(node).__name__ == 'RowsMaxSum'
            if fused and any(e.data.data in sh for e in state.in_edges(node)):
                raise NotImplementedError('%s reads the sharded array(s) %s: call shard_reductions() BEFORE the fusing '
                                          'rewrites (dace_b200.auto.apply_b200(sdfg, fuse=False), then shard, then fuse)' %
                                          (node.label, sorted(e.data.data for e in state.in_edges(node) if e.data.data in sh)))
    for state in list(sdfg.states()):
        for node in list(state.nodes()):
            if isinstance(node, Reduce) and state.in_edges(node)[0].data.data in sh:
                if not _reduce_needs_combine(node, state, sh):
                    counts['local'] += 1
                    continue
                if getattr(node, '_b200_combined', False):
                    continue
                acc = state.out_edges(node)[0].dst
                t = ShardedReduceCombine()
                t.setup_match(sdfg, sdfg.cfg_id, sdfg.node_id(state),
                              {ShardedReduceCombine.reduce: state.node_id(node),
                               ShardedReduceCombine.out: state.node_id(acc)}, 0)
                t.sharded = sh
                if not t.can_be_applied(state, 0, sdfg):
                    raise NotImplementedError('cannot combine %s over sharded %s (identity None, custom operator or '
                                              'partial output memlet)' % (node, state.in_edges(node)[0].data))
                t.apply(state, sdfg)
                counts['reduce'] += 1
            elif isinstance(node, ScatterReduce):
                idx = [e for e in state.in_edges(node) if e.dst_conn == '_idx'][0]
                if idx.data.data in sh and not getattr(node, '_b200_combined', False):
                    _shard_scatter(sdfg, state, node, node.wcr)
                    node._b200_combined = True
                    counts['scatter'] += 1
    return counts
