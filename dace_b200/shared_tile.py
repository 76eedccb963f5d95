"""DaCe plug-in: privatised shared-memory WCR tiles flushed by ``include/dace_b200_shared_accum.cuh``.

The reference's only "privatise, then flush with atomics" primitive is a GPU_Shared transient between two
map exits whose outgoing memlet carries the WCR - what ``AccumulateTransient`` (reference
dace/transformation/dataflow/stream_transient.py:136) builds and what the CUDA code generator lowers to
``dace::ResetShared<...>::Reset`` + ``dace::SharedToGlobal1D<...>::template Accum<REDTYPE>`` (reference
dace/codegen/targets/cuda.py:744-749, :1379-1417; dace/runtime/include/dace/cuda/copy.cuh:706-812).

* :func:`use_b200_shared_tiles` puts the device header into the program's .cu file.  The header specialises
  those two runtime templates for the arithmetic element types, so the generated call sites stay exactly as
  the reference emits them and pick up the sm_100 flush (neutral entries skipped, ``red.global``, vector fp32
  reductions, no compare-and-swap loops for float min / max).  For every tile it finds it also emits a
  ``static_assert`` on the specialisation, so a program that silently fell back to the reference's version
  does not compile.
* :func:`privatise_wcr_in_shared` is the convenience wrapper around the reference's own transformation:
  ``AccumulateTransient`` between a thread-block map and its GPU_Device parent, transient moved to shared
  memory and zeroed on allocation (the steps of reference samples/optimization/matmul.py:174-181).
"""
import os

import dace
from dace import dtypes, nodes

from . import _lib

HEADER = os.path.join(_lib.INCLUDE_DIR, 'dace_b200_shared_accum.cuh')
_MARK = '/* dace_b200 shared tiles */'
_SPECIALISED = {'float', 'double', 'int', 'unsigned int', 'long long', 'unsigned long long'}


def shared_wcr_tiles(sdfg):
    """(state, access node) of every GPU_Shared transient that is flushed through a WCR memlet or reset on
    allocation."""
    found = []
    for sd in sdfg.all_sdfgs_recursive():
        for state in sd.states():
            for n in state.data_nodes():
                desc = n.desc(sd)
                if desc.storage != dtypes.StorageType.GPU_Shared:
                    continue
                if n.setzero or any(e.data.wcr is not None for e in state.out_edges(n)):
                    found.append((state, n))
    return found


def use_b200_shared_tiles(sdfg) -> int:
    """Route the program's shared-tile resets / WCR flushes to ``dace_b200::shared_reset`` / ``shared_accum``.
    Returns the number of tiles found (0: nothing to do, nothing added)."""
    tiles = shared_wcr_tiles(sdfg)
    if not tiles:
        return 0
    have = ''.join((cb.as_string or '') for cb in sdfg.global_code.values())
    if _MARK not in have:
        sdfg.append_global_code('%s\n#include "%s"\n' % (_MARK, HEADER), 'cuda')
    for state, n in tiles:
        desc = n.desc(state.parent if hasattr(state, 'parent') else sdfg)
        ctype = desc.dtype.ctype
        if ctype not in _SPECIALISED:
            raise NotImplementedError('shared WCR tile %s: element type %s has no sm_100 flush' % (n.data, ctype))
        # one instantiation is enough to prove that the specialisation exists for this element type
        if 'static_assert(dace::SharedToGlobal1D<%s, 32, 1, 1, 32, false>' % ctype in have:
            continue
        have += 'static_assert(dace::SharedToGlobal1D<%s, 32, 1, 1, 32, false>' % ctype
        sdfg.append_global_code('static_assert(dace::SharedToGlobal1D<%s, 32, 1, 1, 32, false>::kDaceB200 == 1 && '
                                'dace::ResetShared<%s, 32, 1, 1, 32, 1, false>::kDaceB200 == 1, '
                                '"dace_b200_shared_accum.cuh must come after <dace/dace.h>");\n' % (ctype, ctype), 'cuda')
    return len(tiles)


def privatise_wcr_in_shared(sdfg, state, inner_exit: nodes.MapExit, outer_exit: nodes.MapExit, array: str = None):
    """WCR output of a thread-block map -> shared-memory tile, reset on allocation, flushed once per block.
    Returns the tile's access node."""
    from dace.transformation.dataflow import AccumulateTransient
    from .redtypes import ReductionType, detect_reduction_type
    for e in state.edges_between(inner_exit, outer_exit):
        if e.data.wcr is not None and (array is None or e.data.data == array):
            rt = detect_reduction_type(e.data.wcr)
            if rt not in (ReductionType.Sum, ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor):
                # the tile is reset to zero bits on allocation, which is the neutral element of these only
                raise NotImplementedError('privatise_wcr_in_shared: %s needs an explicit neutral fill '
                                          '(AccumulateTransient with identity=...)' % rt.name)
    before = {n for n in state.data_nodes()}
    opts = {'array': array} if array else {}
    AccumulateTransient.apply_to(sdfg, options=opts or None, map_exit=inner_exit, outer_map_exit=outer_exit)
    new = [n for n in state.data_nodes() if n not in before and n.desc(sdfg).transient]
    if len(new) != 1:
        raise RuntimeError('AccumulateTransient created %d transients' % len(new))
    tile = new[0]
    tile.desc(sdfg).storage = dtypes.StorageType.GPU_Shared
    tile.setzero = True
    use_b200_shared_tiles(sdfg)
    return tile
