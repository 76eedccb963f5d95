"""DaCe plug-in: one map with SEVERAL write-conflict outputs -> one map per output.

The reference lowers every WCR output of a map to its own atomic per element (reference
dace/codegen/targets/cpp.py:694-770, dace/codegen/targets/cpu.py:965-1013); a map such as the one in reference
tests/multiple_cr_test.py (``s += a`` and ``p *= a`` from two tasklets of the same map) therefore issues two
same-address atomics per element.  The rewrites of this package (``WCRMapToFusedReduce``, ``WCRMapToScatterReduce``)
take maps with ONE WCR output.  When the outputs of a map are computed by disjoint sets of tasklets, the map is split
here into one map per output - each reads its own inputs again, which costs a second streaming pass and saves one
serialised atomic per element - and every piece is then lowered on its own.

The split is only kept when every piece is taken by one of the two rewrites (checked on a copy of the SDFG first);
otherwise the map is left exactly as it was.
"""
import copy
from typing import Dict, List, Optional

import dace
from dace import nodes


def _slices(state, entry: nodes.MapEntry) -> Optional[List[set]]:
    """Per in-edge of the map exit: the scope nodes its value depends on.  None if the map is not splittable: nested
    scopes, plain (non-WCR) outputs, dynamic map ranges, fewer than two outputs, or outputs that share a node."""
    if state.entry_node(entry) is not None:
        return None
    mexit = state.exit_node(entry)
    inner = state.in_edges(mexit)
    if len(inner) < 2:
        return None
    body = set(state.scope_subgraph(entry, include_entry=False, include_exit=False).nodes())
    if any(isinstance(n, (nodes.EntryNode, nodes.ExitNode)) for n in body):
        return None
    if any(e.dst_conn is None or not e.dst_conn.startswith('IN_') for e in state.in_edges(entry)):
        return None
    if any(e.data.wcr is None or e.dst_conn is None for e in inner):
        return None
    if any(e.data.wcr is None for e in state.out_edges(mexit)):
        return None
    slices = []
    for e in inner:
        seen, todo = set(), [e.src]
        while todo:
            n = todo.pop()
            if n in seen or n is entry:
                continue
            if n not in body:
                return None
            seen.add(n)
            todo.extend(p.src for p in state.in_edges(n))
        slices.append(seen)
    for i in range(len(slices)):
        for j in range(i + 1, len(slices)):
            if slices[i] & slices[j]:
                return None
    if set().union(*slices) != body:
        return None                       # something in the scope feeds no output: not ours to judge
    return slices


def split_map(sdfg, state, entry: nodes.MapEntry) -> List[nodes.MapEntry]:
    """Split a map whose WCR outputs come from disjoint sets of tasklets; returns the map entries (the original one
    first).  The caller has checked ``_slices``."""
    slices = _slices(state, entry)
    mexit = state.exit_node(entry)
    inner_out = state.in_edges(mexit)
    result = [entry]
    for j in range(1, len(slices)):
        w = inner_out[j]
        new_map = nodes.Map('%s_wcr%d' % (entry.map.label, j), list(entry.map.params), copy.deepcopy(entry.map.range),
                            schedule=entry.map.schedule)
        new_map.debuginfo = entry.map.debuginfo
        new_entry, new_exit = nodes.MapEntry(new_map), nodes.MapExit(new_map)
        state.add_node(new_entry)
        state.add_node(new_exit)
        for n in slices[j]:
            for e in list(state.in_edges(n)):
                if e.src is not entry:
                    continue
                if e.src_conn is None:                                  # ordering edge of an input-less tasklet
                    state.add_edge(new_entry, None, n, e.dst_conn, copy.deepcopy(e.data))
                else:
                    name = e.src_conn[4:]
                    if 'IN_' + name not in new_entry.in_connectors:
                        new_entry.add_in_connector('IN_' + name)
                        new_entry.add_out_connector('OUT_' + name)
                        for o in state.in_edges(entry):
                            if o.dst_conn == 'IN_' + name:
                                state.add_edge(o.src, o.src_conn, new_entry, 'IN_' + name, copy.deepcopy(o.data))
                    state.add_edge(new_entry, e.src_conn, n, e.dst_conn, copy.deepcopy(e.data))
                state.remove_edge(e)
        name = w.dst_conn[3:]
        new_exit.add_in_connector('IN_' + name)
        new_exit.add_out_connector('OUT_' + name)
        state.add_edge(w.src, w.src_conn, new_exit, 'IN_' + name, copy.deepcopy(w.data))
        for o in list(state.out_edges(mexit)):
            if o.src_conn == 'OUT_' + name:
                state.add_edge(new_exit, 'OUT_' + name, o.dst, o.dst_conn, copy.deepcopy(o.data))
                state.remove_edge(o)
        state.remove_edge(w)
        mexit.remove_in_connector('IN_' + name)
        mexit.remove_out_connector('OUT_' + name)
        result.append(new_entry)
    # inputs of the original map that only the moved tasklets used
    for conn in list(entry.out_connectors):
        if not any(e.src_conn == conn for e in state.out_edges(entry)):
            for o in list(state.in_edges(entry)):
                if o.dst_conn == 'IN_' + conn[4:]:
                    state.remove_edge(o)
            entry.remove_out_connector(conn)
            entry.remove_in_connector('IN_' + conn[4:])
    return result


def _taken(sd, state, entry) -> bool:
    from . import fused_expansion as _fe
    from .auto import scatter_is_eligible
    return _fe._match_wcr(state, sd, entry) is not None or scatter_is_eligible(sd, state, entry)


def split_wcr_outputs(sdfg) -> int:
    """Split every top-level map with several independent WCR outputs whose pieces the sm_100a rewrites take
    (call after ``apply_gpu_transformations``, before / as part of ``apply_b200``).  Returns the number of maps split."""
    done = 0
    for sd_index, sd in enumerate(list(sdfg.all_sdfgs_recursive())):
        for state_index, state in enumerate(sd.states()):
            for entry in [n for n in state.nodes() if isinstance(n, nodes.MapEntry)]:
                if entry not in state.nodes() or _slices(state, entry) is None:
                    continue
                # dry run on a copy: keep the split only if every piece is lowered afterwards
                trial = copy.deepcopy(sdfg)
                tsd = list(trial.all_sdfgs_recursive())[sd_index]
                tstate = list(tsd.states())[state_index]
                tentry = tstate.node(state.node_id(entry))
                if not isinstance(tentry, nodes.MapEntry) or _slices(tstate, tentry) is None:
                    continue
                pieces = split_map(tsd, tstate, tentry)
                if not all(_taken(tsd, tstate, p) for p in pieces):
                    continue
                split_map(sd, state, entry)
                done += 1
    return done
