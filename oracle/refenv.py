"""Make the read-only reference checkout (/root/reference) importable here.

TEST INFRASTRUCTURE ONLY.  Nothing under ``dace_b200/`` may import this module;
it is used by ``oracle/make_golden.py`` (golden-vector generation),
``oracle/build_ref.py`` (compiling the reference's own CPU expansions into
``oracle/_ref/``) and by the CPU-side tests that exercise the drop-in
registration against the real ``dace.library`` registry.

The reference does not import out of the box in this image: ``astunparse``,
``ordered_set``, ``ply`` and ``fparser`` are missing and networkx 3.6 dropped
the ``start -> start`` entry from ``immediate_dominators`` which
reference dace/sdfg/analysis/cfg.py:409 indexes.  ``oracle/shims`` provides
stand-ins; this module wires them in.  /root/reference does not exist on the
GPU box, so everything that calls :func:`import_reference` must tolerate
``None``.
"""
import os
import sys

REFERENCE_ROOT = os.environ.get('DACE_REFERENCE_ROOT', '/root/reference')
_SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'shims')


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, 'dace'))


def _patch_networkx():
    import networkx as nx
    if getattr(nx, '_dace_b200_idom_patched', False):
        return
    orig = nx.immediate_dominators

    def immediate_dominators(G, start):
        idom = dict(orig(G, start))
        idom.setdefault(start, start)
        return idom

    nx.immediate_dominators = immediate_dominators
    try:
        import networkx.algorithms.dominance as dom
        dom.immediate_dominators = immediate_dominators
    except Exception:  # pragma: no cover
        pass
    nx._dace_b200_idom_patched = True


def import_reference(cache_dir: str = None):
    """Return the reference ``dace`` module, or ``None`` when unavailable."""
    if not reference_available():
        return None
    # The image's default g++ wrapper cannot find libgomp.spec; use the system one.
    os.environ['CC'] = '/usr/bin/gcc'
    os.environ['CXX'] = '/usr/bin/g++'
    os.environ['CUDAHOSTCXX'] = '/usr/bin/g++'
    os.environ.setdefault('DACE_compiler_cuda_cuda_arch', '100')
    if cache_dir is not None:
        os.environ['DACE_default_build_folder'] = cache_dir
    for p in (REFERENCE_ROOT, _SHIMS):
        if p not in sys.path:
            sys.path.insert(0, p)
    _patch_networkx()
    import dace  # noqa: E402
    return dace
