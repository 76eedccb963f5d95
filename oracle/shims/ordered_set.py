"""Loads the `ordered_set` module that setuptools vendors (MIT, third party).

TEST INFRASTRUCTURE ONLY: the reference imports ``ordered_set`` (reference
dace/sdfg/scope.py:11, dace/sdfg/graph.py:11) and the package is not installed
in this image.
"""
import glob
import importlib.util
import sys

_cands = (glob.glob('/usr/lib/python3/dist-packages/setuptools/_vendor/ordered_set.py') +
          glob.glob(sys.prefix + '/lib/python3*/site-packages/setuptools/_vendor/ordered_set.py') +
          glob.glob(sys.prefix + '/lib/python3*/site-packages/setuptools/_vendor/ordered_set/__init__.py') +
          glob.glob(sys.prefix + '/lib/python3*/site-packages/pkg_resources/_vendor/ordered_set.py'))
if not _cands:
    raise ImportError('no vendored ordered_set found for the reference shim')
_spec = importlib.util.spec_from_file_location('_vendored_ordered_set', _cands[0])
_mod = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mod)
OrderedSet = _mod.OrderedSet
__all__ = ['OrderedSet']
