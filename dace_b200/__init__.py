"""dace_b200 - B200 (sm_100a) reduction backend for DaCe's Reduce / WCR hot path.

Public host API (mirrors the reference operator interface, see api.py):
``reduce``, ``argmax``, ``argmin``, ``rows_max_sum``, ``wcr_scatter``, ``histogram``, ``Workspace``,
``ReductionType``, ``detect_reduction_type``.  The DaCe plug-in (library
expansion ``'B200'``) lives in ``dace_b200.expansion`` and is imported only when
DaCe itself is importable.
"""
from .redtypes import ReductionType, detect_reduction_type, lower_custom, resolve  # noqa: F401
from ._lib import B200LibraryError, EXPORTED_SYMBOLS, LIB_PATH, build, load  # noqa: F401


def __getattr__(name):
    # torch-backed entry points are imported lazily so that `import dace_b200`
    # (e.g. for build()) does not pay for `import torch`.
    if name in ('reduce', 'rows_max_sum', 'wcr_scatter', 'histogram', 'Workspace', 'argmax', 'argmin'):
        from . import api as _r
        return getattr(_r, name)
    raise AttributeError(name)
