"""Minimal stand-in for the `astunparse` package, built on the stdlib unparser.

TEST INFRASTRUCTURE ONLY (oracle side): lets the read-only reference checkout
import in this image, where `astunparse` is not installed.  The reference
subclasses ``astunparse.Unparser`` and overrides ``_Constant``, ``_Attribute``
and ``_Subscript`` (reference dace/frontend/python/astutils.py:210-241), and
calls ``astunparse.unparse`` in a few other places.
"""
import ast
import sys
from io import StringIO


class Unparser(ast._Unparser):
    """``Unparser(tree, file)`` writes the source of ``tree`` to ``file``."""

    def __init__(self, tree, file=sys.stdout):
        super().__init__()
        self._source = []
        self.dispatch(tree)
        file.write(''.join(self._source) + '\n')
        file.flush()

    # astunparse API used by subclasses -------------------------------------
    def dispatch(self, tree):
        if isinstance(tree, list):
            for t in tree:
                self.dispatch(t)
            return
        self.traverse(tree)

    def visit(self, node):
        # Route through subclass hooks named ``_<ClassName>`` when present.
        hook = getattr(self, '_' + node.__class__.__name__, None)
        if hook is not None:
            return hook(node)
        return super().visit(node)

    # Default hooks forward to the stdlib visitors --------------------------
    def _Constant(self, t):
        return super().visit_Constant(t)

    def _Attribute(self, t):
        return super().visit_Attribute(t)

    def _Subscript(self, t):
        return super().visit_Subscript(t)


def unparse(tree):
    v = StringIO()
    Unparser(tree, file=v)
    return v.getvalue()
