"""Reduction-type classification for WCR lambdas.

Mirrors ``dace.frontend.operations.detect_reduction_type`` (reference
dace/frontend/operations.py:130-185) and the ``dace.dtypes.ReductionType`` enum
(reference dace/dtypes.py:93-113; C++ mirror dace/runtime/include/dace/types.h:301-317)
without importing DaCe or sympy: the reference evaluates the lambda on two sympy
symbols and compares the result with ``a + b``, ``Max(a, b)`` ...; here the same
decision is taken structurally on the lambda's AST.  Numeric values equal the
C++ enum, which is what the C ABI takes.
"""
import ast
import functools
import enum
from typing import Optional, Union


class ReductionType(enum.IntEnum):
    Custom = 0
    Min = 1
    Max = 2
    Sum = 3
    Product = 4
    Logical_And = 5
    Bitwise_And = 6
    Logical_Or = 7
    Bitwise_Or = 8
    Logical_Xor = 9
    Bitwise_Xor = 10
    Min_Location = 11
    Max_Location = 12
    Exchange = 13
    # OpenMP-only extensions in the reference (dace/dtypes.py:110-112)
    Sub = 14
    Div = 15


_BINOPS = {
    ast.Add: ReductionType.Sum,
    ast.Mult: ReductionType.Product,
    ast.BitAnd: ReductionType.Bitwise_And,
    ast.BitOr: ReductionType.Bitwise_Or,
    ast.BitXor: ReductionType.Bitwise_Xor,
}


def _lambda_of(wcr: Union[str, ast.AST]):
    tree = ast.parse(wcr.strip()) if isinstance(wcr, str) else wcr
    node = tree.body[0].value if isinstance(tree, ast.Module) else tree
    if not isinstance(node, ast.Lambda) or len(node.args.args) != 2:
        raise ValueError('WCR must be a two-argument lambda, got: %r' % (wcr, ))
    return node, node.args.args[0].arg, node.args.args[1].arg


def _is_pair(x, y, a, b):
    return (isinstance(x, ast.Name) and isinstance(y, ast.Name) and {x.id, y.id} == {a, b} and x.id != y.id)


def detect_reduction_type(wcr_str: Optional[str], openmp: bool = False) -> Optional[ReductionType]:
    """Classify a WCR lambda string; same answers as the reference's function."""
    if wcr_str == '' or wcr_str is None:
        return None
    lam, a, b = _lambda_of(wcr_str)
    body = lam.body
    if isinstance(body, ast.Call) and isinstance(body.func, ast.Name) and body.func.id in ('max', 'min'):
        return ReductionType.Max if body.func.id == 'max' else ReductionType.Min
    if isinstance(body, ast.BinOp) and _is_pair(body.left, body.right, a, b):
        t = _BINOPS.get(type(body.op))
        if t is not None:
            return t
        if openmp and isinstance(body.op, ast.Sub) and body.left.id == a:
            return ReductionType.Sub
        if openmp and isinstance(body.op, ast.Div) and body.left.id == a:
            return ReductionType.Div
        return ReductionType.Custom
    if isinstance(body, ast.BoolOp) and isinstance(body.op, ast.And):
        return ReductionType.Logical_And
    if isinstance(body, ast.BoolOp) and isinstance(body.op, ast.Or):
        return ReductionType.Logical_Or
    if isinstance(body, ast.Compare) and isinstance(body.ops[0], ast.NotEq):
        return ReductionType.Logical_Xor
    if isinstance(body, ast.Name) and body.id == b:
        return ReductionType.Exchange
    return ReductionType.Custom


def lower_custom(wcr_str: str) -> ReductionType:
    """Recognise hand-written min/max conditionals that the reference classifies
    as Custom (e.g. ``lambda a, b: a if a < b else b``, reference
    tests/custom_reduce_test.py:7-17) so the precompiled kernels can serve them.
    Returns ``ReductionType.Custom`` when the lambda is genuinely custom."""
    lam, a, b = _lambda_of(wcr_str)
    body = lam.body
    if not (isinstance(body, ast.IfExp) and isinstance(body.test, ast.Compare) and len(body.test.ops) == 1):
        return ReductionType.Custom
    t, yes, no = body.test, body.body, body.orelse
    left, right = t.left, t.comparators[0]
    if not (_is_pair(left, right, a, b) and _is_pair(yes, no, a, b)):
        return ReductionType.Custom
    op = type(t.ops[0])
    if op in (ast.Lt, ast.LtE):
        picks_smaller = yes.id == left.id
    elif op in (ast.Gt, ast.GtE):
        picks_smaller = yes.id == right.id
    else:
        return ReductionType.Custom
    return ReductionType.Min if picks_smaller else ReductionType.Max


@functools.lru_cache(maxsize=256)
def _classify_lambda(op: str) -> ReductionType:
    # parsing a lambda costs ~20 us; hot loops pass the same string on every call
    rt = detect_reduction_type(op)
    if rt == ReductionType.Custom:
        rt = lower_custom(op)
    return rt


def resolve(op) -> ReductionType:
    """Accept a ReductionType, its name, or a lambda string; reject what the
    precompiled sm_100a kernels cannot run (no CPU fallback)."""
    if isinstance(op, ReductionType):
        rt = op
    elif isinstance(op, int):
        rt = ReductionType(op)
    elif isinstance(op, str) and op in ReductionType.__members__:
        rt = ReductionType[op]
    elif isinstance(op, str):
        rt = _classify_lambda(op)
    else:
        raise TypeError('reduction operator must be a ReductionType, its name or a lambda string')
    if rt is None or rt < ReductionType.Min or rt > ReductionType.Bitwise_Xor:
        raise NotImplementedError('reduction type %s has no precompiled sm_100a kernel '
                                  '(custom lambdas / argmin / argmax / exchange are not supported)' % (rt, ))
    return rt
