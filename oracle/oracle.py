"""NumPy-facing wrapper of the CPU oracle (oracle/reduce_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu_baseline / ``--impl reference`` legs; never by ``dace_b200``.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, 'reduce_oracle.c')
LIB = os.path.join(_HERE, 'build', 'libreduce_oracle.so')

F32, F64, I32, I64, U32, U64, BF16, F16, I8, U8, I16, U16 = range(12)
MIN, MAX, SUM, PRODUCT, LAND, BAND, LOR, BOR, LXOR, BXOR = range(1, 11)
MIN_LOCATION, MAX_LOCATION = 11, 12

_NP2DT = {np.dtype(np.float32): F32, np.dtype(np.float64): F64, np.dtype(np.int32): I32, np.dtype(np.int64): I64,
          np.dtype(np.uint32): U32, np.dtype(np.uint64): U64, np.dtype(np.float16): F16, np.dtype(np.int8): I8,
          np.dtype(np.uint8): U8, np.dtype(np.int16): I16, np.dtype(np.uint16): U16, np.dtype(np.bool_): U8}

_lib = None


def build(force: bool = False) -> str:
    """gcc -O2 -fwrapv (signed overflow wraps like the reference's two's-complement machines)."""
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(['/usr/bin/gcc', '-O2', '-fwrapv', '-fPIC', '-shared', '-o', LIB, SRC])
    return LIB


def load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB)
        _lib.oracle_reduce_nd.restype = ctypes.c_int
        _lib.oracle_reduce_nd.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64),
                                          ctypes.POINTER(ctypes.c_int64), ctypes.c_uint32, ctypes.c_void_p, ctypes.c_int]
        _lib.oracle_wcr_scatter.restype = ctypes.c_int
        _lib.oracle_wcr_scatter.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]
        _lib.oracle_reduce_arg.restype = ctypes.c_int
        _lib.oracle_reduce_arg.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_void_p] + [ctypes.c_int64] * 8 + [ctypes.c_void_p]
    return _lib


def dtype_code(a: np.ndarray, bf16: bool = False) -> int:
    if bf16:
        assert a.dtype == np.uint16
        return BF16
    return _NP2DT[a.dtype]


def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16 bit patterns (uint16)."""
    w = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    w = w + 0x7fff + ((w >> 16) & 1)
    return (w >> 16).astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


def _scalar(value, dt: int) -> bytes:
    if dt == BF16:
        return f32_to_bf16_bits(np.array([value], np.float32)).tobytes()
    npdt = {F32: np.float32, F64: np.float64, I32: np.int32, I64: np.int64, U32: np.uint32, U64: np.uint64,
            F16: np.float16, I8: np.int8, U8: np.uint8, I16: np.int16, U16: np.uint16}[dt]
    return np.array(value).astype(npdt, casting='unsafe').tobytes()


def reduce(red: int, a: np.ndarray, axes=None, identity=None, out: np.ndarray = None, wide: bool = False,
           in_bf16: bool = False, out_dtype=None, out_bf16: bool = False) -> np.ndarray:
    """Fold ``a`` over ``axes`` (None = all) like the reference's Reduce node.
    ``a`` may be any strided view.  bf16 data travels as uint16 bit patterns
    (``in_bf16`` / ``out_bf16``).  ``out`` given + ``identity=None`` folds into it."""
    lib = load()
    nd = a.ndim
    axes = tuple(range(nd)) if axes is None else tuple(sorted(x % nd for x in ([axes] if np.isscalar(axes) else axes)))
    kept = [i for i in range(nd) if i not in axes]
    dt_in = dtype_code(a, in_bf16)
    if out is None:
        if identity is None:
            raise ValueError('identity required without out')
        odt = np.dtype(out_dtype) if out_dtype is not None else (np.dtype(np.uint16) if in_bf16 else a.dtype)
        if odt == np.bool_:
            odt = np.dtype(np.uint8)
        out = np.empty([a.shape[i] for i in kept], dtype=odt)
        if in_bf16 and out_dtype is None:
            out_bf16 = True
    dt_out = dtype_code(out, out_bf16)
    isz = a.dtype.itemsize
    shape = (ctypes.c_int64 * max(nd, 1))(*a.shape)
    istr = (ctypes.c_int64 * max(nd, 1))(*[s // isz for s in a.strides])
    osq = [s // out.dtype.itemsize for n, s in zip(out.shape, out.strides) if n != 1]
    it = iter(osq)
    ostr_l = [0 if a.shape[i] == 1 else next(it) for i in kept]
    ostr = (ctypes.c_int64 * max(len(kept), 1))(*ostr_l)
    mask = 0
    for x in axes:
        mask |= 1 << x
    ident = ctypes.create_string_buffer(_scalar(identity, dt_out), 8) if identity is not None else None
    rc = lib.oracle_reduce_nd(red, dt_in, dt_out, a.ctypes.data, out.ctypes.data, nd, shape, istr, ostr, mask, ident,
                              1 if wide else 0)
    if rc != 0:
        raise RuntimeError('oracle_reduce_nd failed with %d' % rc)
    return out


def wcr_scatter(red: int, bins: np.ndarray, idx: np.ndarray, values=None) -> np.ndarray:
    """In-place ``bins[idx[i]] = wcr(bins[idx[i]], values[i] | const | 1)``."""
    lib = load()
    assert bins.flags.c_contiguous and idx.flags.c_contiguous
    dt_val, dt_idx = dtype_code(bins), dtype_code(idx)
    vptr, cbuf = None, None
    if isinstance(values, np.ndarray):
        values = np.ascontiguousarray(values, dtype=bins.dtype)
        vptr = values.ctypes.data
    elif values is not None:
        cbuf = ctypes.create_string_buffer(_scalar(values, dt_val), 8)
    rc = lib.oracle_wcr_scatter(red, dt_idx, dt_val, idx.ctypes.data, vptr, cbuf, idx.size, bins.ctypes.data, bins.size)
    if rc != 0:
        raise RuntimeError('oracle_wcr_scatter failed with %d' % rc)
    return bins


def default_arg_init(red: int, a: np.ndarray, in_bf16: bool = False):
    """What the frontend's init state writes: dtypes.min_value / max_value of the input type
    (reference dace/frontend/python/replacements/reduction.py:303-304, dace/dtypes.py:426-449)."""
    if in_bf16:
        return -3.3895313892515355e38 if red == MAX_LOCATION else 3.3895313892515355e38
    info = np.finfo(a.dtype) if np.issubdtype(a.dtype, np.floating) else np.iinfo(a.dtype)
    return info.min if red == MAX_LOCATION else info.max


def reduce_arg(red: int, a: np.ndarray, axis: int = None, init='default', out_idx: np.ndarray = None,
               out_val: np.ndarray = None, idx_dtype=np.int32, in_bf16: bool = False):
    """argmin / argmax of ``a`` along ``axis`` (None = flattened) with the reference's fold rule.
    Returns (idx, val).  ``init=None`` folds into the given ``out_idx`` / ``out_val``."""
    lib = load()
    if axis is None:
        a = a.reshape(-1)
        axis = 0
    axis %= a.ndim
    a = np.ascontiguousarray(a)
    outer = int(np.prod(a.shape[:axis], dtype=np.int64))
    reduced = a.shape[axis]
    inner = int(np.prod(a.shape[axis + 1:], dtype=np.int64))
    kept = a.shape[:axis] + a.shape[axis + 1:]
    dt = dtype_code(a, in_bf16)
    if out_idx is None:
        out_idx = np.empty(kept, dtype=idx_dtype)
    if out_val is None:
        out_val = np.empty(kept, dtype=a.dtype)
    assert out_idx.flags.c_contiguous and out_val.flags.c_contiguous
    if isinstance(init, str):
        init = default_arg_init(red, a, in_bf16)
    ibuf = ctypes.create_string_buffer(_scalar(init, dt), 8) if init is not None else None
    rc = lib.oracle_reduce_arg(red, dt, out_idx.dtype.itemsize, a.ctypes.data, out_idx.ctypes.data, out_val.ctypes.data,
                               outer, reduced, inner, reduced * inner, inner, 1, 0, 0, ibuf)
    if rc != 0:
        raise RuntimeError('oracle_reduce_arg failed with %d' % rc)
    return out_idx, out_val
