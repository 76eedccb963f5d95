"""Host-side mirror of the reference's Reduce operator on the B200 C ABI.

``reduce`` has the argument meaning of ``dace.reduce(op, in_array, out_array,
axis, identity)`` (reference dace/frontend/operations.py:240-258) and of the
``Reduce`` library node's properties ``wcr`` / ``axes`` / ``identity``
(reference dace/libraries/standard/nodes/reduce.py:1346-1411):

* ``identity=None`` folds into the values already in ``out_array``;
* ``axis=None`` reduces all axes; negative axes count from the end;
* views with arbitrary strides are accepted on both sides (memlet subsets);
* the output dtype is the input dtype unless ``out_array`` says otherwise
  (bf16/f16 inputs may write fp32; accumulation is fp32 either way).

torch is used only to own device memory and streams.  Every call goes through
``libdace_b200_reduce.so``; nothing here computes on the CPU.
"""
import ctypes
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib
from .redtypes import ReductionType, resolve

F32, F64, I32, I64, U32, U64, BF16, F16, I8, U8, I16, U16 = range(12)

_TORCH2DT = {
    torch.float32: F32, torch.float64: F64, torch.int32: I32, torch.int64: I64,
    torch.bfloat16: BF16, torch.float16: F16, torch.int8: I8, torch.uint8: U8, torch.int16: I16,
    torch.bool: U8,
}
for _name, _dt in (('uint32', U32), ('uint64', U64), ('uint16', U16)):
    if hasattr(torch, _name):
        _TORCH2DT[getattr(torch, _name)] = _dt

_NP_OF_DT = {F32: np.float32, F64: np.float64, I32: np.int32, I64: np.int64, U32: np.uint32, U64: np.uint64,
             I8: np.int8, U8: np.uint8, I16: np.int16, U16: np.uint16}


def dtype_code(t: torch.dtype) -> int:
    try:
        return _TORCH2DT[t]
    except KeyError:
        raise NotImplementedError('dtype %s has no sm_100a reduction kernel' % (t, ))


def scalar_bytes(value, dt: int) -> bytes:
    """Encode a Python/NumPy scalar as one element of C-ABI dtype ``dt``."""
    if dt == BF16:
        return torch.tensor(float(value), dtype=torch.bfloat16).view(torch.int16).numpy().tobytes()
    if dt == F16:
        with np.errstate(over='ignore'):      # e.g. -3e38 as "minus infinity" for a max
            return np.float16(value).tobytes()
    npdt = _NP_OF_DT[dt]
    if np.issubdtype(npdt, np.integer):
        if isinstance(value, (bool, np.bool_)):
            value = int(value)
        return np.array(value).astype(npdt, casting='unsafe').tobytes()
    return npdt(value).tobytes()


_raw_stream = getattr(torch._C, '_cuda_getCurrentRawStream', None)


class _on_device:
    """``with torch.cuda.device(d)`` only when ``d`` is not already current: the context manager costs
    several microseconds per call, which is the size of a small reduction."""
    __slots__ = ('ctx', )

    def __init__(self, device):
        self.ctx = None if device.index is None or device.index == torch.cuda.current_device() \
            else torch.cuda.device(device)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            return self.ctx.__exit__(*exc)
        return False


def _stream_ptr(stream) -> int:
    if stream is None:
        if _raw_stream is not None:
            return _raw_stream(torch.cuda.current_device())
        return torch.cuda.current_stream().cuda_stream
    if isinstance(stream, int):
        return stream
    return stream.cuda_stream


class Workspace:
    """Device workspace for cross-CTA combining (tickets + partials, plus optional extra scratch for the
    two-pass plans): allocated by the CALLER - here with torch - once, reused; the protocol of the
    reference's CUB temp storage (reference dace/libraries/standard/nodes/reduce.py:550-570).  The C ABI never
    allocates.  One workspace serves one stream at a time."""

    def __init__(self, device=None, stream=None, nbytes: Optional[int] = None, zero_init: bool = True):
        lib = _lib.load()
        base = int(lib.dace_b200_workspace_bytes())
        self.nbytes = max(int(nbytes), base if zero_init else 256) if nbytes is not None else base
        self.buffer = torch.empty(self.nbytes, dtype=torch.uint8, device=device or 'cuda')
        if zero_init:
            _lib.check(lib.dace_b200_workspace_init(self.buffer.data_ptr(), self.nbytes, _stream_ptr(stream)),
                       'dace_b200_workspace_init')

    @property
    def ptr(self) -> int:
        return self.buffer.data_ptr()


# Default workspaces of this host API: one per (device, stream), created on first use and grown on demand
# (before the C call, never inside it).  Outgrown buffers stay alive: kernels enqueued earlier may still use them.
_WORKSPACES = {}
_SCRATCH = {}
_RETIRED = []


def _default_workspace(cache, device: torch.device, stream_ptr: int, need: int, zero_init: bool) -> Workspace:
    key = (device.index if device.index is not None else torch.cuda.current_device(), stream_ptr)
    ws = cache.get(key)
    if ws is None or ws.nbytes < need:
        if ws is not None:
            _RETIRED.append(ws)
        with _on_device(device):
            ws = Workspace(device, stream_ptr, nbytes=need, zero_init=zero_init)
        cache[key] = ws
    return ws


def _require_cuda(t: torch.Tensor, what: str):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        # Same rule as the reference's GPU expansions, which insist on GPU_Global /
        # CPU_Pinned storage (reference reduce.py:499-507) - but no CPU fallback here.
        raise ValueError('%s must be a CUDA tensor: the B200 backend has no CPU fallback' % what)


def _normalize_axes(axis, ndim: int):
    if axis is None:
        return tuple(range(ndim))
    if isinstance(axis, int):
        axis = (axis, )
    out = []
    for a in axis:
        a = int(a)
        if a < -ndim or a >= ndim:
            raise ValueError('axis %d out of range for %d dimensions' % (a, ndim))
        out.append(a % ndim)
    if len(set(out)) != len(out):
        raise ValueError('duplicate reduction axis')
    return tuple(sorted(out))


def reduce(op, in_array: torch.Tensor, out_array: Optional[torch.Tensor] = None,
           axis: Union[None, int, Sequence[int]] = None, identity=None, stream=None,
           workspace: Optional[Workspace] = None) -> Optional[torch.Tensor]:
    """``out_array[kept] = fold(op, identity, in_array[kept, reduced])`` on the GPU.

    Returns the new tensor when ``out_array`` is None (then ``identity`` is
    required), else None - the reference's convention."""
    lib = _lib.load()
    rt = resolve(op)
    _require_cuda(in_array, 'in_array')
    axes = _normalize_axes(axis, in_array.dim())
    kept = [i for i in range(in_array.dim()) if i not in axes]
    kept_shape = [in_array.shape[i] for i in kept]
    created = out_array is None
    if created:
        if identity is None:
            raise ValueError('identity is required when out_array is not given')
        out_dtype = in_array.dtype
        out_array = torch.empty(kept_shape, dtype=out_dtype, device=in_array.device)
    _require_cuda(out_array, 'out_array')
    if out_array.device != in_array.device:
        raise ValueError('in_array and out_array live on different devices')
    # The output may keep unit dims (e.g. C[0:N, 5, 5, 0:C]) - squeeze like the
    # reference does (reference reduce.py:40-54) and match the kept extents.
    out_sq_shape = [s for s in out_array.shape if s != 1]
    kept_sq_shape = [s for s in kept_shape if s != 1]
    if out_sq_shape != kept_sq_shape:
        raise ValueError('out_array shape %s does not match the kept axes %s' %
                         (tuple(out_array.shape), tuple(kept_shape)))
    out_sq_strides = [st for s, st in zip(out_array.shape, out_array.stride()) if s != 1]
    out_strides = []
    it = iter(out_sq_strides)
    for s in kept_shape:
        out_strides.append(0 if s == 1 else next(it))

    dt_in, dt_out = dtype_code(in_array.dtype), dtype_code(out_array.dtype)
    nd = in_array.dim()
    shape = (ctypes.c_int64 * max(nd, 1))(*in_array.shape)
    istr = (ctypes.c_int64 * max(nd, 1))(*in_array.stride())
    ostr = (ctypes.c_int64 * max(len(kept), 1))(*out_strides)
    mask = 0
    for a in axes:
        mask |= 1 << a
    ident_buf = None
    if identity is not None:
        ident_buf = ctypes.create_string_buffer(scalar_bytes(identity, dt_out), 8)
    sptr = _stream_ptr(stream)
    if workspace is None:
        # multi-pass plans (interleaved axes, padded views) want scratch behind the base workspace
        need = int(lib.dace_b200_reduce_nd_workspace_bytes(dt_in, dt_out, nd, shape, istr, mask)) if nd >= 2 else 0
        workspace = _default_workspace(_WORKSPACES, in_array.device, sptr, need, True)
    with _on_device(in_array.device):
        status = lib.dace_b200_reduce_nd(int(rt), dt_in, dt_out, in_array.data_ptr(), out_array.data_ptr(), nd,
                                         shape, istr, ostr, mask, ident_buf, workspace.ptr, workspace.nbytes, sptr)
    _lib.check(status, 'dace_b200_reduce_nd')
    return out_array if created else None


def rows_max_sum(in_array: torch.Tensor, out_max: Optional[torch.Tensor] = None,
                 out_sum: Optional[torch.Tensor] = None, stream=None):
    """Row-wise max and sum of a 2-D array in one pass (fp32 outputs)."""
    lib = _lib.load()
    _require_cuda(in_array, 'in_array')
    if in_array.dim() != 2 or in_array.stride(1) != 1:
        raise ValueError('rows_max_sum expects a 2-D array with contiguous rows')
    rows, n = in_array.shape
    if out_max is None:
        out_max = torch.empty(rows, dtype=torch.float32, device=in_array.device)
    if out_sum is None:
        out_sum = torch.empty(rows, dtype=torch.float32, device=in_array.device)
    for o in (out_max, out_sum):
        _require_cuda(o, 'output')
        if o.dtype != torch.float32 or not o.is_contiguous() or o.numel() != rows:
            raise ValueError('outputs must be contiguous float32[rows]')
    with _on_device(in_array.device):
        status = lib.dace_b200_reduce_rows_max_sum(dtype_code(in_array.dtype), in_array.data_ptr(),
                                                   out_max.data_ptr(), out_sum.data_ptr(), rows, n,
                                                   in_array.stride(0), _stream_ptr(stream))
    _lib.check(status, 'dace_b200_reduce_rows_max_sum')
    return out_max, out_sum


def _arg_reduce(red: int, in_array: torch.Tensor, axis, out_idx, out_val, init, idx_dtype, stream, workspace):
    lib = _lib.load()
    _require_cuda(in_array, 'in_array')
    a = in_array
    if axis is None:
        a = a.reshape(-1)      # a view when possible, like the frontend's flat() (reduction.py:281-283)
        axis = 0
    axis = _normalize_axes(axis, a.dim())[0]
    # [outer, reduced, inner] with one stride each: the kept axes, in order, must merge into at most two dense groups
    # (contiguous arrays: leading and trailing axes; windows of larger arrays: e.g. [A, B, C] with a pitch after
    # every row and every slice, reduced over C -> outer = A, inner = B)
    groups = []
    for i in [d for d in range(a.dim()) if d != axis]:
        if a.shape[i] == 1:
            continue
        if groups and groups[-1][2] == i - 1 and i - 1 != axis and groups[-1][1] == a.stride(i) * a.shape[i]:
            groups[-1] = (groups[-1][0] * a.shape[i], a.stride(i), i)
        else:
            groups.append((a.shape[i], a.stride(i), i))
    if len(groups) > 2:
        raise NotImplementedError('argmin/argmax: the kept dimensions must merge into at most two dense groups')
    reduced = a.shape[axis]
    sr = a.stride(axis)
    if len(groups) == 2:
        (outer, so, _), (inner, si, _) = groups
    elif len(groups) == 1 and groups[0][2] > axis:
        outer, so, (inner, si, _) = 1, 0, groups[0]
    elif len(groups) == 1:
        (outer, so, _), inner, si = groups[0], 1, 1
    else:
        outer, so, inner, si = 1, 0, 1, 1
    outer, inner = int(outer), int(inner)
    kept = tuple(a.shape[:axis]) + tuple(a.shape[axis + 1:])
    created = out_idx is None
    if created:
        if init is None:
            raise ValueError('init=None folds into existing outputs: out_idx and out_val are required')
        out_idx = torch.empty(kept, dtype=idx_dtype, device=a.device)
    if out_val is None and init is None:
        raise ValueError('init=None folds into existing outputs: out_val is required')
    for o, what in ((out_idx, 'out_idx'), (out_val, 'out_val')):
        if o is None:
            continue
        _require_cuda(o, what)
        if not o.is_contiguous() or o.numel() != outer * inner:
            raise ValueError('%s must be contiguous with one element per kept position' % what)
    if out_idx.dtype not in (torch.int32, torch.int64):
        raise NotImplementedError('index dtype %s' % out_idx.dtype)
    if out_val is not None and out_val.dtype != a.dtype:
        raise ValueError('out_val must have the input dtype')
    dt_in = dtype_code(a.dtype)
    if isinstance(init, str):
        if a.dtype.is_floating_point:
            init = torch.finfo(a.dtype).min if red == int(ReductionType.Max_Location) else torch.finfo(a.dtype).max
        else:
            info = torch.iinfo(a.dtype) if a.dtype != torch.bool else None
            init = (info.min if red == int(ReductionType.Max_Location) else info.max) if info else \
                (0 if red == int(ReductionType.Max_Location) else 1)
    ibuf = ctypes.create_string_buffer(scalar_bytes(init, dt_in), 8) if init is not None else None
    sptr = _stream_ptr(stream)
    if workspace is None:
        workspace = _default_workspace(_WORKSPACES, a.device, sptr, 0, True)
    with _on_device(a.device):
        status = lib.dace_b200_reduce_arg(red, dt_in, dtype_code(out_idx.dtype), a.data_ptr(), out_idx.data_ptr(),
                                          out_val.data_ptr() if out_val is not None else None, outer, reduced, inner,
                                          so, sr, si, 0, 0, ibuf, workspace.ptr, workspace.nbytes, sptr)
    _lib.check(status, 'dace_b200_reduce_arg')
    return out_idx if created else None


def argmax(in_array: torch.Tensor, axis: Optional[int] = None, out_idx: Optional[torch.Tensor] = None,
           out_val: Optional[torch.Tensor] = None, init='default', idx_dtype=torch.int32, stream=None,
           workspace: Optional[Workspace] = None):
    """``np.argmax(in_array, axis)`` with the fold rule of the reference frontend's lowering
    (reference dace/frontend/python/replacements/reduction.py:245-348): first maximum wins; the
    pair starts as ``(init, -1)`` with ``init = min_value(dtype)`` by default, so a slice with no
    element above ``init`` yields -1.  ``init=None`` folds into ``out_val`` / ``out_idx``."""
    return _arg_reduce(int(ReductionType.Max_Location), in_array, axis, out_idx, out_val, init, idx_dtype, stream,
                       workspace)


def argmin(in_array: torch.Tensor, axis: Optional[int] = None, out_idx: Optional[torch.Tensor] = None,
           out_val: Optional[torch.Tensor] = None, init='default', idx_dtype=torch.int32, stream=None,
           workspace: Optional[Workspace] = None):
    """``np.argmin`` - see :func:`argmax`."""
    return _arg_reduce(int(ReductionType.Min_Location), in_array, axis, out_idx, out_val, init, idx_dtype, stream,
                       workspace)


def wcr_scatter(op, bins: torch.Tensor, indices: torch.Tensor, values=None, stream=None,
                workspace: Optional[Workspace] = None) -> None:
    """``bins[indices[i]] = op(bins[indices[i]], values[i])`` - the WCR map
    ``out >> bins(1, wcr)[:]; out[idx] = v`` of reference
    samples/explicit/histogram.py:13-26 and ``hist[bin] += w`` of reference
    tests/npbench/misc/azimint_hist_test.py:39-59.  ``values`` may be a tensor,
    a scalar (applied for every index) or None (== 1).  ``bins`` is updated in
    place and must be initialised by the caller (WCR semantics)."""
    lib = _lib.load()
    rt = resolve(op)
    _require_cuda(bins, 'bins')
    _require_cuda(indices, 'indices')
    if not bins.is_contiguous() or not indices.is_contiguous():
        raise ValueError('bins and indices must be contiguous')
    dt_val, dt_idx = dtype_code(bins.dtype), dtype_code(indices.dtype)
    vptr, cbuf = None, None
    if isinstance(values, torch.Tensor):
        _require_cuda(values, 'values')
        if values.dtype != bins.dtype or values.numel() != indices.numel() or not values.is_contiguous():
            raise ValueError('values must be contiguous, of bins.dtype and as long as indices')
        vptr = values.data_ptr()
    elif values is not None:
        cbuf = ctypes.create_string_buffer(scalar_bytes(values, dt_val), 8)
    sptr = _stream_ptr(stream)
    ws_ptr, ws_bytes = None, 0
    if workspace is None and indices.numel() >= (1 << 22):
        # large bin arrays: scratch of the radix-partition path (0 = the path does not apply)
        need = int(lib.dace_b200_wcr_scatter_workspace_bytes(int(rt), dt_idx, dt_val, 1 if vptr else 0,
                                                             indices.numel(), bins.numel()))
        if need:
            workspace = _default_workspace(_SCRATCH, bins.device, sptr, need, False)
    if workspace is not None:
        ws_ptr, ws_bytes = workspace.ptr, workspace.nbytes
    with _on_device(bins.device):
        status = lib.dace_b200_wcr_scatter(int(rt), dt_idx, dt_val, indices.data_ptr(), vptr, cbuf,
                                           indices.numel(), bins.data_ptr(), bins.numel(), ws_ptr, ws_bytes, sptr)
    _lib.check(status, 'dace_b200_wcr_scatter')


def histogram(indices: torch.Tensor, nbins: int, dtype=torch.int32, stream=None) -> torch.Tensor:
    """Counts per bin: zero-initialised bins + Sum WCR scatter of ones."""
    bins = torch.zeros(nbins, dtype=dtype, device=indices.device)
    wcr_scatter(ReductionType.Sum, bins, indices, None, stream=stream)
    return bins
