"""Outer-dimension sharding of a reduction over the GPUs of one box.

One process per GPU (DaCe's own rule, reference dace/library.py:206-219), plumbing
through ``torch.distributed`` (NCCL over NVLink / NVSwitch).  The reference has no
NCCL path; its analogue is the MPI library node that all-reduces a buffer after
the local computation (reference dace/libraries/mpi/nodes/allreduce.py:9-60).

* sharded axis is REDUCED (full sum, column sum over row shards, histogram over
  key shards): every rank folds its shard into a full-shape partial, then ONE
  all-reduce combines the partials on the same stream as the local kernel;
* sharded axis is KEPT (row reductions over row shards): outputs are disjoint
  slices and no collective is issued.

The user identity is folded exactly once (rank 0); the other ranks start from
the operator's neutral element, so non-neutral identities and ``identity=None``
(fold into the existing output) keep their single-GPU meaning.
"""
from typing import Callable, Optional, Sequence, Tuple, Union

import torch

from .redtypes import ReductionType, resolve


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block ``[lo, hi)`` of an axis of extent ``n`` owned by ``rank``
    (first ``n % world`` ranks get one extra element, like ``numpy.array_split``)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError('bad rank/world')
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def needs_combine(axis: Union[None, int, Sequence[int]], ndim: int, shard_axis: int = 0) -> bool:
    """True when the sharded axis is one of the reduced axes (partials must be combined)."""
    if axis is None:
        return True
    axes = (axis, ) if isinstance(axis, int) else tuple(axis)
    return (shard_axis % ndim) in {a % ndim for a in axes}


def neutral_element(rt: ReductionType, dtype: torch.dtype):
    """The operator's neutral element for ``dtype`` (what non-root ranks start from)."""
    if rt in (ReductionType.Sum, ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor, ReductionType.Logical_Or,
              ReductionType.Logical_Xor):
        return 0
    if rt in (ReductionType.Product, ReductionType.Logical_And):
        return 1
    if rt == ReductionType.Bitwise_And:
        return -1 if dtype.is_signed else torch.iinfo(dtype).max
    if rt in (ReductionType.Min, ReductionType.Max):
        if dtype.is_floating_point:
            return float('inf') if rt == ReductionType.Min else float('-inf')
        info = torch.iinfo(dtype)
        return info.max if rt == ReductionType.Min else info.min
    raise NotImplementedError(rt)


def collective_op(rt: ReductionType):
    """The all-reduce operator that combines per-rank partials of ``rt``."""
    import torch.distributed as dist
    table = {
        ReductionType.Sum: dist.ReduceOp.SUM, ReductionType.Product: dist.ReduceOp.PRODUCT,
        ReductionType.Min: dist.ReduceOp.MIN, ReductionType.Max: dist.ReduceOp.MAX,
        ReductionType.Bitwise_And: dist.ReduceOp.BAND, ReductionType.Bitwise_Or: dist.ReduceOp.BOR,
        ReductionType.Bitwise_Xor: dist.ReduceOp.BXOR,
        # logical partials are 0/1, so the bitwise collectives are exact
        ReductionType.Logical_And: dist.ReduceOp.BAND, ReductionType.Logical_Or: dist.ReduceOp.BOR,
        ReductionType.Logical_Xor: dist.ReduceOp.BXOR,
    }
    return table[rt]


class PeerCombiner:
    """One-shot all-reduce of small partials through NVLink peer memory
    (``dace_b200_peer_allreduce``, dace_b200/csrc/peer_combine.cu).

    Every rank allocates one exchange buffer in the kernel library, ships its CUDA IPC
    handle to the other ranks of ``group`` (``all_gather_object``: plumbing only) and maps
    theirs.  :meth:`allreduce` then needs no NCCL call: one small kernel pushes the partial
    to all peers, raises a flag and folds what arrived in rank order.  Payloads up to
    ``cap_bytes``; larger ones and unsupported dtypes/operators stay with NCCL."""

    _DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.uint32: 4, torch.uint64: 5}
    _OPS = (ReductionType.Sum, ReductionType.Product, ReductionType.Min, ReductionType.Max, ReductionType.Bitwise_And,
            ReductionType.Bitwise_Or, ReductionType.Bitwise_Xor)

    def __init__(self, group=None, cap_bytes: int = 64 << 10):
        import ctypes
        import torch.distributed as dist
        from . import _lib
        self._lib = _lib.load()
        self._check = _lib.check
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 16:
            raise ValueError('PeerCombiner supports up to 16 GPUs of one box')
        self.cap_bytes = int(cap_bytes)
        self.epoch = 0
        nbytes = self._lib.dace_b200_peer_buffer_bytes(self.world, self.cap_bytes)
        own = ctypes.c_void_p()
        self._check(self._lib.dace_b200_peer_alloc(nbytes, ctypes.byref(own)), 'dace_b200_peer_alloc')
        self._own = own.value
        handle = ctypes.create_string_buffer(64)
        self._check(self._lib.dace_b200_peer_export(self._own, handle), 'dace_b200_peer_export')
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self._opened = []
        bases = (ctypes.c_void_p * self.world)()
        for r, h in enumerate(handles):
            if r == self.rank:
                bases[r] = self._own
                continue
            ptr = ctypes.c_void_p()
            self._check(self._lib.dace_b200_peer_open(ctypes.create_string_buffer(h, 64), ctypes.byref(ptr)),
                        'dace_b200_peer_open')
            bases[r] = ptr.value
            self._opened.append(ptr.value)
        self._bases = bases
        dist.barrier(group=group)   # everybody has mapped everybody before the first push

    def supports(self, partial: torch.Tensor, rt: ReductionType) -> bool:
        return (partial.is_cuda and partial.is_contiguous() and partial.dtype in self._DT and rt in self._OPS
                and 0 < partial.numel() * partial.element_size() <= self.cap_bytes)

    def allreduce(self, partial: torch.Tensor, rt: ReductionType, stream=None) -> torch.Tensor:
        self.epoch += 1
        s = stream if stream is not None else torch.cuda.current_stream(partial.device)
        self._check(self._lib.dace_b200_peer_allreduce(int(rt), self._DT[partial.dtype], partial.data_ptr(),
                                                       partial.numel(), self._bases, self.rank, self.world, self.epoch,
                                                       self.cap_bytes, s.cuda_stream), 'dace_b200_peer_allreduce')
        return partial

    def close(self):
        """Unmap the peers and free the own buffer (all ranks' streams must be idle)."""
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        for p in self._opened:
            self._lib.dace_b200_peer_close(p)
        self._opened = []
        dist.barrier(group=self.group)
        if self._own:
            self._lib.dace_b200_peer_free(self._own)
            self._own = None


_PEER = {}


def enable_peer_combine(group=None, cap_bytes: int = 64 << 10) -> PeerCombiner:
    """Route small all-reduces of :func:`combine` through NVLink peer memory for ``group``
    (collective call: every rank of the group must make it)."""
    pc = PeerCombiner(group, cap_bytes)
    _PEER[group] = pc
    return pc


def disable_peer_combine(group=None):
    pc = _PEER.pop(group, None)
    if pc is not None:
        pc.close()


def combine(partial: torch.Tensor, op, group=None) -> torch.Tensor:
    """All-reduce ``partial`` in place across the ranks (enqueued on the current stream): the
    one-shot peer-memory kernel when enabled and applicable, else one NCCL / gloo all-reduce."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return partial
    rt = resolve(op)
    pc = _PEER.get(group)
    if pc is not None and pc.supports(partial, rt):
        return pc.allreduce(partial, rt)
    dist.all_reduce(partial, op=collective_op(rt), group=group)
    return partial


def sharded_reduce(op, local_shard: torch.Tensor, out: torch.Tensor, axis=None, identity=None, shard_axis: int = 0,
                   group=None, local_reduce: Optional[Callable] = None) -> torch.Tensor:
    """Reduce an array that is sharded along ``shard_axis`` over the ranks of ``group``.

    ``local_shard`` is this rank's block; ``out`` is the full-shape result when the
    sharded axis is reduced (replicated on every rank after the call), or this
    rank's slice of the result when it is kept.  ``local_reduce`` defaults to the
    CUDA path (:func:`dace_b200.reduce`); tests inject a CPU stand-in to exercise
    this host logic over gloo."""
    import torch.distributed as dist
    if local_reduce is None:
        from .api import reduce as local_reduce
    rt = resolve(op)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world == 1 or not needs_combine(axis, local_shard.dim(), shard_axis):
        local_reduce(op, local_shard, out, axis=axis, identity=identity)
        return out
    if rank == 0:
        if identity is None:
            partial = out.clone()
            local_reduce(op, local_shard, partial, axis=axis, identity=None)
        else:
            partial = torch.empty_like(out)
            local_reduce(op, local_shard, partial, axis=axis, identity=identity)
    else:
        partial = torch.empty_like(out)
        local_reduce(op, local_shard, partial, axis=axis, identity=neutral_element(rt, out.dtype))
    combine(partial, rt, group)
    out.copy_(partial)
    return out


def sharded_wcr_scatter(op, bins: torch.Tensor, local_indices: torch.Tensor, values=None, group=None,
                        local_scatter: Optional[Callable] = None) -> torch.Tensor:
    """Histogram / scatter-add over key shards: each rank scatters its keys into a
    neutral-initialised private copy of ``bins``; copies are all-reduced and folded
    into ``bins`` once (integer counts stay bit-exact)."""
    import torch.distributed as dist
    if local_scatter is None:
        from .api import wcr_scatter as local_scatter
    rt = resolve(op)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        local_scatter(op, bins, local_indices, values)
        return bins
    rank = dist.get_rank(group)
    private = bins.clone() if rank == 0 else torch.full_like(bins, neutral_element(rt, bins.dtype))
    local_scatter(op, private, local_indices, values)
    combine(private, rt, group)
    bins.copy_(private)
    return bins
