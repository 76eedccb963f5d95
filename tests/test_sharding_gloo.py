"""Host logic of the multi-GPU path on CPU: world_size-2 gloo groups.

The local fold is injected (the oracle on CPU tensors stands in for the CUDA
kernel - test infrastructure only); what is under test is dace_b200/sharding.py:
shard bounds, when a collective is needed, identity handling across ranks and the
all-reduce operator per reduction type."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dace_b200 import sharding
from dace_b200.redtypes import ReductionType, resolve


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 8, 1000, 2**22):
        for world in (1, 2, 3, 8):
            blocks = [sharding.shard_bounds(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
            ref = [len(x) for x in np.array_split(np.arange(n), world)]
            assert sizes == ref


def test_needs_combine():
    assert sharding.needs_combine(None, 2)
    assert sharding.needs_combine(0, 2) and sharding.needs_combine((0, 1), 3)
    assert not sharding.needs_combine(1, 2) and not sharding.needs_combine((-1, ), 2)
    assert sharding.needs_combine(-2, 2)


def test_neutral_elements():
    assert sharding.neutral_element(ReductionType.Sum, torch.float32) == 0
    assert sharding.neutral_element(ReductionType.Product, torch.int64) == 1
    assert sharding.neutral_element(ReductionType.Min, torch.int32) == 2**31 - 1
    assert sharding.neutral_element(ReductionType.Max, torch.float64) == float('-inf')
    assert sharding.neutral_element(ReductionType.Bitwise_And, torch.int32) == -1


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_reduce(op, a, out, axis=None, identity=None):
    from oracle import oracle as O
    red = {'Sum': O.SUM, 'Product': O.PRODUCT, 'Max': O.MAX, 'Min': O.MIN, 'Bitwise_Or': O.BOR,
           'Bitwise_And': O.BAND, 'Bitwise_Xor': O.BXOR}[resolve(op).name]
    o = out.numpy()
    if identity is None:
        O.reduce(red, a.numpy(), axis, identity=None, out=o, wide=True)
    else:
        if isinstance(identity, float) and np.isinf(identity):
            identity = np.finfo(o.dtype).max * np.sign(identity)
        o[...] = O.reduce(red, a.numpy(), axis, identity=identity, wide=True).reshape(o.shape)


def _oracle_scatter(op, bins, idx, values=None):
    from oracle import oracle as O
    O.wcr_scatter(O.SUM, bins.numpy(), idx.numpy(), None if values is None else values.numpy())


def _worker(rank, world, port, q):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        full = rng.integers(-50, 50, (101, 37)).astype(np.int64)
        lo, hi = sharding.shard_bounds(101, world, rank)
        shard = torch.from_numpy(full[lo:hi].copy())
        res = {}
        # sharded axis reduced (column sum over row shards, C5 shape class): combine
        out = torch.zeros(37, dtype=torch.int64)
        sharding.sharded_reduce('lambda a, b: a + b', shard, out, axis=0, identity=100, local_reduce=_oracle_reduce)
        res['col_sum'] = bool(np.array_equal(out.numpy(), 100 + full.sum(0)))
        # full reduction (C2 shape class), max with identity None folding into the existing output
        out = torch.tensor(7, dtype=torch.int64)
        sharding.sharded_reduce('lambda a, b: max(a, b)', shard, out, axis=None, identity=None,
                                local_reduce=_oracle_reduce)
        res['full_max_noident'] = int(out.item()) == max(7, int(full.max()))
        out = torch.tensor(0, dtype=torch.int64)
        sharding.sharded_reduce('lambda a, b: a ^ b', shard, out, axis=None, identity=0, local_reduce=_oracle_reduce)
        res['full_xor'] = int(out.item()) == int(np.bitwise_xor.reduce(full.reshape(-1)))
        # sharded axis kept (row sums over row shards, C1/C3 shape class): no collective, local slice
        out = torch.zeros(hi - lo, dtype=torch.int64)
        sharding.sharded_reduce('lambda a, b: a + b', shard, out, axis=1, identity=0, local_reduce=_oracle_reduce)
        res['row_sum_local'] = bool(np.array_equal(out.numpy(), full[lo:hi].sum(1)))
        # float column sums: fp32 within 1e-5 of fp64
        ff = rng.random((64, 16)).astype(np.float32)
        flo, fhi = sharding.shard_bounds(64, world, rank)
        out = torch.zeros(16, dtype=torch.float32)
        sharding.sharded_reduce('lambda a, b: a + b', torch.from_numpy(ff[flo:fhi].copy()), out, axis=0, identity=0,
                                local_reduce=_oracle_reduce)
        res['col_sum_f32'] = bool(np.allclose(out.numpy(), ff.astype(np.float64).sum(0), rtol=1e-5))
        # histogram over key shards (C4 shape class): bit-exact, existing bin contents counted once
        keys = rng.integers(0, 500, 10001).astype(np.int32)
        klo, khi = sharding.shard_bounds(10001, world, rank)
        bins = torch.full((500, ), 3, dtype=torch.int32)
        sharding.sharded_wcr_scatter('lambda a, b: a + b', bins, torch.from_numpy(keys[klo:khi].copy()),
                                     local_scatter=_oracle_scatter)
        res['hist'] = bool(np.array_equal(bins.numpy(), 3 + np.bincount(keys, minlength=500)))
        q.put((rank, res))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world', [2, 3])
def test_sharded_paths_over_gloo(world):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, res in results:
        assert all(res.values()), (rank, res)
