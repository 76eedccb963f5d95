"""Multi-GPU (NCCL) check of the sharded paths; needs >= 2 visible GPUs, else skipped.
Launches one process per GPU like bench.py --gpus N does."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["REPO_ROOT"])
import dace_b200 as db
from dace_b200 import sharding
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
rng = np.random.default_rng(0)
full = rng.integers(-50, 50, (4099, 520)).astype(np.int32)
lo, hi = sharding.shard_bounds(full.shape[0], world, rank)
shard = torch.from_numpy(full[lo:hi].copy()).to(dev)
out = torch.zeros(520, dtype=torch.int32, device=dev)
sharding.sharded_reduce("lambda a, b: a + b", shard, out, axis=0, identity=5)
assert np.array_equal(out.cpu().numpy(), 5 + full.sum(0)), "column sum over row shards"
tot = torch.zeros((), dtype=torch.int32, device=dev)
sharding.sharded_reduce("lambda a, b: max(a, b)", shard, tot, axis=None, identity=-2**31)
assert tot.item() == full.max()
rows = torch.zeros(hi - lo, dtype=torch.int32, device=dev)
sharding.sharded_reduce("lambda a, b: a + b", shard, rows, axis=1, identity=0)
assert np.array_equal(rows.cpu().numpy(), full[lo:hi].sum(1))
keys = rng.integers(0, 1 << 20, 3000001).astype(np.int32)
klo, khi = sharding.shard_bounds(keys.size, world, rank)
bins = torch.zeros(1 << 20, dtype=torch.int32, device=dev)
sharding.sharded_wcr_scatter("lambda a, b: a + b", bins, torch.from_numpy(keys[klo:khi].copy()).to(dev))
assert np.array_equal(bins.cpu().numpy(), np.bincount(keys, minlength=1 << 20))
x = torch.rand(1 << 22, device=dev, generator=torch.Generator(device=dev).manual_seed(rank))
s = torch.zeros((), device=dev)
sharding.sharded_reduce("lambda a, b: a + b", x, s, axis=None, identity=0)
ref = torch.tensor([x.double().sum().item()], dtype=torch.float64, device=dev)
dist.all_reduce(ref)
assert abs(s.item() - ref.item()) <= 1e-5 * ref.item()
# ---- the same partials through the one-shot NVLink peer-memory kernel instead of NCCL ----
pc = sharding.enable_peer_combine()
for it in range(40):                      # many epochs: both slots, flags only ever grow
    part = torch.full((520,), rank + it, dtype=torch.int32, device=dev)
    assert pc.supports(part, sharding.resolve("lambda a, b: a + b"))
    sharding.combine(part, "lambda a, b: a + b")
    assert bool((part == sum(r + it for r in range(world))).all()), "peer all-reduce, epoch %d" % it
assert pc.epoch == 40
out2 = torch.zeros(520, dtype=torch.int32, device=dev)
sharding.sharded_reduce("lambda a, b: a + b", shard, out2, axis=0, identity=5)
assert torch.equal(out2, out), "peer-combined column sum == NCCL-combined column sum"
tot2 = torch.zeros((), dtype=torch.int32, device=dev)
sharding.sharded_reduce("lambda a, b: max(a, b)", shard, tot2, axis=None, identity=-2**31)
assert tot2.item() == full.max()
s2 = torch.zeros((), device=dev)
sharding.sharded_reduce("lambda a, b: a + b", x, s2, axis=None, identity=0)
assert abs(s2.item() - ref.item()) <= 1e-5 * ref.item()
gathered = [torch.zeros((), device=dev) for _ in range(world)]
dist.all_gather(gathered, s2)
assert all(g.item() == s2.item() for g in gathered), "rank-ordered fold: bit-identical on every rank"
d64 = torch.full((8192,), 0.5 + rank, dtype=torch.float64, device=dev)   # 64 KiB: the largest peer payload
sharding.combine(d64, "lambda a, b: min(a, b)")
assert bool((d64 == 0.5).all())
big = torch.ones(1 << 20, dtype=torch.int32, device=dev)                 # 4 MiB: stays with NCCL
assert not pc.supports(big, sharding.resolve("lambda a, b: a + b"))
sharding.combine(big, "lambda a, b: a + b")
assert bool((big == world).all())
sharding.disable_peer_combine()
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_sharded_reduce_nccl(tmp_path):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs >= 2 GPUs')
    n = min(n, 4)
    script = tmp_path / 'worker.py'
    script.write_text(WORKER)
    env = dict(os.environ, REPO_ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(n),
                        '--master-addr', '127.0.0.1', '--master-port', '29577', str(script)],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count(' ok') == n


# ---- native communicator (dace_b200_comm_*) and the DaCe-built sharded program, one process per GPU ----

COMM_WORKER = r'''
import ctypes, os, sys
isolate = os.environ["ISOLATE"] == "1"
if isolate:     # DaCe's process model: the rank's GPU is device 0 of the process (reference codegen/targets/cuda.py:446-455)
    os.environ["CUDA_VISIBLE_DEVICES"] = os.environ["LOCAL_RANK"]
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["REPO_ROOT"])
from dace_b200 import _lib
from oracle import ref_loader
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = 0 if isolate else int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("gloo")          # plumbing for the explicit-id route and for the reference values
lib = _lib.load()
# (a) explicit rendezvous: id made on rank 0, shipped by the host
ident = ctypes.create_string_buffer(128)
if rank == 0:
    _lib.check(lib.dace_b200_comm_unique_id(ident), "unique_id")
box = [ident.raw]
dist.broadcast_object_list(box, src=0)
comm = ctypes.c_void_p()
_lib.check(lib.dace_b200_comm_create(rank, world, ctypes.create_string_buffer(box[0], 128), ctypes.byref(comm)), "comm_create")
assert lib.dace_b200_comm_rank(comm) == rank and lib.dace_b200_comm_world(comm) == world
peer = lib.dace_b200_comm_uses_peer_memory(comm)
want_peer = os.environ.get("DACE_B200_NO_PEER") != "1"
if not want_peer:
    assert peer == 0
elif not isolate:
    assert peer == 1, "peer-memory path must come up when every GPU is visible to every rank"
print("rank", rank, "peer-memory path:", peer, "isolated:", isolate)
DT = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3}
s = torch.cuda.current_stream().cuda_stream
def allreduce(t, red):
    _lib.check(lib.dace_b200_comm_allreduce(comm, red, DT[t.dtype], t.data_ptr(), t.numel(), s), "comm_allreduce")
    torch.cuda.synchronize()
    return t
def gathered(t):
    out = [torch.zeros_like(t.cpu()) for _ in range(world)]
    dist.all_gather(out, t.cpu())
    return torch.stack(out)
g = torch.Generator(device=dev).manual_seed(100 + rank)
for n in (1, 4096, 16384, 1 << 20):                     # 4 B ... 64 KiB: peer kernel; 4 MiB: NCCL
    x = torch.randint(-1000, 1000, (n,), device=dev, dtype=torch.int32, generator=g)
    ref = gathered(x)
    assert torch.equal(allreduce(x.clone(), 3).cpu(), ref.sum(0).to(torch.int32)), "int32 sum n=%d" % n
    assert torch.equal(allreduce(x.clone(), 2).cpu(), ref.max(0).values), "int32 max n=%d" % n
    f = torch.rand(n, device=dev, dtype=torch.float64, generator=g)[: max(1, n // 2)]
    rf = gathered(f)
    got = allreduce(f.clone(), 3).cpu()
    assert torch.allclose(got, rf.sum(0), rtol=1e-12), "f64 sum"
    assert torch.equal(allreduce(f.clone(), 1).cpu(), rf.min(0).values), "f64 min"
    allg = gathered(got)
    assert bool((allg == allg[0]).all()), "result identical on every rank"
b = torch.full((100,), 1 << rank, device=dev, dtype=torch.int32)
if peer:
    assert bool((allreduce(b.clone(), 8) == (1 << world) - 1).all()), "bitwise or on the peer path"
    big = torch.ones(1 << 20, device=dev, dtype=torch.int32)
    rc = lib.dace_b200_comm_allreduce(comm, 8, 2, big.data_ptr(), big.numel(), s)
    assert rc == 801, "bitwise or beyond the peer cap: cudaErrorNotSupported, got %d" % rc
l = torch.full((10,), 1 if rank == 0 else 0, device=dev, dtype=torch.int32)
assert bool((allreduce(l.clone(), 7) == 1).all()) and bool((allreduce(l.clone(), 5) == (1 if world == 1 else 0)).all())
if not isolate:
    torch.cuda.synchronize(); dist.barrier()
    lib.dace_b200_comm_destroy(comm)
    dist.barrier(); dist.destroy_process_group()
    print("rank", rank, "ok")
    sys.exit(0)
# (b) the DaCe-built program: communicator from the environment (RANK/WORLD_SIZE + rendezvous file)
built = os.path.join(os.environ["REPO_ROOT"], "tests", "dace_e2e", "_built")
rng = np.random.default_rng(7)
R, M = 4000 * world + 3, 300
full = rng.standard_normal((R, M)).astype(np.float32)
keys = rng.integers(0, M, R).astype(np.int32)
lo, hi = (R * rank) // world, (R * (rank + 1)) // world
a, k = np.ascontiguousarray(full[lo:hi]), np.ascontiguousarray(keys[lo:hi])
cs, tot, rs = np.zeros(M, np.float32), np.zeros(1, np.float32), np.zeros(hi - lo, np.float32)
bins = np.full(M, 7, np.int32)
p = ref_loader.RefProgram("b200_sharded", built)
for rep in range(3):
    p(A=a, keys=k, cs=cs, tot=tot, rs=rs, bins=bins, N=hi - lo, M=M)
    scale = np.abs(full.astype(np.float64)).sum(0)
    assert np.all(np.abs(cs - full.astype(np.float64).sum(0)) <= 1e-5 * scale), "sharded column sum"
    assert tot[0] == max(full.max(), np.float32(0.25)), "sharded max"
    assert np.allclose(rs, a.astype(np.float64).sum(1), rtol=1e-5, atol=1e-4), "row sums stay local"
    assert np.array_equal(bins, 7 + (rep + 1) * np.bincount(keys, minlength=M)), "sharded histogram"
neg = np.ascontiguousarray(-np.abs(a) - 1)
p(A=neg, keys=k, cs=cs, tot=tot, rs=rs, bins=bins, N=hi - lo, M=M)
assert tot[0] == np.float32(0.25), "non-neutral identity folded exactly once, got %r" % tot[0]
dflt = ctypes.c_void_p()
_lib.check(lib.dace_b200_comm_default(ctypes.byref(dflt)), "comm_default")
assert lib.dace_b200_comm_world(dflt) == world and lib.dace_b200_comm_rank(dflt) == rank
del p
torch.cuda.synchronize(); dist.barrier()
lib.dace_b200_comm_destroy(dflt); lib.dace_b200_comm_destroy(comm)
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
'''


@pytest.mark.parametrize('isolate,no_peer', [('0', '0'), ('1', '0'), ('1', '1')])
def test_native_communicator_and_sharded_dace_program(tmp_path, isolate, no_peer):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs >= 2 GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(os.path.join(root, 'tests', 'dace_e2e', '_built', 'libb200_sharded.so')):
        pytest.skip('tests/dace_e2e/_built not present')
    n = min(n, 4)
    script = tmp_path / 'comm_worker.py'
    script.write_text(COMM_WORKER)
    env = dict(os.environ, REPO_ROOT=root, DACE_B200_NO_PEER=no_peer, DACE_B200_DEBUG='1', ISOLATE=isolate,
               DACE_B200_RENDEZVOUS=str(tmp_path / 'nccl.id'))
    r = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(n),
                        '--master-addr', '127.0.0.1', '--master-port', '29578', str(script)],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count('ok') == n


# ---- the same DaCe-built program under dace_b200.launch: no torch in the worker, NCCL found by dlopen ----

LAUNCH_WORKER = r'''
import ctypes, os, sys, numpy as np
sys.path.insert(0, os.environ["REPO_ROOT"])
from dace_b200 import _lib
from oracle import ref_loader
assert "torch" not in sys.modules
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
assert os.environ["CUDA_VISIBLE_DEVICES"] == str(rank)
built = os.path.join(os.environ["REPO_ROOT"], "tests", "dace_e2e", "_built")
rng = np.random.default_rng(9)
R, M = 3000 * world + 1, 256
full = rng.standard_normal((R, M)).astype(np.float32)
keys = rng.integers(0, M, R).astype(np.int32)
lo, hi = (R * rank) // world, (R * (rank + 1)) // world
a, k = np.ascontiguousarray(full[lo:hi]), np.ascontiguousarray(keys[lo:hi])
cs, tot, rs = np.zeros(M, np.float32), np.zeros(1, np.float32), np.zeros(hi - lo, np.float32)
bins = np.zeros(M, np.int32)
p = ref_loader.RefProgram("b200_sharded", built)
p(A=a, keys=k, cs=cs, tot=tot, rs=rs, bins=bins, N=hi - lo, M=M)
scale = np.abs(full.astype(np.float64)).sum(0)
assert np.all(np.abs(cs - full.astype(np.float64).sum(0)) <= 1e-5 * scale)
assert tot[0] == max(full.max(), np.float32(0.25))
assert np.array_equal(bins, np.bincount(keys, minlength=M))
assert np.allclose(rs, a.astype(np.float64).sum(1), rtol=1e-5, atol=1e-4)
sq = np.zeros(M, np.float32)
pf = ref_loader.RefProgram("b200_sharded_fused", built)       # shard -> fuse: fused kernel for the partial, then the combine
pf(A=a, sq=sq, N=hi - lo, M=M)
ref_sq = (full.astype(np.float64) ** 2).sum(0)
assert np.all(np.abs(sq - ref_sq) <= 1e-5 * ref_sq), "sharded + fused column sum of squares"
del pf
lib = _lib.load()
c = ctypes.c_void_p()
_lib.check(lib.dace_b200_comm_default(ctypes.byref(c)), "comm_default")
assert lib.dace_b200_comm_world(c) == world and lib.dace_b200_comm_rank(c) == rank
sys.stdout.write("rank %d ok peer=%d\n" % (rank, lib.dace_b200_comm_uses_peer_memory(c)))   # one write: ranks share the pipe
sys.stdout.flush()
del p
lib.dace_b200_comm_destroy(c)
'''


def test_launcher_runs_the_sharded_dace_program(tmp_path):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs >= 2 GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.exists(os.path.join(root, 'tests', 'dace_e2e', '_built', 'libb200_sharded.so')):
        pytest.skip('tests/dace_e2e/_built not present')
    n = min(n, 4)
    script = tmp_path / 'launch_worker.py'
    script.write_text(LAUNCH_WORKER)
    env = dict(os.environ, REPO_ROOT=root, DACE_B200_DEBUG='1')
    env.pop('CUDA_VISIBLE_DEVICES', None)
    r = subprocess.run([sys.executable, '-m', 'dace_b200.launch', '--nproc', str(n), '--timeout', '300', str(script)],
                       env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count('ok peer=') == n, r.stdout
