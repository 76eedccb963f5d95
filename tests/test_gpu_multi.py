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
    assert r.stdout.count('ok') == n
