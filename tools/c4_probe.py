"""Ad-hoc probe: C4-shaped histogram under different key distributions - the default path (device-side
selection between the radix partition, the warp-merged global atomics and the hot-key cache) against
the global-atomic kernel alone (a workspace too small for the partition forces it)."""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
from dace_b200.api import Workspace
dev = torch.device('cuda:0')
nb, n = 1 << 20, 1 << 28
g = torch.Generator(device=dev); g.manual_seed(0)
def zipf(wrap=True):
    r = np.random.default_rng(0)
    z = r.zipf(1.1, n // 16)
    k = (z % nb) if wrap else np.minimum(z - 1, nb - 1)
    return torch.from_numpy(np.tile(k.astype(np.int32), 16)).to(dev)
def runs(length):
    base = torch.randint(0, nb, (n // length, ), device=dev, dtype=torch.int32, generator=g)
    return base.repeat_interleave(length)
dists = {
    'uniform': lambda: torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g),
    'zipf1.1_wrapped': lambda: zipf(True),
    'zipf1.1_clipped': lambda: zipf(False),
    'all_equal': lambda: torch.full((n,), 12345, device=dev, dtype=torch.int32),
    'sorted': lambda: torch.sort(torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g)).values,
    'one_bucket': lambda: torch.randint(0, 32768, (n,), device=dev, dtype=torch.int32, generator=g),
    'runs_of_64': lambda: runs(64),
    'half_hot': lambda: torch.where(torch.rand(n, device=dev, generator=g) < 0.5, torch.full((n,), 777, device=dev, dtype=torch.int32),
                                    torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g)),
}
tiny = Workspace(dev, nbytes=256, zero_init=False)
only = sys.argv[1:]
for name, mk in dists.items():
    if only and name not in only:
        continue
    keys = mk()
    ref = torch.bincount(keys, minlength=nb).to(torch.int32)
    for label, ws in (('selected', None), ('atomic  ', tiny)):
        bins = torch.zeros(nb, device=dev, dtype=torch.int32)
        db.wcr_scatter('lambda a, b: a + b', bins, keys, None, workspace=ws)
        torch.cuda.synchronize()
        ok = torch.equal(bins, ref)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): db.wcr_scatter('lambda a, b: a + b', bins, keys, None, workspace=ws)
        e1.record(); torch.cuda.synchronize()
        print('%-16s %s  %s  %.3f ms' % (name, label, 'ok ' if ok else 'BAD', e0.elapsed_time(e1) / 5), flush=True)
    del keys, ref
