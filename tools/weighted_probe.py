"""Ad-hoc probe: weighted histogram (per-element float values) of 2^28 keys into 2^20 bins."""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
from dace_b200.api import Workspace
dev = torch.device('cuda:0')
nb, n = 1 << 20, 1 << 28
g = torch.Generator(device=dev); g.manual_seed(0)
w = torch.rand(n, device=dev, generator=g)
r = np.random.default_rng(0)
z = torch.from_numpy(np.tile((r.zipf(1.1, n // 16) % nb).astype(np.int32), 16)).to(dev)
u = torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g)
# DACE_B200_NO_PARTITION=1 (read when the library is loaded) turns the device-side selection off: run this probe
# once with and once without it to compare the hot-key cache with plain global atomics.
label = 'global atomics only' if os.environ.get('DACE_B200_NO_PARTITION') else 'device-side selection'
wi = torch.randint(0, 5, (n,), device=dev, dtype=torch.int32, generator=g)
for name, keys in (('uniform', u), ('zipf1.1', z)):
    for wname, wt in (('f32', w), ('i32', wi)):
        bins = torch.zeros(nb, device=dev, dtype=wt.dtype)
        db.wcr_scatter('lambda a, b: a + b', bins, keys, wt)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): db.wcr_scatter('lambda a, b: a + b', bins, keys, wt)
        e1.record(); torch.cuda.synchronize()
        print('weighted %s %-8s %-22s %.3f ms' % (wname, name, label, e0.elapsed_time(e1) / 3), flush=True)
