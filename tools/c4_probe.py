"""Ad-hoc probe: C4-shaped histogram under different key distributions, partitioned vs atomic path."""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dace_b200 as db
dev = torch.device('cuda:0')
nb, n = 1 << 20, 1 << 28
g = torch.Generator(device=dev); g.manual_seed(0)
def zipf():
    r = np.random.default_rng(0)
    k = r.zipf(1.1, n // 16).astype(np.int64) % nb
    return torch.from_numpy(np.tile(k.astype(np.int32), 16)).to(dev)
dists = {
    'uniform': lambda: torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g),
    'zipf1.1': zipf,
    'all_equal': lambda: torch.full((n,), 12345, device=dev, dtype=torch.int32),
    'sorted': lambda: torch.sort(torch.randint(0, nb, (n,), device=dev, dtype=torch.int32, generator=g)).values,
    'one_bucket': lambda: torch.randint(0, 32768, (n,), device=dev, dtype=torch.int32, generator=g),
}
for name, mk in dists.items():
    keys = mk()
    ref = torch.bincount(keys.long(), minlength=nb).to(torch.int32)
    for nopart in ('0', '1'):
        if nopart == '1': os.environ['DACE_B200_NO_PARTITION'] = '1'
        else: os.environ.pop('DACE_B200_NO_PARTITION', None)
        bins = torch.zeros(nb, device=dev, dtype=torch.int32)
        db.wcr_scatter('lambda a, b: a + b', bins, keys, None)
        torch.cuda.synchronize()
        ok = torch.equal(bins, ref)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): db.wcr_scatter('lambda a, b: a + b', bins, keys, None)
        e1.record(); torch.cuda.synchronize()
        print('%-10s %s  %s  %.3f ms' % (name, 'atomic     ' if nopart == '1' else 'partitioned', 'ok ' if ok else 'BAD', e0.elapsed_time(e1) / 5), flush=True)
    del keys, ref
